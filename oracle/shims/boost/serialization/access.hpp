// Stand-in for <boost/serialization/access.hpp> (test infrastructure, see oracle/Makefile): the
// reference's base/PlannerData.h only befriends this class.
#pragma once
namespace boost
{
    namespace serialization
    {
        class access;
    }
}
