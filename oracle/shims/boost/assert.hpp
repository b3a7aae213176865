// Stand-in for <boost/assert.hpp> (Boost is absent).  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <cassert>
#define BOOST_ASSERT(x) assert(x)
#define BOOST_ASSERT_MSG(x, m) assert((x) && (m))
