#pragma once
#define BOOST_CONCEPT_ASSERT(x) static_assert(true, "")
namespace boost { template <class A, class B> struct Convertible {}; }
