// Stand-in for <boost/math/constants/constants.hpp> (Boost is absent).  TEST
// INFRASTRUCTURE ONLY.  The decimal expansions below are the published values;
// each literal rounds to the nearest double exactly as Boost's own tables do.
#pragma once
#include <boost/assert.hpp>
namespace boost { namespace math {
namespace constants {
template <class T> constexpr T pi() { return static_cast<T>(3.141592653589793238462643383279502884e+00L); }
template <class T> constexpr T two_pi() { return static_cast<T>(6.283185307179586476925286766559005768e+00L); }
template <class T> constexpr T half_pi() { return static_cast<T>(1.570796326794896619231321691639751442e+00L); }
template <class T> constexpr T third() { return static_cast<T>(3.333333333333333333333333333333333333e-01L); }
template <class T> constexpr T root_three() { return static_cast<T>(1.732050807568877293527446341505872367e+00L); }
}
namespace double_constants {
constexpr double pi = 3.141592653589793238462643383279502884e+00;
constexpr double two_pi = 6.283185307179586476925286766559005768e+00;
constexpr double half_pi = 1.570796326794896619231321691639751442e+00;
constexpr double third = 3.333333333333333333333333333333333333e-01;
constexpr double root_three = 1.732050807568877293527446341505872367e+00;
}
}}
