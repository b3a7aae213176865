// dmath.cuh -- FP64 transcendentals used by the exact path (SO3 arc length and
// slerp, kinematic-chain forward kinematics).
//
// The reference calls glibc's libm (acos/sin/cos; < 1 ulp, usually but not
// always correctly rounded).  Round 1 uses CUDA's own double-precision
// routines (documented max error: sin/cos 2 ulp over the full range, acos
// 1 ulp), which agree with glibc in all but the last bit of a minority of
// arguments.  A last-bit difference changes an observable result only when a
// state sits within ~1e-16 of an obstacle boundary, or distance /
// longestValidSegment within ~1e-16 of an integer;
// RealVector, SO2, SE2 and the hypercube / circle checkers use none of these
// and are bit-exact by construction.  tests/ measure the booleans on seeded
// batches; DESIGN.md records the status.
#pragma once

namespace omplcu
{
    __device__ __forceinline__ double omplcu_sin(double x)
    {
        return sin(x);
    }
    __device__ __forceinline__ double omplcu_cos(double x)
    {
        return cos(x);
    }
    __device__ __forceinline__ double omplcu_acos(double x)
    {
        return acos(x);
    }
}
