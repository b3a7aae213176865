// common.cuh -- shared host/device definitions for libomplcu.
//
// Exact FP64 state-space math.  The reference x86-64 build has no FMA
// contraction (CMakeModules/CompilerSettings.cmake:5-9), so every multiply and
// add rounds on its own; the device code spells each operation with the
// round-to-nearest intrinsics (__dmul_rn/__dadd_rn/...), which nvcc never
// fuses, and keeps the reference's evaluation order.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdexcept>
#include <string>

#include "../../include/omplcu.h"

namespace omplcu
{
    struct Error : std::runtime_error
    {
        int code;
        Error(int c, const std::string &m) : std::runtime_error(m), code(c)
        {
        }
    };

#define CUDA_CHECK(expr)                                                                                     \
    do                                                                                                       \
    {                                                                                                        \
        cudaError_t e__ = (expr);                                                                            \
        if (e__ != cudaSuccess)                                                                              \
            throw ::omplcu::Error(OMPLCU_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));     \
    } while (0)

#define REQUIRE(cond, msg)                                                                                   \
    do                                                                                                       \
    {                                                                                                        \
        if (!(cond))                                                                                         \
            throw ::omplcu::Error(OMPLCU_ERR_INVALID, msg);                                                  \
    } while (0)

    // Device-side copy of the space description (passed to kernels by value).
    struct SpaceDesc
    {
        int ncomp;
        int compound;
        int stride;
        int _pad;
        omplcu_component comp[OMPLCU_MAX_COMPONENTS];
    };

    // largest state row kept in registers / local memory by the generic kernels
    constexpr int MAX_STRIDE = 128;

    struct DeviceGuard
    {
        int prev;
        explicit DeviceGuard(int dev)
        {
            cudaGetDevice(&prev);
            if (prev != dev)
                cudaSetDevice(dev);
        }
        ~DeviceGuard()
        {
            int cur;
            cudaGetDevice(&cur);
            if (cur != prev)
                cudaSetDevice(prev);
        }
    };

    // Simple owning device buffer that only grows.
    template <typename T>
    struct DevBuf
    {
        T *p = nullptr;
        size_t cap = 0;
        void reserve(size_t n, bool keep = false, cudaStream_t s = 0)
        {
            if (n <= cap)
                return;
            size_t ncap = cap ? cap : 1;
            while (ncap < n)
                ncap *= 2;
            T *np = nullptr;
            CUDA_CHECK(cudaMalloc(&np, ncap * sizeof(T)));
            if (keep && p && cap)
            {
                CUDA_CHECK(cudaMemcpyAsync(np, p, cap * sizeof(T), cudaMemcpyDeviceToDevice, s));
                CUDA_CHECK(cudaStreamSynchronize(s));
            }
            if (p)
                cudaFree(p);
            p = np;
            cap = ncap;
        }
        void release()
        {
            if (p)
                cudaFree(p);
            p = nullptr;
            cap = 0;
        }
        ~DevBuf()
        {
            release();
        }
        DevBuf() = default;
        DevBuf(const DevBuf &) = delete;
        DevBuf &operator=(const DevBuf &) = delete;
        DevBuf(DevBuf &&o) noexcept : p(o.p), cap(o.cap)
        {
            o.p = nullptr;
            o.cap = 0;
        }
    };

    // Page-locked host staging that only grows (small host-pointer calls: one asynchronous copy each way
    // instead of the driver's synchronous pageable path).
    struct PinBuf
    {
        unsigned char *p = nullptr;
        size_t cap = 0;
        void reserve(size_t n)
        {
            if (n <= cap)
                return;
            size_t ncap = cap ? cap : 4096;
            while (ncap < n)
                ncap *= 2;
            if (p)
                cudaFreeHost(p);
            p = nullptr;
            cap = 0;
            CUDA_CHECK(cudaMallocHost(reinterpret_cast<void **>(&p), ncap));
            cap = ncap;
        }
        ~PinBuf()
        {
            if (p)
                cudaFreeHost(p);
        }
        PinBuf() = default;
        PinBuf(const PinBuf &) = delete;
        PinBuf &operator=(const PinBuf &) = delete;
    };

    // ------------------------------------------------------------------
    // exact FP64 device math
    // ------------------------------------------------------------------
#ifdef __CUDACC__
    __device__ __forceinline__ double dmul(double a, double b)
    {
        return __dmul_rn(a, b);
    }
    __device__ __forceinline__ double dadd(double a, double b)
    {
        return __dadd_rn(a, b);
    }
    __device__ __forceinline__ double dsub(double a, double b)
    {
        return __dsub_rn(a, b);
    }
    __device__ __forceinline__ double ddiv(double a, double b)
    {
        return __ddiv_rn(a, b);
    }
    __device__ __forceinline__ double dsqrt(double a)
    {
        return __dsqrt_rn(a);
    }

    // boost::math::constants::pi<double>() (nearest double)
    __device__ constexpr double kPi = 3.14159265358979323846;
    __device__ constexpr double kTwoPi = 2.0 * 3.14159265358979323846;  // exact doubling
    __device__ constexpr double kDblEps = 2.220446049250313e-16;
    __device__ constexpr double kFltEps = 1.1920928955078125e-07;
    __device__ constexpr double kMaxQuatNormErr = 1e-9;

    // Transcendentals: see dmath.cuh (kept behind these names so the
    // implementation can be swapped without touching callers).
    __device__ __forceinline__ double omplcu_sin(double x);
    __device__ __forceinline__ double omplcu_cos(double x);
    __device__ __forceinline__ double omplcu_acos(double x);
}
#include "dmath.cuh"
namespace omplcu
{
    // RealVectorStateSpace::distance, RealVectorStateSpace.cpp:230-242
    template <typename PA, typename PB>
    __device__ __forceinline__ double rv_distance(PA a, PB b, int d)
    {
        double dist = 0.0;
        for (int i = 0; i < d; ++i)
        {
            double diff = dsub(a[i], b[i]);
            dist = dadd(dist, dmul(diff, diff));
        }
        return dsqrt(dist);
    }

    // SO2StateSpace::distance, SO2StateSpace.cpp:116-128
    __device__ __forceinline__ double so2_distance(double a, double b)
    {
        double d = fabs(dsub(a, b));
        return (d > kPi) ? dsub(kTwoPi, d) : d;
    }

    // SO2StateSpace::interpolate, SO2StateSpace.cpp:136-155 (also the per-joint
    // rule of KinematicChainSpace::interpolate, demos/KinematicChain.h:150-176)
    __device__ __forceinline__ double so2_interpolate(double from, double to, double t)
    {
        double diff = dsub(to, from);
        if (fabs(diff) <= kPi)
            return dadd(from, dmul(diff, t));
        if (diff > 0.0)
            diff = dsub(kTwoPi, diff);
        else
            diff = dsub(-kTwoPi, diff);
        double v = dsub(from, dmul(diff, t));
        if (v > kPi)
            v = dsub(v, kTwoPi);
        else if (v < -kPi)
            v = dadd(v, kTwoPi);
        return v;
    }

    // raw quaternion dot product in the reference's order (SO3StateSpace.cpp:258)
    template <typename PA, typename PB>
    __device__ __forceinline__ double quat_dot(PA a, PB b)
    {
        return dadd(dadd(dadd(dmul(a[0], b[0]), dmul(a[1], b[1])), dmul(a[2], b[2])), dmul(a[3], b[3]));
    }

    // arcLength, SO3StateSpace.cpp:254-262
    template <typename PA, typename PB>
    __device__ __forceinline__ double so3_arc_length(PA a, PB b)
    {
        double dq = fabs(quat_dot(a, b));
        if (dq > 1.0 - kMaxQuatNormErr)
            return 0.0;
        return omplcu_acos(dq);
    }

    // SO3StateSpace::interpolate, SO3StateSpace.cpp:289-318
    template <typename PA, typename PB, typename PO>
    __device__ __forceinline__ void so3_interpolate(PA from, PB to, double t, PO out)
    {
        double theta = so3_arc_length(from, to);
        if (theta > kDblEps)
        {
            double d = ddiv(1.0, omplcu_sin(theta));
            double s0 = omplcu_sin(dmul(dsub(1.0, t), theta));
            double s1 = omplcu_sin(dmul(t, theta));
            double dq = quat_dot(from, to);
            if (dq < 0)
                s1 = -s1;
            for (int i = 0; i < 4; ++i)
                out[i] = dmul(dadd(dmul(from[i], s0), dmul(to[i], s1)), d);
        }
        else
        {
            for (int i = 0; i < 4; ++i)
                out[i] = from[i];
        }
    }

    // SO3StateSpace::interpolate split in two: the part that depends only on the end
    // points (theta, 1/sin(theta), sign of the dot product) and the part that depends
    // on t.  Same operations in the same order as so3_interpolate, so the states are
    // bit-identical; the motion kernel computes the first part once per segment.
    struct So3Seg
    {
        double theta, d;
        bool neg;
    };
    template <typename PA, typename PB>
    __device__ __forceinline__ So3Seg so3_prepare(PA from, PB to)
    {
        So3Seg g;
        g.theta = so3_arc_length(from, to);
        g.d = 0.0;
        g.neg = false;
        if (g.theta > kDblEps)
        {
            g.d = ddiv(1.0, omplcu_sin(g.theta));
            g.neg = quat_dot(from, to) < 0;
        }
        return g;
    }
    template <typename PA, typename PB, typename PO>
    __device__ __forceinline__ void so3_interpolate_prepared(const So3Seg &g, PA from, PB to, double t, PO out)
    {
        if (g.theta > kDblEps)
        {
            double s0 = omplcu_sin(dmul(dsub(1.0, t), g.theta));
            double s1 = omplcu_sin(dmul(t, g.theta));
            if (g.neg)
                s1 = -s1;
            for (int i = 0; i < 4; ++i)
                out[i] = dmul(dadd(dmul(from[i], s0), dmul(to[i], s1)), g.d);
        }
        else
        {
            for (int i = 0; i < 4; ++i)
                out[i] = from[i];
        }
    }

    // KinematicChainSpace::distance, demos/KinematicChain.h:105-121
    template <typename PA, typename PB>
    __device__ __forceinline__ double chain_distance(PA a, PB b, int d, double linkLength)
    {
        double theta1 = 0., theta2 = 0., dx = 0., dy = 0., dist = 0.;
        for (int i = 0; i < d; ++i)
        {
            theta1 = dadd(theta1, a[i]);
            theta2 = dadd(theta2, b[i]);
            dx = dadd(dx, dsub(omplcu_cos(theta1), omplcu_cos(theta2)));
            dy = dadd(dy, dsub(omplcu_sin(theta1), omplcu_sin(theta2)));
            dist = dadd(dist, dsqrt(dadd(dmul(dx, dx), dmul(dy, dy))));
        }
        return dmul(dist, linkLength);
    }

    template <typename PA, typename PB>
    __device__ __forceinline__ double comp_distance(const omplcu_component &c, PA a, PB b)
    {
        switch (c.type)
        {
            case OMPLCU_RV:
                return rv_distance(a, b, c.dim);
            case OMPLCU_SO2:
                return so2_distance(a[0], b[0]);
            case OMPLCU_SO3:
                return so3_arc_length(a, b);
            default:
                return chain_distance(a, b, c.dim, c.param);
        }
    }

    // CompoundStateSpace::distance, StateSpace.cpp:1068-1076
    template <typename PA, typename PB>
    __device__ __forceinline__ double space_distance(const SpaceDesc &sp, PA a, PB b)
    {
        if (!sp.compound)
            return comp_distance(sp.comp[0], a, b);
        double dist = 0.0;
        int off = 0;
        for (int i = 0; i < sp.ncomp; ++i)
        {
            dist = dadd(dist, dmul(sp.comp[i].weight, comp_distance(sp.comp[i], a + off, b + off)));
            off += sp.comp[i].dim;
        }
        return dist;
    }

    // CompoundStateSpace::interpolate, StateSpace.cpp:1109-1116
    template <typename PA, typename PB, typename PO>
    __device__ __forceinline__ void space_interpolate(const SpaceDesc &sp, PA from, PB to, double t, PO out)
    {
        int off = 0;
        for (int c = 0; c < sp.ncomp; ++c)
        {
            const omplcu_component &cc = sp.comp[c];
            if (cc.type == OMPLCU_RV)
            {
                // RealVectorStateSpace::interpolate, RealVectorStateSpace.cpp:257-265
                for (int i = 0; i < cc.dim; ++i)
                    out[off + i] = dadd(from[off + i], dmul(dsub(to[off + i], from[off + i]), t));
            }
            else if (cc.type == OMPLCU_SO2)
                out[off] = so2_interpolate(from[off], to[off], t);
            else if (cc.type == OMPLCU_SO3)
                so3_interpolate(from + off, to + off, t, out + off);
            else
                for (int i = 0; i < cc.dim; ++i)
                    out[off + i] = so2_interpolate(from[off + i], to[off + i], t);
            off += cc.dim;
        }
    }

    // The same with the SO(3) components' per-segment constants supplied by the caller
    // (prep[k] belongs to the k-th SO(3) component, k < MAX_SO3_PREP; further ones are
    // computed in place).
    constexpr int MAX_SO3_PREP = 2;
    template <typename PA, typename PB>
    __device__ __forceinline__ void space_prepare(const SpaceDesc &sp, PA from, PB to, So3Seg (&prep)[MAX_SO3_PREP])
    {
        int off = 0, k = 0;
        for (int c = 0; c < sp.ncomp; ++c)
        {
            if (sp.comp[c].type == OMPLCU_SO3 && k < MAX_SO3_PREP)
                prep[k++] = so3_prepare(from + off, to + off);
            off += sp.comp[c].dim;
        }
    }
    template <typename PA, typename PB, typename PO>
    __device__ __forceinline__ void space_interpolate_prepared(const SpaceDesc &sp, const So3Seg (&prep)[MAX_SO3_PREP],
                                                               PA from, PB to, double t, PO out)
    {
        int off = 0, k = 0;
        for (int c = 0; c < sp.ncomp; ++c)
        {
            const omplcu_component &cc = sp.comp[c];
            if (cc.type == OMPLCU_RV)
            {
                for (int i = 0; i < cc.dim; ++i)
                    out[off + i] = dadd(from[off + i], dmul(dsub(to[off + i], from[off + i]), t));
            }
            else if (cc.type == OMPLCU_SO2)
                out[off] = so2_interpolate(from[off], to[off], t);
            else if (cc.type == OMPLCU_SO3)
            {
                if (k < MAX_SO3_PREP)
                    so3_interpolate_prepared(prep[k++], from + off, to + off, t, out + off);
                else
                    so3_interpolate(from + off, to + off, t, out + off);
            }
            else
                for (int i = 0; i < cc.dim; ++i)
                    out[off + i] = so2_interpolate(from[off + i], to[off + i], t);
            off += cc.dim;
        }
    }

    // StateSpace::validSegmentCount, StateSpace.cpp:851-854, and the compound
    // max rule, StateSpace.cpp:1085-1097.  (unsigned int)ceil(x) for x >= 0.
    template <typename PA, typename PB>
    __device__ __forceinline__ unsigned space_segment_count(const SpaceDesc &sp, PA a, PB b)
    {
        unsigned sc = 0;
        int off = 0;
        for (int i = 0; i < sp.ncomp; ++i)
        {
            double x = ceil(ddiv(comp_distance(sp.comp[i], a + off, b + off), sp.comp[i].lvs));
            unsigned sci = sp.comp[i].factor * (unsigned)x;
            if (!sp.compound)
                return sci;
            if (sci > sc)
                sc = sci;
            off += sp.comp[i].dim;
        }
        return sc;
    }
#endif  // __CUDACC__
}  // namespace omplcu
