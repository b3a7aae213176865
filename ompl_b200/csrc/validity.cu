// validity.cu -- batched state validity and motion validity.
//
// Replaces, for batches:
//   StateValidityChecker::isValid            base/StateValidityChecker.h:111
//   DiscreteMotionValidator::checkMotion     base/src/DiscreteMotionValidator.cpp:48-145
// with the demo checkers as device code:
//   hypercube   demos/HypercubeBenchmark.cpp:57-72
//   chain       demos/KinematicChain.h:200-276
//   circles     tests/resources/circles2D.h:139-149
//   se3         not in the reference (demos/RigidBodyPlanning.cpp:48-62 is always
//               true); specified in DESIGN.md, oracle in oracle/ompl_oracle.c.
//
// A motion (s1, s2) is valid iff s2 and every interpolated state j/nd,
// j = 1..nd-1, nd = validSegmentCount(s1, s2), is valid.  The reference visits
// them in bisection order and stops at the first failure
// (DiscreteMotionValidator.cpp:115-134); the checkers are pure, so the result
// does not depend on the order.  Here a GROUP of G lanes (G = 4..32, chosen from
// the mean state count per segment) owns one segment at a time: the lanes take
// the states j = 1, 2, ... in increasing order, G per pass, then s2; one warp
// ballot per pass finds the first failing state, which ends the segment (early
// exit) and is exactly what the lastValid overload reports
// (DiscreteMotionValidator.cpp:60-72).  Without lastValid any failing state
// decides, and the lanes visit the states spread out over the segment instead
// (see `spread` in motion_group_kernel).  Groups pull segments from a global
// counter, so long and short segments balance across the persistent grid.
//
// Interpolation and the deciding predicate are FP64 with the reference's
// operation order (common.cuh).  The SE3 checker puts an FP32 bounding-sphere
// filter in front of the FP64 predicate (se3_valid_fast); the filter is
// conservative by a proven margin, so the booleans are those of the exact test.
#include "validity.cuh"

#include <algorithm>
#include <cstdlib>
#include <limits>
#include <vector>

#include <cub/cub.cuh>

namespace omplcu
{
    namespace
    {
        constexpr unsigned CODE_VALID = 0x7FFFFFFFu;
        constexpr unsigned CODE_S2 = 0x7FFFFFFEu;

        // ---- per-state predicates ----
        template <typename PS>
        __device__ __forceinline__ bool hypercube_valid(const CheckerDesc &c, PS s)
        {
            bool found = false;
            const double hi = dsub(1., c.edge_width);
            for (int i = c.dim - 1; i >= 0; i--)
                if (!found)
                {
                    if (s[i] > c.edge_width)
                        found = true;
                }
                else if (s[i] < hi)
                    return false;
            return true;
        }

        template <typename PS>
        __device__ __forceinline__ bool circles_valid(const CheckerDesc &c, PS s)
        {
            const double x = s[0], y = s[1];
            for (int i = 0; i < c.ncirc; ++i)
            {
                const double dx = dsub(c.circ[3 * i + 0], x);
                const double dy = dsub(c.circ[3 * i + 1], y);
                const double r = c.circ[3 * i + 2];
                if (dadd(dmul(dx, dx), dmul(dy, dy)) < dmul(r, r))
                    return false;
            }
            return true;
        }

        // obstacle arrays may live in shared memory (obs != nullptr) or global
        template <typename PS>
        __device__ __forceinline__ bool se3_valid(const CheckerDesc &c, const double *robot, const double *sph,
                                                  const double *box, PS st)
        {
            const double px = st[0], py = st[1], pz = st[2];
            const double qx = st[3], qy = st[4], qz = st[5], qw = st[6];
            for (int i = 0; i < c.nrobot; ++i)
            {
                const double ox = robot[4 * i + 0], oy = robot[4 * i + 1], oz = robot[4 * i + 2];
                const double rr = robot[4 * i + 3];
                const double tx = dmul(2.0, dsub(dmul(qy, oz), dmul(qz, oy)));
                const double ty = dmul(2.0, dsub(dmul(qz, ox), dmul(qx, oz)));
                const double tz = dmul(2.0, dsub(dmul(qx, oy), dmul(qy, ox)));
                const double cx = dadd(px, dadd(dadd(ox, dmul(qw, tx)), dsub(dmul(qy, tz), dmul(qz, ty))));
                const double cy = dadd(py, dadd(dadd(oy, dmul(qw, ty)), dsub(dmul(qz, tx), dmul(qx, tz))));
                const double cz = dadd(pz, dadd(dadd(oz, dmul(qw, tz)), dsub(dmul(qx, ty), dmul(qy, tx))));
                for (int j = 0; j < c.nsph; ++j)
                {
                    const double dx = dsub(sph[4 * j + 0], cx);
                    const double dy = dsub(sph[4 * j + 1], cy);
                    const double dz = dsub(sph[4 * j + 2], cz);
                    const double R = dadd(sph[4 * j + 3], rr);
                    if (dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz)) < dmul(R, R))
                        return false;
                }
                const double rr2 = dmul(rr, rr);
                for (int j = 0; j < c.nbox; ++j)
                {
                    double ex = dsub(fabs(dsub(cx, box[6 * j + 0])), box[6 * j + 3]);
                    double ey = dsub(fabs(dsub(cy, box[6 * j + 1])), box[6 * j + 4]);
                    double ez = dsub(fabs(dsub(cz, box[6 * j + 2])), box[6 * j + 5]);
                    if (ex < 0.0)
                        ex = 0.0;
                    if (ey < 0.0)
                        ey = 0.0;
                    if (ez < 0.0)
                        ez = 0.0;
                    if (dadd(dadd(dmul(ex, ex), dmul(ey, ey)), dmul(ez, ez)) < rr2)
                        return false;
                }
            }
            return true;
        }

        // KinematicChainValidityChecker::intersectionTest, demos/KinematicChain.h:251-276
        __device__ __forceinline__ bool seg_intersect(double ax0, double ay0, double ax1, double ay1, double bx0,
                                                      double by0, double bx1, double by1)
        {
            const double s10_x = dsub(ax1, ax0);
            const double s10_y = dsub(ay1, ay0);
            const double s32_x = dsub(bx1, bx0);
            const double s32_y = dsub(by1, by0);
            const double denom = dsub(dmul(s10_x, s32_y), dmul(s32_x, s10_y));
            if (fabs(denom) < kDblEps)
                return false;
            const bool denomPositive = denom > 0;
            const double s02_x = dsub(ax0, bx0);
            const double s02_y = dsub(ay0, by0);
            const double s_numer = dsub(dmul(s10_x, s02_y), dmul(s10_y, s02_x));
            if ((s_numer < kFltEps) == denomPositive)
                return false;
            const double t_numer = dsub(dmul(s32_x, s02_y), dmul(s32_y, s02_x));
            if ((t_numer < kFltEps) == denomPositive)
                return false;
            if (((dsub(s_numer, denom) > -kFltEps) == denomPositive) ||
                ((dsub(t_numer, denom) > kFltEps) == denomPositive))
                return false;
            return true;
        }

        // KinematicChainValidityChecker::isValidImpl, demos/KinematicChain.h:209-249.
        // Joint positions are kept as a polyline: segment i = (xs[i],ys[i])-(xs[i+1],ys[i+1]).
        template <typename PS>
        __device__ bool chain_valid(const CheckerDesc &c, PS s)
        {
            const int n = c.nlinks;
            double xs[MAX_CHAIN_LINKS + 2], ys[MAX_CHAIN_LINKS + 2];
            double theta = 0., x = 0., y = 0.;
            xs[0] = 0.;
            ys[0] = 0.;
            for (int i = 0; i < n; ++i)
            {
                theta = dadd(theta, s[i]);
                x = dadd(x, dmul(omplcu_cos(theta), c.link_length));
                y = dadd(y, dmul(omplcu_sin(theta), c.link_length));
                xs[i + 1] = x;
                ys[i + 1] = y;
            }
            xs[n + 1] = dadd(x, dmul(omplcu_cos(theta), 0.001));
            ys[n + 1] = dadd(y, dmul(omplcu_sin(theta), 0.001));
            for (int i = 0; i < n + 1; ++i)
                for (int j = i + 1; j < n + 1; ++j)
                    if (seg_intersect(xs[i], ys[i], xs[i + 1], ys[i + 1], xs[j], ys[j], xs[j + 1], ys[j + 1]))
                        return false;
            for (int i = 0; i < n + 1; ++i)
                for (int j = 0; j < c.nenv; ++j)
                    if (seg_intersect(xs[i], ys[i], xs[i + 1], ys[i + 1], c.env[4 * j + 0], c.env[4 * j + 1],
                                      c.env[4 * j + 2], c.env[4 * j + 3]))
                        return false;
            return true;
        }

        template <int KIND, typename PS>
        __device__ __forceinline__ bool state_valid(const CheckerDesc &c, const double *robot, const double *sph,
                                                    const double *box, PS s)
        {
            if (KIND == OMPLCU_CHK_HYPERCUBE)
                return hypercube_valid(c, s);
            if (KIND == OMPLCU_CHK_CHAIN)
                return chain_valid(c, s);
            if (KIND == OMPLCU_CHK_SE3)
                return se3_valid(c, robot, sph, box, s);
            if (KIND == OMPLCU_CHK_CIRCLES)
                return circles_valid(c, s);
            return true;  // AllValidStateValidityChecker, base/StateValidityChecker.h:165-182
        }

        // ---- SE3: FP32 bounding-sphere filter in front of the FP64 predicate ----
        //
        // Every obstacle j gets a bounding sphere (centre o_j, radius rho_j: the sphere
        // itself, or |half extents| for a box) and the robot one around the body origin p
        // (radius rho_rob = max_i(|offset_i| + r_i), inflated by 2^-10).  A robot sphere
        // can touch obstacle j only if |o_j - p| < rho_j + rho_rob =: sqrt(W_j), i.e. only if
        //     (|o_j|^2 - W_j) - 2 o_j.p  <  -|p|^2.
        // The left side is 3 FP32 FMAs per obstacle on the row (o_j, |o_j|^2 - W_j) kept in
        // shared memory; the right side is one value per state.  With A >= every |o_c|,
        // |p_c| and sqrt(W_j), the FP32 evaluation of both sides is off by less than
        // 64 u A^2 (u = 2^-24; 12 u A^2 from rounding o and p, 27 u A^2 + 3 u W from the three
        // FMAs, 3 u A^2 + u W from the stored norm, 18 u A^2 from the right side); the test
        // uses a margin of 2^-16 A^2, four times that, so an obstacle that fails the filter
        // provably cannot collide.  Survivors (about two per state in the benchmark
        // world) go through the exact FP64 expressions of se3_valid for that obstacle.
        // The body rotation maps offset o to (1-s) o + s R o with s = |q|^2, so
        // |c_i - p| <= (|1-s| + s)|o_i|; states with |s - 1| > 2e-4 skip the filter.
        // robot sphere i in the world frame (the expressions of se3_valid, same order)
        template <typename PS>
        __device__ __forceinline__ void se3_robot_centre(const CheckerDesc &c, int i, PS st, double &cx, double &cy,
                                                         double &cz, double &rr)
        {
            const double px = st[0], py = st[1], pz = st[2];
            const double qx = st[3], qy = st[4], qz = st[5], qw = st[6];
            const double ox = c.robot[4 * i + 0], oy = c.robot[4 * i + 1], oz = c.robot[4 * i + 2];
            rr = c.robot[4 * i + 3];
            const double tx = dmul(2.0, dsub(dmul(qy, oz), dmul(qz, oy)));
            const double ty = dmul(2.0, dsub(dmul(qz, ox), dmul(qx, oz)));
            const double tz = dmul(2.0, dsub(dmul(qx, oy), dmul(qy, ox)));
            cx = dadd(px, dadd(dadd(ox, dmul(qw, tx)), dsub(dmul(qy, tz), dmul(qz, ty))));
            cy = dadd(py, dadd(dadd(oy, dmul(qw, ty)), dsub(dmul(qz, tx), dmul(qx, tz))));
            cz = dadd(pz, dadd(dadd(oz, dmul(qw, tz)), dsub(dmul(qx, ty), dmul(qy, tx))));
        }

        // exact test of one robot sphere against obstacle j (spheres first, then boxes)
        __device__ __forceinline__ bool se3_obstacle_hits(const CheckerDesc &c, int j, double cx, double cy, double cz,
                                                          double rr)
        {
            if (j < c.nsph)
            {
                const double *sp = c.sph + 4 * j;
                const double dx = dsub(sp[0], cx);
                const double dy = dsub(sp[1], cy);
                const double dz = dsub(sp[2], cz);
                const double R = dadd(sp[3], rr);
                return dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz)) < dmul(R, R);
            }
            const double *bx = c.box + 6 * (j - c.nsph);
            double ex = dsub(fabs(dsub(cx, bx[0])), bx[3]);
            double ey = dsub(fabs(dsub(cy, bx[1])), bx[4]);
            double ez = dsub(fabs(dsub(cz, bx[2])), bx[5]);
            if (ex < 0.0)
                ex = 0.0;
            if (ey < 0.0)
                ey = 0.0;
            if (ez < 0.0)
                ez = 0.0;
            return dadd(dadd(dmul(ex, ex), dmul(ey, ey)), dmul(ez, ez)) < dmul(rr, rr);
        }

        constexpr int SURV_CAP = 12;  // survivors kept per lane before the exact pass

        // F: filter rows in shared memory; surv_base: a [SURV_CAP][blockDim.x] u16 array in shared
        // memory (one column per thread) followed by one float4 of scratch per thread.  Survivors are only recorded while the rows stream by and
        // are examined afterwards, all lanes of the warp together: the FP64 work of one lane
        // does not stall the other 31 in the middle of the filter loop.
        template <typename PS>
        __device__ __forceinline__ bool se3_valid_fast(const CheckerDesc &c, const float4 *F,
                                                       unsigned short *surv_base, PS st)
        {
            float px = (float)st[0], py = (float)st[1], pz = (float)st[2];
            const float qx = (float)st[3], qy = (float)st[4], qz = (float)st[5], qw = (float)st[6];
            const float s = qx * qx + qy * qy + qz * qz + qw * qw;
            if (!(fabsf(s - 1.0f) <= 2e-4f))
                return se3_valid(c, c.robot, c.sph, c.box, st);
            const float A = fmaxf(c.filt_amax, fmaxf(fabsf(px), fmaxf(fabsf(py), fabsf(pz)))) * 1.0001f;
            const float pn = px * px + py * py + pz * pz;
            // The four loop constants take a round trip through shared memory: ptxas otherwise
            // re-derives them from the FP64 state in every iteration of the filter loop.
            volatile float4 *scr = reinterpret_cast<volatile float4 *>(surv_base + SURV_CAP * blockDim.x) + threadIdx.x;
            scr->x = -2.0f * px;
            scr->y = -2.0f * py;
            scr->z = -2.0f * pz;
            scr->w = A * A * 1.52587890625e-05f - pn;
            const float mx = scr->x, my = scr->y, mz = scr->z, tau = scr->w;
            unsigned short *surv = surv_base + threadIdx.x;
            const int stride = blockDim.x;
            // Broad phase: a uniform grid over the obstacles' bounding spheres.  Cell lists are
            // built on the host with slack for the FP32 cell index; a state outside the grid is
            // farther from every obstacle than its sphere reaches.
            const float fx = (px - c.grid_x0) * c.grid_inv_h;
            const float fy = (py - c.grid_y0) * c.grid_inv_h;
            const float fz = (pz - c.grid_z0) * c.grid_inv_h;
            if (!(fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < (float)c.grid_nx && fy < (float)c.grid_ny &&
                  fz < (float)c.grid_nz))
                return true;
            const int cell = ((int)fz * c.grid_ny + (int)fy) * c.grid_nx + (int)fx;
            const uint32_t kb = __ldg(c.cell_start + cell), ke = __ldg(c.cell_start + cell + 1);
            int ns = 0;
            bool hit = false;
            for (uint32_t k = kb; k < ke; ++k)
            {
                const int jj = (int)__ldg(c.cell_items + k);
                const float4 o = F[jj];
                const float sv = fmaf(mz, o.z, fmaf(my, o.y, fmaf(mx, o.x, o.w)));
                if (sv <= tau)
                {
                    if (ns < SURV_CAP)
                        surv[(ns++) * stride] = (unsigned short)jj;
                    else if (!hit)
                    {
                        // list full (dense clutter): examine this one right away
                        for (int i = 0; i < c.nrobot && !hit; ++i)
                        {
                            double cx, cy, cz, rr;
                            se3_robot_centre(c, i, st, cx, cy, cz, rr);
                            hit = se3_obstacle_hits(c, jj, cx, cy, cz, rr);
                        }
                    }
                }
            }
            if (hit)
                return false;
            if (ns == 0)
                return true;
            // Narrow phase, FP32 first.  The FP32 vector v_f the distance is taken of (obstacle centre minus robot
            // sphere centre; for a box the clamped excess over the half extents, a 1-Lipschitz map of it) differs from
            // the vector v of the FP64 reference expressions by less than 90 u A per coordinate (u = 2^-24; the centre
            // c = p + (o + qw t + q x t), t = 2 q x o: 80 u |o| + 2 u A with p, q, o rounded to FP32 and |o| <= rho_rob
            // <= A; 3 u A for the difference, 2 u A more for a box), so | |v_f| - |v| | < e := 2^-15 A = 512 u A, more
            // than three times sqrt(3) 90 u A.  With D_f = fl(|v_f|^2) = |v_f|^2 (1 +- 4 u), R_f the FP32 sum of the
            // radii = R (1 +- 2 u) and T_f = fl(R_f^2), take the band
            //     m = 2.2 R_f e + 1.2 e^2 + 40 u T_f   >=   (R_f (1 + 16 u) + 1.04 e)^2 - R_f^2 .
            // D_f < T_f - m gives sqrt(D_f) < R_f (1 - 16 u) - 1.04 e, hence |v| <= sqrt(D_f) (1 + 2 u) + e < R_f (1 - 13 u)
            // < R: a hit in FP64 too.  D_f > T_f + m gives sqrt(D_f) > (R_f (1 + 16 u) + 1.04 e) (1 - u), hence
            // |v| >= sqrt(D_f) (1 - 2 u) - e > R_f (1 + 12 u) > R: a miss in FP64 too.  (The FP64 expressions' own
            // roundings, 2^-53 relative, disappear in the slack.)  Inside the band the FP64 expressions decide.  The
            // band is per pair and proportional to R_f: 4e-6 for a 5 cm pair in the benchmark world, where one width
            // for all pairs from |v| <= 2 A (2^-15 A^2 = 4.4e-5, the same size as the smallest thresholds) sent a
            // quarter of the states into the FP64 expressions.
            const float eband = A * 3.0517578125e-05f;
            const float e2band = 1.2f * eband * eband;
            for (int i = 0; i < c.nrobot; ++i)
            {
                const float4 ro = __ldg(c.robot32 + i);
                const float tx = 2.0f * (qy * ro.z - qz * ro.y);
                const float ty = 2.0f * (qz * ro.x - qx * ro.z);
                const float tz = 2.0f * (qx * ro.y - qy * ro.x);
                const float cxf = px + (ro.x + qw * tx + (qy * tz - qz * ty));
                const float cyf = py + (ro.y + qw * ty + (qz * tx - qx * tz));
                const float czf = pz + (ro.z + qw * tz + (qx * ty - qy * tx));
                unsigned amb = 0;
                for (int k = 0; k < ns; ++k)
                {
                    const int j = (int)surv[k * stride];
                    float D, T, m2;
                    if (j < c.nsph)
                    {
                        const float4 o = __ldg(c.sph32 + j);
                        const float dx = o.x - cxf, dy = o.y - cyf, dz = o.z - czf;
                        const float R = o.w + ro.w;
                        D = dx * dx + dy * dy + dz * dz;
                        T = R * R;
                        m2 = fmaf(2.2f * R, eband, fmaf(T, 2.384185791015625e-06f, e2band));
                    }
                    else
                    {
                        const float4 b0 = __ldg(c.box32 + 2 * (j - c.nsph));
                        const float4 b1 = __ldg(c.box32 + 2 * (j - c.nsph) + 1);
                        const float ex = fmaxf(fabsf(cxf - b0.x) - b0.w, 0.f);
                        const float ey = fmaxf(fabsf(cyf - b0.y) - b1.x, 0.f);
                        const float ez = fmaxf(fabsf(czf - b0.z) - b1.y, 0.f);
                        D = ex * ex + ey * ey + ez * ez;
                        T = ro.w * ro.w;
                        m2 = fmaf(2.2f * ro.w, eband, fmaf(T, 2.384185791015625e-06f, e2band));
                    }
                    if (D < T - m2)
                        return false;  // certain hit
                    if (!(D > T + m2))
                        amb |= 1u << k;  // too close to call in FP32 (or NaN)
                }
                if (amb)
                {
                    double cx, cy, cz, rr;
                    se3_robot_centre(c, i, st, cx, cy, cz, rr);
                    for (int k = 0; k < ns; ++k)
                        if ((amb >> k) & 1u)
                            if (se3_obstacle_hits(c, (int)surv[k * stride], cx, cy, cz, rr))
                                return false;
                }
            }
            return true;
        }

        // Shared-memory layout of the SE3 kernels: filter rows, then the survivor lists.
        template <int KIND>
        struct Staged
        {
            const float4 *F;
            unsigned short *surv;
        };

        template <int KIND>
        __device__ __forceinline__ Staged<KIND> stage_filter(const CheckerDesc &c, float4 *smem)
        {
            Staged<KIND> r{nullptr, nullptr};
            if (KIND != OMPLCU_CHK_SE3 || !c.stage_smem)
                return r;
            for (int i = threadIdx.x; i < c.nfilt_pad; i += blockDim.x)
                smem[i] = c.filt[i];
            __syncthreads();
            r.F = smem;
            r.surv = reinterpret_cast<unsigned short *>(smem + c.nfilt_pad);
            return r;
        }

        template <int KIND, typename PS>
        __device__ __forceinline__ bool state_valid_f(const CheckerDesc &c, const Staged<KIND> &sm, PS s)
        {
            if (KIND == OMPLCU_CHK_SE3)
            {
                if (c.stage_smem)
                    return se3_valid_fast(c, sm.F, sm.surv, s);
                return se3_valid(c, c.robot, c.sph, c.box, s);  // world too large for shared memory
            }
            return state_valid<KIND>(c, c.robot, c.sph, c.box, s);
        }

        template <int KIND, int MAXS>
        __global__ void __launch_bounds__(128) check_states_kernel(CheckerDesc c, int stride, const double *states,
                                                                   size_t n, uint8_t *valid)
        {
            extern __shared__ float4 smem_filt[];
            const Staged<KIND> sm = stage_filter<KIND>(c, smem_filt);
            for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (size_t)blockDim.x)
            {
                double st[MAXS];
                for (int d = 0; d < stride && d < MAXS; ++d)
                    st[d] = states[i * stride + d];
                valid[i] = state_valid_f<KIND>(c, sm, st) ? 1 : 0;
            }
        }

        // nd and the number of states to check per segment: (nd-1 interior if nd >= 2) + s2
        // prep (optional): the per-segment constants of SO3StateSpace::interpolate (so3_prepare) of the first
        // MAX_SO3_PREP SO(3) components, [component][segment] as (theta, d with the sign of `neg`): one thread per
        // segment here, all lanes busy, instead of a lane group of motion_group_kernel repeating them while the
        // other groups of its warp wait.
        __global__ void segment_count_kernel(SpaceDesc sp, const double *s1, const double *s2, size_t n, uint32_t *nd,
                                             unsigned long long *cnt, double2 *prep)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            const unsigned v = space_segment_count(sp, s1 + i * sp.stride, s2 + i * sp.stride);
            nd[i] = v;
            if (cnt)
                cnt[i] = (v >= 2 ? (unsigned long long)(v - 1) : 0ull) + 1ull;
            if (prep && v >= 2)
            {
                So3Seg g[MAX_SO3_PREP] = {};
                space_prepare(sp, s1 + i * sp.stride, s2 + i * sp.stride, g);
                int k = 0;
                for (int c = 0; c < sp.ncomp && k < MAX_SO3_PREP; ++c)
                    if (sp.comp[c].type == OMPLCU_SO3)
                    {
                        prep[(size_t)k * n + i] = make_double2(g[k].theta, g[k].neg ? -g[k].d : g[k].d);
                        ++k;
                    }
            }
        }

        // ctr[0]: next segment to hand out, ctr[1]: states evaluated (statistics)
        // SPEC = 1: the space is exactly SE(3) (RealVector(3) x SO(3), stride 7).  The state is then built by straight
        // code on constant indices and lives in registers; the general path walks the component table with run-time
        // offsets, which puts the state into local memory (86 STL / 22 LDL in the SASS of the general kernel).
        template <int KIND, int MAXS, int G, int SPEC>
        __global__ void __launch_bounds__(128, SPEC == 1 ? 8 : 1) motion_group_kernel(CheckerDesc c, SpaceDesc sp, const double *s1,
                                                                   const double *s2, size_t n, const uint32_t *nd,
                                                                   const double2 *prep_in, int nprep, int spread,
                                                                   unsigned *code, unsigned long long *ctr)
        {
            extern __shared__ float4 smem_filt[];
            const Staged<KIND> sm = stage_filter<KIND>(c, smem_filt);
            constexpr unsigned FULL = 0xffffffffu;
            const unsigned lane = threadIdx.x & 31u, gl = lane % G, gbase = lane - gl;
            const unsigned gmask = (G == 32) ? FULL : (((1u << G) - 1u) << gbase);
            size_t seg = 0;
            unsigned cnt = 0, base = 0, ndv = 0;
            // spread (the bool overload: any failing state decides, DiscreteMotionValidator.cpp:106-134 visits them in
            // bisection order for the same reason): lane gl owns the gl-th block of blk consecutive items and walks it
            // in steps of ~0.62 blk, so the G states of a pass, and those of successive passes, lie far apart on the
            // segment.  Neighbouring states fail together; states far apart fail independently, and a colliding
            // segment is found out in fewer passes.
            unsigned blk = 0, off = 0, stepc = 1;
            So3Seg prep[MAX_SO3_PREP] = {};
            unsigned long long evaluated = 0;
            bool need = true, done = false;
            while (true)
            {
                const bool fetch = need && !done;
                unsigned long long sidx = 0;
                if (fetch && gl == 0)
                    sidx = atomicAdd(&ctr[0], 1ull);
                sidx = __shfl_sync(FULL, sidx, gbase);
                if (fetch)
                {
                    if (sidx >= n)
                        done = true;
                    else
                    {
                        seg = (size_t)sidx;
                        ndv = nd[seg];
                        cnt = (ndv >= 2 ? ndv - 1 : 0u) + 1u;
                        base = 0;
                        need = false;
                        if (spread)
                        {
                            blk = (cnt + G - 1) / G;
                            off = 0;
                            stepc = (blk * 5u + 4u) / 8u;
                            if (stepc == 0)
                                stepc = 1;
                            for (;; ++stepc)  // coprime with blk: the walk visits every item of the block
                            {
                                unsigned x = stepc, y = blk;
                                while (y)
                                {
                                    const unsigned r = x % y;
                                    x = y;
                                    y = r;
                                }
                                if (x == 1 || blk == 1)
                                    break;
                            }
                        }
                        if (cnt > 1)  // acos, 1/sin(theta): once per segment, not once per state
                        {
                            if (SPEC == 1)
                            {
                                if (prep_in)
                                {
                                    const double2 g = prep_in[seg];
                                    prep[0].theta = g.x;
                                    prep[0].d = fabs(g.y);
                                    prep[0].neg = g.y < 0.0;
                                }
                                else
                                    prep[0] = so3_prepare(s1 + seg * 7 + 3, s2 + seg * 7 + 3);
                            }
                            else if (prep_in)
                            {
                                // computed by segment_count_kernel
                                for (int k = 0; k < MAX_SO3_PREP; ++k)
                                    if (k < nprep)
                                    {
                                        const double2 g = prep_in[(size_t)k * n + seg];
                                        prep[k].theta = g.x;
                                        prep[k].d = fabs(g.y);
                                        prep[k].neg = g.y < 0.0;
                                    }
                            }
                            else
                                space_prepare(sp, s1 + seg * sp.stride, s2 + seg * sp.stride, prep);
                        }
                    }
                }
                if (__all_sync(FULL, done))
                    break;
                // item m < cnt-1 is the interior state j = m+1, item cnt-1 is s2
                const unsigned m = spread ? gl * blk + off : base + gl;
                bool bad = false;
                if (!done && m < cnt)
                {
                    const double *a = s1 + seg * sp.stride, *b = s2 + seg * sp.stride;
                    double st[MAXS];
                    if (SPEC == 1)
                    {
                        if (m == cnt - 1)
                        {
#pragma unroll
                            for (int d = 0; d < 7; ++d)
                                st[d] = b[d];
                        }
                        else
                        {
                            const double t = ddiv((double)(m + 1), (double)ndv);
                            // RealVectorStateSpace::interpolate, then SO3StateSpace::interpolate: the expressions of
                            // space_interpolate_prepared for this component table
#pragma unroll
                            for (int d = 0; d < 3; ++d)
                                st[d] = dadd(a[d], dmul(dsub(b[d], a[d]), t));
                            so3_interpolate_prepared(prep[0], a + 3, b + 3, t, st + 3);
                        }
                    }
                    else if (m == cnt - 1)
                    {
                        for (int d = 0; d < sp.stride && d < MAXS; ++d)
                            st[d] = b[d];
                    }
                    else
                    {
                        // t = (double)j / (double)nd, DiscreteMotionValidator.cpp:63,121
                        const double t = ddiv((double)(m + 1), (double)ndv);
                        space_interpolate_prepared(sp, prep, a, b, t, st);
                    }
                    bad = !state_valid_f<KIND>(c, sm, st);
                    evaluated++;
                }
                const unsigned hit = __ballot_sync(FULL, bad) & gmask;
                if (!done)
                {
                    if (hit)
                    {
                        const unsigned first = base + (unsigned)(__ffs(hit) - 1) - gbase;
                        if (gl == 0)  // spread: only "not valid" is read (motion_finalize_kernel without last_t)
                            code[seg] = spread ? 1u : ((first == cnt - 1) ? CODE_S2 : first + 1);
                        need = true;
                    }
                    else
                    {
                        if (spread)
                        {
                            base += 1;
                            off += stepc;
                            if (off >= blk)
                                off -= blk;
                        }
                        else
                            base += G;
                        if (base >= (spread ? blk : cnt))
                        {
                            if (gl == 0)
                                code[seg] = CODE_VALID;
                            need = true;
                        }
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1)
                evaluated += __shfl_down_sync(FULL, evaluated, o);
            if (lane == 0 && evaluated)
                atomicAdd(&ctr[1], evaluated);
        }

        // valid flag, lastValid.second and lastValid.first per segment
        __global__ void motion_finalize_kernel(SpaceDesc sp, const double *s1, const double *s2, size_t n,
                                               const uint32_t *nd, const unsigned *code, uint8_t *valid,
                                               double *last_t, double *last_state)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            const unsigned cd = code[i];
            valid[i] = (cd == CODE_VALID) ? 1 : 0;
            if (!last_t)
                return;
            double t = 1.0;
            if (cd != CODE_VALID)
            {
                const int ndi = (int)nd[i];
                // (double)(j - 1) / (double)nd resp. (double)(nd - 1) / (double)nd,
                // DiscreteMotionValidator.cpp:66,81 (int arithmetic, then FP64 division)
                const int num = (cd == CODE_S2) ? (ndi - 1) : ((int)cd - 1);
                t = ddiv((double)num, (double)ndi);
                if (last_state)
                    space_interpolate(sp, s1 + i * sp.stride, s2 + i * sp.stride, t, last_state + i * sp.stride);
            }
            last_t[i] = t;
        }

        __global__ void space_distance_kernel(SpaceDesc sp, const double *a, const double *b, size_t n, double *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i < n)
                out[i] = space_distance(sp, a + i * sp.stride, b + i * sp.stride);
        }

        __global__ void space_interpolate_kernel(SpaceDesc sp, const double *a, const double *b, const double *t,
                                                 size_t n, double *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i < n)
                space_interpolate(sp, a + i * sp.stride, b + i * sp.stride, t[i], out + i * sp.stride);
        }

        inline unsigned nblk(size_t n, unsigned bs)
        {
            return (unsigned)((n + bs - 1) / bs);
        }
    }  // namespace

    // ==================================================================
    // host side
    // ==================================================================
    // ---- scratch of omplcu_check_motions_device_noalloc (no counterpart in the reference; added
    // for callers that cannot call cudaMalloc) ----
    // Layout: [0, 16) the two work counters of motion_group_kernel, [256, 256 + 4 M) segment counts,
    // [256 + 4 M, 256 + 8 M) first-failure codes, M = OMPLCU_NOALLOC_MAX_SEGMENTS.  Every GPU that loads
    // the library holds these 2 MiB + 256 B until exit; all checkers and calls on a device share them.
    constexpr size_t NOALLOC_HEAD = 256;
    constexpr size_t NOALLOC_BYTES = NOALLOC_HEAD + 8 * (size_t)OMPLCU_NOALLOC_MAX_SEGMENTS;
    static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

    namespace
    {
        // In host code the array's name is a host-side stand-in, not a device address: the address a
        // kernel may be given comes from cudaGetSymbolAddress, asked once per device with that device
        // current (the caller holds a DeviceGuard).
        unsigned char *noalloc_scratch_address(int device)
        {
            static std::mutex mu;
            static unsigned char *addr[64] = {};
            std::lock_guard<std::mutex> lk(mu);
            REQUIRE(device >= 0 && device < 64, "device index out of range");
            if (!addr[device])
            {
                void *p = nullptr;
                CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
                REQUIRE(p != nullptr, "cudaGetSymbolAddress returned NULL for the no-alloc scratch");
                addr[device] = static_cast<unsigned char *>(p);
            }
            return addr[device];
        }
    }  // namespace

    Checker::Checker(int device, const CheckerDesc &proto, const double *env, const double *robot, const double *sph,
                     const double *box, const double *circ)
      : device_(device), desc_(proto)
    {
        DeviceGuard g(device_);
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        noalloc_scratch_ = noalloc_scratch_address(device_);  // set-up: the calls themselves only use the pointer
        const size_t n_env = 4 * (size_t)desc_.nenv, n_rob = 4 * (size_t)desc_.nrobot, n_sph = 4 * (size_t)desc_.nsph,
                     n_box = 6 * (size_t)desc_.nbox, n_circ = 3 * (size_t)desc_.ncirc;
        const size_t tot = n_env + n_rob + n_sph + n_box + n_circ;
        params_.reserve(std::max<size_t>(tot, 1));
        double *p = params_.p;
        auto put = [&](const double *src, size_t cnt) -> const double * {
            const double *at = p;
            if (cnt)
                CUDA_CHECK(cudaMemcpyAsync(p, src, cnt * sizeof(double), cudaMemcpyHostToDevice, stream_));
            p += cnt;
            return at;
        };
        desc_.env = put(env, n_env);
        desc_.robot = put(robot, n_rob);
        desc_.sph = put(sph, n_sph);
        desc_.box = put(box, n_box);
        desc_.circ = put(circ, n_circ);
        CUDA_CHECK(cudaStreamSynchronize(stream_));
        smem_bytes_ = 0;
        desc_.stage_smem = 0;
        desc_.filt = nullptr;
        desc_.nfilt_pad = 0;
        desc_.filt_amax = 0.f;
        if (desc_.kind == OMPLCU_CHK_SE3)
            build_se3_filter(robot, sph, box);
        int dev_sms = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, device_));
        sms_ = dev_sms;
        ctr_.reserve(2);
    }

    // Bounding-sphere rows for se3_valid_fast (see the comment there).  Everything is
    // computed in FP64 and rounded so that the FP32 row can only widen the sphere.
    void Checker::build_se3_filter(const double *robot, const double *sph, const double *box)
    {
        double rho_rob = 0.0;
        for (int i = 0; i < desc_.nrobot; ++i)
        {
            const double *o = robot + 4 * i;
            rho_rob = std::max(rho_rob, std::sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]) + std::fabs(o[3]));
        }
        rho_rob *= 1.0 + 1.0 / 1024.0;
        const int nf = desc_.nsph + desc_.nbox;
        const int pad = (nf + 3) / 4 * 4;
        std::vector<float> rows(4 * (size_t)std::max(pad, 4));
        std::vector<double> centre(3 * (size_t)std::max(nf, 1)), reach(std::max(nf, 1));
        double amax = 0.0;
        for (int j = 0; j < pad; ++j)
        {
            float *r = rows.data() + 4 * (size_t)j;
            if (j >= nf)
            {
                r[0] = r[1] = r[2] = 0.f;
                r[3] = std::numeric_limits<float>::infinity();  // never passes the filter
                continue;
            }
            const double *o = j < desc_.nsph ? sph + 4 * j : box + 6 * (j - desc_.nsph);
            double rho;
            if (j < desc_.nsph)
                rho = std::fabs(o[3]);
            else
                rho = std::sqrt(o[3] * o[3] + o[4] * o[4] + o[5] * o[5]);
            const double w = (rho + rho_rob) * (rho + rho_rob) * (1.0 + 1e-6);
            r[0] = (float)o[0];
            r[1] = (float)o[1];
            r[2] = (float)o[2];
            r[3] = (float)(o[0] * o[0] + o[1] * o[1] + o[2] * o[2] - w);
            for (int a = 0; a < 3; ++a)
                centre[3 * (size_t)j + a] = o[a];
            reach[j] = std::sqrt(w);
            for (double v : {std::fabs(o[0]), std::fabs(o[1]), std::fabs(o[2]), std::sqrt(w)})
                amax = std::max(amax, v);
        }
        filt_.reserve(rows.size() / 4);
        CUDA_CHECK(cudaMemcpyAsync(filt_.p, rows.data(), rows.size() * sizeof(float), cudaMemcpyHostToDevice, stream_));
        CUDA_CHECK(cudaStreamSynchronize(stream_));
        desc_.filt = filt_.p;
        desc_.nfilt_pad = pad;
        desc_.filt_amax = (float)(amax * (1.0 + 1e-6));
        build_se3_grid(centre, reach, nf);
        {
            // FP32 copies of the robot and the obstacles for the narrow phase
            std::vector<float> n32(4 * (size_t)(desc_.nrobot + desc_.nsph + 2 * desc_.nbox) + 4, 0.f);
            float *q = n32.data();
            for (int i = 0; i < 4 * desc_.nrobot; ++i)
                *q++ = (float)robot[i];
            for (int i = 0; i < 4 * desc_.nsph; ++i)
                *q++ = (float)sph[i];
            for (int j = 0; j < desc_.nbox; ++j)
            {
                const double *b = box + 6 * j;
                const float v[8] = {(float)b[0], (float)b[1], (float)b[2], (float)b[3], (float)b[4], (float)b[5], 0.f, 0.f};
                for (float x : v)
                    *q++ = x;
            }
            narrow32_.reserve(n32.size() / 4);
            CUDA_CHECK(cudaMemcpyAsync(narrow32_.p, n32.data(), n32.size() * sizeof(float), cudaMemcpyHostToDevice,
                                       stream_));
            CUDA_CHECK(cudaStreamSynchronize(stream_));
            desc_.robot32 = narrow32_.p;
            desc_.sph32 = narrow32_.p + desc_.nrobot;
            desc_.box32 = narrow32_.p + desc_.nrobot + desc_.nsph;
        }
        // filter rows + [SURV_CAP][128] u16 survivor lists + one float4 of scratch per thread
        const size_t need = (size_t)pad * sizeof(float4) + SURV_CAP * 128 * sizeof(unsigned short) + 128 * sizeof(float4);
        if (need <= 200 * 1024 && nf <= 65535)
        {
            smem_bytes_ = need;
            desc_.stage_smem = 1;
        }
    }

    // Uniform grid over the bounding spheres (centre o_j, radius sqrt(W_j) = rho_j + rho_rob).
    // Cell (ix,iy,iz) lists every obstacle whose sphere comes within `slack` of the cell's
    // box; the slack covers the FP32 evaluation of the cell index on the device (error
    // below 2^-21 * A in length units, A = largest coordinate).
    void Checker::build_se3_grid(const std::vector<double> &centre, const std::vector<double> &R, int nf)
    {
        double lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}, meanR = 0.0, amax = desc_.filt_amax;
        for (int j = 0; j < nf; ++j)
        {
            const double *r = centre.data() + 3 * (size_t)j;
            meanR += R[j] / nf;
            for (int a = 0; a < 3; ++a)
            {
                lo[a] = j ? std::min(lo[a], r[a] - R[j]) : r[a] - R[j];
                hi[a] = j ? std::max(hi[a], r[a] + R[j]) : r[a] + R[j];
            }
        }
        double ext = 0.0;
        for (int a = 0; a < 3; ++a)
            ext = std::max(ext, hi[a] - lo[a]);
        double h = std::max({1.5 * meanR, ext / 48.0, 1e-30});
        std::vector<uint32_t> start;
        std::vector<uint16_t> items;
        int n[3] = {1, 1, 1};
        double slack = 0.0;
        for (int attempt = 0; attempt < 8; ++attempt, h *= 2.0)
        {
            slack = 1e-3 * h + amax * 9.5367431640625e-07;  // 2^-20 * A
            for (int a = 0; a < 3; ++a)
                n[a] = std::max(1, (int)std::ceil((hi[a] - lo[a] + 2.0 * slack) / h));
            const size_t ncell = (size_t)n[0] * n[1] * n[2];
            std::vector<std::vector<uint16_t>> cells(ncell);
            size_t total = 0;
            for (int j = 0; j < nf; ++j)
            {
                const double *r = centre.data() + 3 * (size_t)j;
                int c0[3], c1[3];
                for (int a = 0; a < 3; ++a)
                {
                    c0[a] = std::max(0, (int)std::floor((r[a] - R[j] - slack - (lo[a] - slack)) / h));
                    c1[a] = std::min(n[a] - 1, (int)std::floor((r[a] + R[j] + slack - (lo[a] - slack)) / h));
                }
                for (int iz = c0[2]; iz <= c1[2]; ++iz)
                    for (int iy = c0[1]; iy <= c1[1]; ++iy)
                        for (int ix = c0[0]; ix <= c1[0]; ++ix)
                        {
                            // squared distance from the sphere centre to the cell's box
                            const int ic[3] = {ix, iy, iz};
                            double d2 = 0.0;
                            for (int a = 0; a < 3; ++a)
                            {
                                const double b0 = lo[a] - slack + ic[a] * h, b1 = b0 + h;
                                const double e = std::max({b0 - r[a], r[a] - b1, 0.0});
                                d2 += e * e;
                            }
                            if (d2 <= (R[j] + slack) * (R[j] + slack))
                            {
                                cells[((size_t)iz * n[1] + iy) * n[0] + ix].push_back((uint16_t)j);
                                ++total;
                            }
                        }
            }
            if (total > (size_t)8 << 20 && attempt < 7)
                continue;  // too fine for this world: coarsen
            start.assign(ncell + 1, 0);
            items.clear();
            items.reserve(total);
            for (size_t cidx = 0; cidx < ncell; ++cidx)
            {
                start[cidx] = (uint32_t)items.size();
                items.insert(items.end(), cells[cidx].begin(), cells[cidx].end());
            }
            start[ncell] = (uint32_t)items.size();
            break;
        }
        cell_start_.reserve(start.size());
        cell_items_.reserve(std::max<size_t>(items.size(), 1));
        CUDA_CHECK(cudaMemcpyAsync(cell_start_.p, start.data(), start.size() * sizeof(uint32_t),
                                   cudaMemcpyHostToDevice, stream_));
        if (!items.empty())
            CUDA_CHECK(cudaMemcpyAsync(cell_items_.p, items.data(), items.size() * sizeof(uint16_t),
                                       cudaMemcpyHostToDevice, stream_));
        CUDA_CHECK(cudaStreamSynchronize(stream_));
        desc_.cell_start = cell_start_.p;
        desc_.cell_items = cell_items_.p;
        desc_.grid_x0 = (float)(lo[0] - slack);
        desc_.grid_y0 = (float)(lo[1] - slack);
        desc_.grid_z0 = (float)(lo[2] - slack);
        desc_.grid_inv_h = (float)(1.0 / h);
        desc_.grid_nx = n[0];
        desc_.grid_ny = n[1];
        desc_.grid_nz = n[2];
        grid_cells_ = (long)start.size() - 1;
        grid_items_ = (long)items.size();
    }

    Checker::~Checker()
    {
        DeviceGuard g(device_);
        if (stream_)
        {
            cudaStreamSynchronize(stream_);
            cudaStreamDestroy(stream_);
        }
        if (copy_stream_)
            cudaStreamDestroy(copy_stream_);
    }

    long Checker::get_stat(const char *key) const
    {
        if (!strcmp(key, "launches"))
            return launches_;
        if (!strcmp(key, "last_states"))
            return last_states_;
        if (!strcmp(key, "last_states_evaluated"))
            return last_evaluated_;
        if (!strcmp(key, "last_group"))
            return last_group_;
        if (!strcmp(key, "grid_cells"))
            return grid_cells_;
        if (!strcmp(key, "grid_items"))
            return grid_items_;
        throw Error(OMPLCU_ERR_INVALID, std::string("unknown stat ") + key);
    }

    namespace
    {
        template <int KIND, int MAXS>
        void launch_states(const CheckerDesc &c, size_t smem, int sms, int stride, const double *st, size_t n,
                           uint8_t *v, cudaStream_t s)
        {
            if (smem > 48 * 1024)
                CUDA_CHECK(cudaFuncSetAttribute(check_states_kernel<KIND, MAXS>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            const unsigned grid = std::min<unsigned>(nblk(n, 128), (unsigned)sms * 16u);
            check_states_kernel<KIND, MAXS><<<grid, 128, smem, s>>>(c, stride, st, n, v);
        }

        template <int KIND, int MAXS, int G, int SPEC>
        void launch_motion_g(const CheckerDesc &c, size_t smem, int sms, const SpaceDesc &sp, const double *s1,
                             const double *s2, size_t n, const uint32_t *nd, const double2 *prep, int nprep, int spread,
                             unsigned *code, unsigned long long *ctr, cudaStream_t s)
        {
            if (smem > 48 * 1024)
                CUDA_CHECK(cudaFuncSetAttribute(motion_group_kernel<KIND, MAXS, G, SPEC>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            // persistent grid: a multiple of the SM count, no more groups than segments
            const size_t groups_per_cta = 128 / G;
            const unsigned want = (unsigned)std::min<size_t>((n + groups_per_cta - 1) / groups_per_cta,
                                                             (size_t)sms * 8u);
            motion_group_kernel<KIND, MAXS, G, SPEC><<<std::max(want, 1u), 128, smem, s>>>(c, sp, s1, s2, n, nd, prep, nprep,
                                                                                     spread, code, ctr);
        }

        template <int KIND, int MAXS>
        void launch_motion(const CheckerDesc &c, size_t smem, int sms, int group, const SpaceDesc &sp,
                           const double *s1, const double *s2, size_t n, const uint32_t *nd, const double2 *prep,
                           int nprep, int spread, unsigned *code, unsigned long long *ctr, cudaStream_t s)
        {
            // exactly SE(3): RealVector(3) x SO(3), nothing else in the state
            constexpr bool can_spec = KIND == OMPLCU_CHK_SE3 && MAXS == 8;
            const bool spec = can_spec && sp.compound && sp.ncomp == 2 && sp.stride == 7 && sp.comp[0].type == OMPLCU_RV &&
                              sp.comp[0].dim == 3 && sp.comp[1].type == OMPLCU_SO3 && (prep == nullptr || nprep == 1) &&
                              getenv("OMPLCU_MOTION_GENERIC") == nullptr;
            if (can_spec && spec)
            {
                if (group >= 32)
                    launch_motion_g<KIND, MAXS, 32, can_spec ? 1 : 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
                else if (group >= 16)
                    launch_motion_g<KIND, MAXS, 16, can_spec ? 1 : 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
                else if (group >= 8)
                    launch_motion_g<KIND, MAXS, 8, can_spec ? 1 : 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
                else
                    launch_motion_g<KIND, MAXS, 4, can_spec ? 1 : 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
                return;
            }
            if (group >= 32)
                launch_motion_g<KIND, MAXS, 32, 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
            else if (group >= 16)
                launch_motion_g<KIND, MAXS, 16, 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
            else if (group >= 8)
                launch_motion_g<KIND, MAXS, 8, 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
            else
                launch_motion_g<KIND, MAXS, 4, 0>(c, smem, sms, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
        }

        template <int KIND>
        void states_by_stride(const CheckerDesc &c, size_t smem, int sms, int stride, const double *st, size_t n,
                              uint8_t *v, cudaStream_t s)
        {
            if (stride <= 8)
                launch_states<KIND, 8>(c, smem, sms, stride, st, n, v, s);
            else if (stride <= 32)
                launch_states<KIND, 32>(c, smem, sms, stride, st, n, v, s);
            else
                launch_states<KIND, MAX_STRIDE>(c, smem, sms, stride, st, n, v, s);
        }

        template <int KIND>
        void motion_by_stride(const CheckerDesc &c, size_t smem, int sms, int group, const SpaceDesc &sp,
                              const double *s1, const double *s2, size_t n, const uint32_t *nd, const double2 *prep,
                              int nprep, int spread, unsigned *code, unsigned long long *ctr, cudaStream_t s)
        {
            if (sp.stride <= 8)
                launch_motion<KIND, 8>(c, smem, sms, group, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
            else if (sp.stride <= 32)
                launch_motion<KIND, 32>(c, smem, sms, group, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
            else
                launch_motion<KIND, MAX_STRIDE>(c, smem, sms, group, sp, s1, s2, n, nd, prep, nprep, spread, code, ctr, s);
        }

        void check_stride(const CheckerDesc &c, int stride)
        {
            REQUIRE(stride >= 1 && stride <= MAX_STRIDE, "state stride out of range (1..128)");
            if (c.kind == OMPLCU_CHK_HYPERCUBE)
                REQUIRE(stride >= c.dim, "hypercube checker: state has fewer values than dim");
            if (c.kind == OMPLCU_CHK_CHAIN)
                REQUIRE(stride >= c.nlinks, "chain checker: state has fewer values than links");
            if (c.kind == OMPLCU_CHK_SE3)
                REQUIRE(stride >= 7, "se3 checker: state needs 7 values (x,y,z,qx,qy,qz,qw)");
            if (c.kind == OMPLCU_CHK_CIRCLES)
                REQUIRE(stride >= 2, "circles checker: state needs 2 values");
        }
    }  // namespace

    void Checker::check_states_device(int stride, const double *d_states, size_t n, uint8_t *d_valid, cudaStream_t s)
    {
        std::lock_guard<std::mutex> lk(mu_);
        DeviceGuard g(device_);
        if (n == 0)
            return;
        REQUIRE(d_states && d_valid, "omplcu_check_states: NULL buffer");
        check_stride(desc_, stride);
        switch (desc_.kind)
        {
            case OMPLCU_CHK_HYPERCUBE:
                states_by_stride<OMPLCU_CHK_HYPERCUBE>(desc_, 0, sms_, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_CHAIN:
                states_by_stride<OMPLCU_CHK_CHAIN>(desc_, 0, sms_, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_SE3:
                states_by_stride<OMPLCU_CHK_SE3>(desc_, smem_bytes_, sms_, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_CIRCLES:
                states_by_stride<OMPLCU_CHK_CIRCLES>(desc_, 0, sms_, stride, d_states, n, d_valid, s);
                break;
            default:
                states_by_stride<OMPLCU_CHK_ALLVALID>(desc_, 0, sms_, stride, d_states, n, d_valid, s);
        }
        launches_++;
        last_states_ = (long)n;
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(s));
    }

    void Checker::check_motions_device(const SpaceDesc &sp, const double *d_s1, const double *d_s2, size_t n,
                                       uint8_t *d_valid, double *d_last_t, double *d_last_state, cudaStream_t s)
    {
        std::lock_guard<std::mutex> lk(mu_);
        DeviceGuard g(device_);
        if (n == 0)
            return;
        REQUIRE(d_s1 && d_s2 && d_valid, "omplcu_check_motions: NULL buffer");
        REQUIRE(!(d_last_state && !d_last_t), "omplcu_check_motions: last_state needs last_t");
        check_stride(desc_, sp.stride);
        nd_.reserve(n);
        cnt_.reserve(n);
        offs_.reserve(1);
        code_.reserve(n);
        // the bool overload: states in spread order (motion_group_kernel)
        int spread = d_last_t == nullptr ? 1 : 0;
        if (const char *e = getenv("OMPLCU_MOTION_SPREAD"))  // tuning experiments only
            spread = spread && atoi(e) != 0;
        int nprep = 0;
        for (int i = 0; i < sp.ncomp; ++i)
            nprep += sp.comp[i].type == OMPLCU_SO3 ? 1 : 0;
        nprep = std::min(nprep, MAX_SO3_PREP);
        if (const char *e = getenv("OMPLCU_MOTION_PREP"))  // tuning experiments only: 0 = per lane group, as before
            if (atoi(e) == 0)
                nprep = 0;
        double2 *prep = nullptr;
        if (nprep)
        {
            prep_.reserve((size_t)nprep * n);
            prep = prep_.p;
        }
        segment_count_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd_.p, cnt_.p, prep);
        CUDA_CHECK(cudaGetLastError());
        size_t tmp = 0;
        CUDA_CHECK(cub::DeviceReduce::Sum(nullptr, tmp, cnt_.p, offs_.p, (int)n, s));
        cub_tmp_.reserve(tmp);
        CUDA_CHECK(cub::DeviceReduce::Sum(cub_tmp_.p, tmp, cnt_.p, offs_.p, (int)n, s));
        unsigned long long total = 0;
        CUDA_CHECK(cudaMemcpyAsync(&total, offs_.p, sizeof(total), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaMemsetAsync(ctr_.p, 0, 2 * sizeof(unsigned long long), s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        // Lanes per segment.  Measured on M3 (nd ~ 72): 32 lanes 2.46 ms, 16 lanes 1.64, 8 lanes 1.25,
        // 4 lanes 1.16 (uniform end points: most segments fail within their first few states, and a
        // pass evaluates all its lanes); PRM-like short edges: 2.55 / 2.38 / 2.33 / 2.52.  Narrow
        // groups waste fewer evaluations after the first failure and in the last pass of a segment,
        // and fewer lanes repeat the per-segment setup; 8 is the best all-rounder.
        const double mean = (double)total / (double)n;
        int group = mean >= 6.0 ? 8 : 4;
        // Feedback from the previous batch on this checker: when fewer than a quarter of the states of its segments had
        // to be looked at (a sampling planner's long motions through clutter end within their first states), the next
        // one takes the narrow groups: M3 1.06 -> 0.95 ms.  The results do not depend on the width.
        if (group == 8 && fb_states_ > 0 && fb_evaluated_ * 4 < fb_states_)
            group = 4;
        if (const char *e = getenv("OMPLCU_MOTION_GROUP"))  // tuning experiments only
            group = atoi(e);
        switch (desc_.kind)
        {
            case OMPLCU_CHK_HYPERCUBE:
                motion_by_stride<OMPLCU_CHK_HYPERCUBE>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd_.p, prep, nprep, spread, code_.p,
                                                       ctr_.p, s);
                break;
            case OMPLCU_CHK_CHAIN:
                motion_by_stride<OMPLCU_CHK_CHAIN>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd_.p, prep, nprep, spread, code_.p, ctr_.p,
                                                   s);
                break;
            case OMPLCU_CHK_SE3:
                motion_by_stride<OMPLCU_CHK_SE3>(desc_, smem_bytes_, sms_, group, sp, d_s1, d_s2, n, nd_.p, prep, nprep, spread, code_.p,
                                                 ctr_.p, s);
                break;
            case OMPLCU_CHK_CIRCLES:
                motion_by_stride<OMPLCU_CHK_CIRCLES>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd_.p, prep, nprep, spread, code_.p, ctr_.p,
                                                     s);
                break;
            default:
                motion_by_stride<OMPLCU_CHK_ALLVALID>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd_.p, prep, nprep, spread, code_.p,
                                                      ctr_.p, s);
        }
        CUDA_CHECK(cudaGetLastError());
        motion_finalize_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd_.p, code_.p, d_valid, d_last_t,
                                                            d_last_state);
        CUDA_CHECK(cudaGetLastError());
        unsigned long long ctr_host[2] = {0, 0};
        CUDA_CHECK(cudaMemcpyAsync(ctr_host, ctr_.p, sizeof(ctr_host), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        launches_ += 4;
        last_states_ = fb_states_ = (long)total;
        last_evaluated_ = fb_evaluated_ = (long)ctr_host[1];
        last_group_ = group;
    }

    // omplcu_check_motions_device_noalloc: the same three kernels as check_motions_device on the same
    // inputs, hence the same outputs, with the per-segment scratch in g_noalloc_scratch.  No cudaMalloc,
    // cudaFree, CUB temporary storage or DevBuf::reserve on this path.  The lane-group width is fixed at 8
    // (the other entry point picks 4 or 8 from the mean state count, which costs a reduction into
    // temporary storage; the results do not depend on it).  The array keeps the previous call's contents:
    // the work counters must start at zero and are cleared here; segment counts and codes of rows < n are
    // overwritten by segment_count_kernel / motion_group_kernel before they are read.  Calls that overlap
    // on one GPU race on the array.
    void Checker::check_motions_device_noalloc(const SpaceDesc &sp, const double *d_s1, const double *d_s2, size_t n,
                                               uint8_t *d_valid, double *d_last_t, double *d_last_state,
                                               cudaStream_t s)
    {
        std::lock_guard<std::mutex> lk(mu_);
        DeviceGuard g(device_);
        if (n == 0)
            return;
        REQUIRE(n <= (size_t)OMPLCU_NOALLOC_MAX_SEGMENTS,
                "omplcu_check_motions_device_noalloc: more than OMPLCU_NOALLOC_MAX_SEGMENTS (262144) segments in one "
                "call; split the batch or use omplcu_check_motions_device");
        REQUIRE(d_s1 && d_s2 && d_valid, "omplcu_check_motions: NULL buffer");
        REQUIRE(!(d_last_state && !d_last_t), "omplcu_check_motions: last_state needs last_t");
        REQUIRE(noalloc_scratch_ != nullptr, "no-alloc scratch address missing");
        check_stride(desc_, sp.stride);
        unsigned long long *ctr = reinterpret_cast<unsigned long long *>(noalloc_scratch_);
        uint32_t *nd = reinterpret_cast<uint32_t *>(noalloc_scratch_ + NOALLOC_HEAD);
        unsigned *code = reinterpret_cast<unsigned *>(noalloc_scratch_ + NOALLOC_HEAD +
                                                      4 * (size_t)OMPLCU_NOALLOC_MAX_SEGMENTS);
        CUDA_CHECK(cudaMemsetAsync(ctr, 0, 2 * sizeof(unsigned long long), s));
        segment_count_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd, nullptr, nullptr);
        CUDA_CHECK(cudaGetLastError());
        const int group = 8;
        const int spread = d_last_t == nullptr ? 1 : 0;
        switch (desc_.kind)
        {
            case OMPLCU_CHK_HYPERCUBE:
                motion_by_stride<OMPLCU_CHK_HYPERCUBE>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd, nullptr, 0, spread, code, ctr, s);
                break;
            case OMPLCU_CHK_CHAIN:
                motion_by_stride<OMPLCU_CHK_CHAIN>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd, nullptr, 0, spread, code, ctr, s);
                break;
            case OMPLCU_CHK_SE3:
                motion_by_stride<OMPLCU_CHK_SE3>(desc_, smem_bytes_, sms_, group, sp, d_s1, d_s2, n, nd, nullptr, 0, spread, code, ctr, s);
                break;
            case OMPLCU_CHK_CIRCLES:
                motion_by_stride<OMPLCU_CHK_CIRCLES>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd, nullptr, 0, spread, code, ctr, s);
                break;
            default:
                motion_by_stride<OMPLCU_CHK_ALLVALID>(desc_, 0, sms_, group, sp, d_s1, d_s2, n, nd, nullptr, 0, spread, code, ctr, s);
        }
        CUDA_CHECK(cudaGetLastError());
        motion_finalize_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd, code, d_valid, d_last_t,
                                                            d_last_state);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(s));
        launches_ += 3;
        last_group_ = group;
    }

    void Checker::check_states_host(int stride, const double *states, size_t n, uint8_t *valid)
    {
        if (n == 0)
            return;
        REQUIRE(states && valid, "omplcu_check_states: NULL buffer");
        REQUIRE(stride >= 1 && stride <= MAX_STRIDE, "state stride out of range (1..128)");
        DeviceGuard g(device_);
        // the staging buffers (st_a_, st_v_, ...) belong to this checker: ONE per-instance mutex for both
        // host entry points, so isValid and checkMotion racing on one checker cannot overwrite or
        // reallocate each other's staging, and checkers on other GPUs are not serialised
        std::lock_guard<std::mutex> hl(host_mu_);
        st_a_.reserve(n * stride);
        st_v_.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(st_a_.p, states, n * stride * sizeof(double), cudaMemcpyHostToDevice, stream_));
        check_states_device(stride, st_a_.p, n, st_v_.p, stream_);
        CUDA_CHECK(cudaMemcpyAsync(valid, st_v_.p, n, cudaMemcpyDeviceToHost, stream_));
        CUDA_CHECK(cudaStreamSynchronize(stream_));
    }

    void Checker::check_motions_host(const SpaceDesc &sp, const double *s1, const double *s2, size_t n, uint8_t *valid,
                                     double *last_t, double *last_state)
    {
        if (n == 0)
            return;
        REQUIRE(s1 && s2 && valid, "omplcu_check_motions: NULL buffer");
        REQUIRE(!(last_state && !last_t), "omplcu_check_motions: last_state needs last_t");
        DeviceGuard g(device_);
        // the staging buffers (st_a_, st_v_, ...) belong to this checker: ONE per-instance mutex for both
        // host entry points, so isValid and checkMotion racing on one checker cannot overwrite or
        // reallocate each other's staging, and checkers on other GPUs are not serialised
        std::lock_guard<std::mutex> hl(host_mu_);
        const size_t st = (size_t)sp.stride;
        st_a_.reserve(n * st);
        st_b_.reserve(n * st);
        st_v_.reserve(n);
        if (last_t)
            st_t_.reserve(n);
        if (last_state)
            st_ls_.reserve(n * st);
        // Large batches are cut into chunks: all host->device copies are queued on a second stream up
        // front (one event per chunk), so chunk c+1 crosses PCIe while chunk c is being checked.
        const size_t CH = (size_t)1 << 18;
        const size_t nch = (n + CH - 1) / CH;
        if (copy_stream_ == nullptr)
            CUDA_CHECK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
        std::vector<cudaEvent_t> ev(nch);
        for (size_t c = 0; c < nch; ++c)
        {
            const size_t o = c * CH, cn = std::min(CH, n - o);
            CUDA_CHECK(cudaEventCreateWithFlags(&ev[c], cudaEventDisableTiming));
            CUDA_CHECK(cudaMemcpyAsync(st_a_.p + o * st, s1 + o * st, cn * st * sizeof(double), cudaMemcpyHostToDevice,
                                       copy_stream_));
            CUDA_CHECK(cudaMemcpyAsync(st_b_.p + o * st, s2 + o * st, cn * st * sizeof(double), cudaMemcpyHostToDevice,
                                       copy_stream_));
            if (last_state)  // rows of valid segments stay as the caller initialised them: send them over first
                CUDA_CHECK(cudaMemcpyAsync(st_ls_.p + o * st, last_state + o * st, cn * st * sizeof(double),
                                           cudaMemcpyHostToDevice, copy_stream_));
            CUDA_CHECK(cudaEventRecord(ev[c], copy_stream_));
        }
        long tot_states = 0, tot_eval = 0;
        for (size_t c = 0; c < nch; ++c)
        {
            const size_t o = c * CH, cn = std::min(CH, n - o);
            CUDA_CHECK(cudaStreamWaitEvent(stream_, ev[c], 0));
            check_motions_device(sp, st_a_.p + o * st, st_b_.p + o * st, cn, st_v_.p + o,
                                 last_t ? st_t_.p + o : nullptr, last_state ? st_ls_.p + o * st : nullptr, stream_);
            tot_states += last_states_;
            tot_eval += last_evaluated_;
            CUDA_CHECK(cudaMemcpyAsync(valid + o, st_v_.p + o, cn, cudaMemcpyDeviceToHost, stream_));
            if (last_t)
                CUDA_CHECK(cudaMemcpyAsync(last_t + o, st_t_.p + o, cn * sizeof(double), cudaMemcpyDeviceToHost, stream_));
            if (last_state)
                CUDA_CHECK(cudaMemcpyAsync(last_state + o * st, st_ls_.p + o * st, cn * st * sizeof(double),
                                           cudaMemcpyDeviceToHost, stream_));
        }
        CUDA_CHECK(cudaStreamSynchronize(stream_));
        last_states_ = tot_states;  // statistics of the whole call, not of its last chunk
        last_evaluated_ = tot_eval;
        for (auto &e : ev)
            cudaEventDestroy(e);
    }

    // ---- batched space math ----
    namespace
    {
        struct TmpStream
        {
            cudaStream_t s = nullptr;
            TmpStream()
            {
                CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            }
            ~TmpStream()
            {
                if (s)
                    cudaStreamDestroy(s);
            }
        };
    }  // namespace

    void space_distance_host(const SpaceDesc &sp, int device, const double *a, const double *b, size_t n, double *out)
    {
        if (n == 0)
            return;
        REQUIRE(a && b && out, "omplcu_space_distance: NULL buffer");
        DeviceGuard g(device);
        TmpStream ts;
        DevBuf<double> da, db, dout;
        const size_t st = (size_t)sp.stride;
        da.reserve(n * st);
        db.reserve(n * st);
        dout.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(da.p, a, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(db.p, b, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        space_distance_kernel<<<nblk(n, 128), 128, 0, ts.s>>>(sp, da.p, db.p, n, dout.p);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, dout.p, n * sizeof(double), cudaMemcpyDeviceToHost, ts.s));
        CUDA_CHECK(cudaStreamSynchronize(ts.s));
    }

    void space_interpolate_host(const SpaceDesc &sp, int device, const double *a, const double *b, const double *t,
                                size_t n, double *out)
    {
        if (n == 0)
            return;
        REQUIRE(a && b && t && out, "omplcu_space_interpolate: NULL buffer");
        DeviceGuard g(device);
        TmpStream ts;
        DevBuf<double> da, db, dt, dout;
        const size_t st = (size_t)sp.stride;
        da.reserve(n * st);
        db.reserve(n * st);
        dt.reserve(n);
        dout.reserve(n * st);
        CUDA_CHECK(cudaMemcpyAsync(da.p, a, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(db.p, b, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(dt.p, t, n * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        space_interpolate_kernel<<<nblk(n, 128), 128, 0, ts.s>>>(sp, da.p, db.p, dt.p, n, dout.p);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, dout.p, n * st * sizeof(double), cudaMemcpyDeviceToHost, ts.s));
        CUDA_CHECK(cudaStreamSynchronize(ts.s));
    }

    namespace
    {
        // StateSpace::satisfiesBounds per component: RealVectorStateSpace.cpp:200-208 (+- DBL_EPSILON slack),
        // SO2StateSpace.cpp:90-94 ([-pi, pi)), SO3StateSpace.cpp:177-181,213-216 (| |q| - 1 | < 1e-9, with
        // the norm taken as exactly 1 when |q|^2 is within DBL_EPSILON of 1), CompoundStateSpace: all of them
        // (StateSpace.cpp:1021-1028).  The kinematic chain space is a RealVectorStateSpace (demos/KinematicChain.h:88).
        __global__ void satisfies_bounds_kernel(SpaceDesc sp, const double *states, size_t n, const double *low,
                                                const double *high, uint8_t *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            const double *s = states + i * sp.stride;
            const double eps = 2.220446049250313e-16;
            bool ok = true;
            int off = 0;
            for (int c = 0; c < sp.ncomp; ++c)
            {
                const omplcu_component &k = sp.comp[c];
                if (k.type == OMPLCU_RV || k.type == OMPLCU_CHAIN)
                {
                    for (int j = 0; j < k.dim; ++j)
                        if (dsub(s[off + j], eps) > high[off + j] || dadd(s[off + j], eps) < low[off + j])
                            ok = false;
                }
                else if (k.type == OMPLCU_SO2)
                    ok = ok && (s[off] < kPi) && (s[off] >= -kPi);
                else
                {
                    // quaternionNormSquared: x*x + y*y + z*z + w*w, left to right
                    const double n2 = dadd(dadd(dadd(dmul(s[off], s[off]), dmul(s[off + 1], s[off + 1])),
                                                dmul(s[off + 2], s[off + 2])),
                                           dmul(s[off + 3], s[off + 3]));
                    const double nrm = (fabs(dsub(n2, 1.0)) > eps) ? dsqrt(n2) : 1.0;
                    ok = ok && (fabs(dsub(nrm, 1.0)) < 1e-9);
                }
                off += k.dim;
            }
            out[i] = ok ? 1 : 0;
        }
    }  // namespace

    void space_satisfies_bounds_host(const SpaceDesc &sp, int device, const double *low, const double *high,
                                     const double *states, size_t n, uint8_t *out)
    {
        if (n == 0)
            return;
        REQUIRE(states && out && low && high, "omplcu_space_satisfies_bounds: NULL buffer");
        DeviceGuard g(device);
        TmpStream ts;
        DevBuf<double> ds, db;
        DevBuf<uint8_t> dout;
        const size_t st = (size_t)sp.stride;
        ds.reserve(n * st);
        db.reserve(2 * st);
        dout.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(ds.p, states, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(db.p, low, st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(db.p + st, high, st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        satisfies_bounds_kernel<<<nblk(n, 128), 128, 0, ts.s>>>(sp, ds.p, n, db.p, db.p + st, dout.p);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, dout.p, n, cudaMemcpyDeviceToHost, ts.s));
        CUDA_CHECK(cudaStreamSynchronize(ts.s));
    }

    void space_segment_count_host(const SpaceDesc &sp, int device, const double *a, const double *b, size_t n,
                                  uint32_t *out)
    {
        if (n == 0)
            return;
        REQUIRE(a && b && out, "omplcu_space_segment_count: NULL buffer");
        DeviceGuard g(device);
        TmpStream ts;
        DevBuf<double> da, db;
        DevBuf<uint32_t> dout;
        const size_t st = (size_t)sp.stride;
        da.reserve(n * st);
        db.reserve(n * st);
        dout.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(da.p, a, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        CUDA_CHECK(cudaMemcpyAsync(db.p, b, n * st * sizeof(double), cudaMemcpyHostToDevice, ts.s));
        segment_count_kernel<<<nblk(n, 128), 128, 0, ts.s>>>(sp, da.p, db.p, n, dout.p, nullptr, nullptr);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, dout.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ts.s));
        CUDA_CHECK(cudaStreamSynchronize(ts.s));
    }
}  // namespace omplcu
