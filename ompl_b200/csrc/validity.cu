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
// does not depend on the order and all states of all segments are checked in
// one flat launch: thread e of the grid owns one (segment, j) pair found by
// binary search in the exclusive prefix sum of per-segment state counts, and
// the per-segment outcome is the MINIMUM failing j (atomicMin), which is what
// the lastValid overload reports (DiscreteMotionValidator.cpp:60-72).
// All arithmetic is FP64 with the reference's operation order (common.cuh).
#include "validity.cuh"

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

        // copy the SE3 obstacle set into shared memory (all threads participate)
        template <int KIND>
        __device__ __forceinline__ void stage_obstacles(const CheckerDesc &c, double *smem, const double *&robot,
                                                        const double *&sph, const double *&box)
        {
            robot = c.robot;
            sph = c.sph;
            box = c.box;
            if (KIND == OMPLCU_CHK_SE3 && c.stage_smem)
            {
                double *r = smem, *sp = r + 4 * c.nrobot, *bx = sp + 4 * c.nsph;
                for (int i = threadIdx.x; i < 4 * c.nrobot; i += blockDim.x)
                    r[i] = c.robot[i];
                for (int i = threadIdx.x; i < 4 * c.nsph; i += blockDim.x)
                    sp[i] = c.sph[i];
                for (int i = threadIdx.x; i < 6 * c.nbox; i += blockDim.x)
                    bx[i] = c.box[i];
                __syncthreads();
                robot = r;
                sph = sp;
                box = bx;
            }
        }

        template <int KIND, int MAXS>
        __global__ void __launch_bounds__(128) check_states_kernel(CheckerDesc c, int stride, const double *states,
                                                                   size_t n, uint8_t *valid)
        {
            extern __shared__ double smem_obs[];
            const double *robot, *sph, *box;
            stage_obstacles<KIND>(c, smem_obs, robot, sph, box);
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            double st[MAXS];
            for (int d = 0; d < stride && d < MAXS; ++d)
                st[d] = states[i * stride + d];
            valid[i] = state_valid<KIND>(c, robot, sph, box, st) ? 1 : 0;
        }

        // nd and the number of states to check per segment: (nd-1 interior if nd >= 2) + s2
        __global__ void segment_count_kernel(SpaceDesc sp, const double *s1, const double *s2, size_t n, uint32_t *nd,
                                             unsigned long long *cnt, unsigned *code)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            const unsigned v = space_segment_count(sp, s1 + i * sp.stride, s2 + i * sp.stride);
            nd[i] = v;
            if (cnt)
                cnt[i] = (v >= 2 ? (unsigned long long)(v - 1) : 0ull) + 1ull;
            if (code)
                code[i] = CODE_VALID;
        }

        template <int KIND, int MAXS>
        __global__ void __launch_bounds__(128) motion_states_kernel(
            CheckerDesc c, SpaceDesc sp, const double *s1, const double *s2, size_t n, const uint32_t *nd,
            const unsigned long long *offs, unsigned long long total, unsigned *code, int want_first)
        {
            extern __shared__ double smem_obs[];
            const double *robot, *sph, *box;
            stage_obstacles<KIND>(c, smem_obs, robot, sph, box);
            const unsigned long long e = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
            if (e >= total)
                return;
            // segment = last index with offs[seg] <= e
            size_t lo = 0, hi = n;
            while (hi - lo > 1)
            {
                const size_t mid = (lo + hi) >> 1;
                if (offs[mid] <= e)
                    lo = mid;
                else
                    hi = mid;
            }
            const size_t seg = lo;
            const unsigned r = (unsigned)(e - offs[seg]);
            const unsigned my = (r == 0) ? CODE_S2 : r;
            // early exit: a failure that already decides this segment
            const unsigned cur = code[seg];
            if (want_first ? (cur <= my) : (cur != CODE_VALID))
                return;
            const double *a = s1 + seg * sp.stride, *b = s2 + seg * sp.stride;
            double st[MAXS];
            if (r == 0)
            {
                for (int d = 0; d < sp.stride && d < MAXS; ++d)
                    st[d] = b[d];
            }
            else
            {
                // t = (double)j / (double)nd, DiscreteMotionValidator.cpp:63,121
                const double t = ddiv((double)r, (double)nd[seg]);
                space_interpolate(sp, a, b, t, st);
            }
            if (!state_valid<KIND>(c, robot, sph, box, st))
                atomicMin(&code[seg], my);
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
    Checker::Checker(int device, const CheckerDesc &proto, const double *env, const double *robot, const double *sph,
                     const double *box, const double *circ)
      : device_(device), desc_(proto)
    {
        DeviceGuard g(device_);
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
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
        if (desc_.kind == OMPLCU_CHK_SE3)
        {
            const size_t need = (n_rob + n_sph + n_box) * sizeof(double);
            if (need <= 96 * 1024)
            {
                smem_bytes_ = need;
                desc_.stage_smem = 1;
            }
        }
    }

    Checker::~Checker()
    {
        DeviceGuard g(device_);
        if (stream_)
        {
            cudaStreamSynchronize(stream_);
            cudaStreamDestroy(stream_);
        }
    }

    long Checker::get_stat(const char *key) const
    {
        if (!strcmp(key, "launches"))
            return launches_;
        if (!strcmp(key, "last_states"))
            return last_states_;
        throw Error(OMPLCU_ERR_INVALID, std::string("unknown stat ") + key);
    }

    namespace
    {
        template <int KIND, int MAXS>
        void launch_states(const CheckerDesc &c, size_t smem, int stride, const double *st, size_t n, uint8_t *v,
                           cudaStream_t s)
        {
            if (smem > 48 * 1024)
                CUDA_CHECK(cudaFuncSetAttribute(check_states_kernel<KIND, MAXS>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            check_states_kernel<KIND, MAXS><<<nblk(n, 128), 128, smem, s>>>(c, stride, st, n, v);
        }

        template <int KIND, int MAXS>
        void launch_motion(const CheckerDesc &c, size_t smem, const SpaceDesc &sp, const double *s1, const double *s2,
                           size_t n, const uint32_t *nd, const unsigned long long *offs, unsigned long long total,
                           unsigned *code, int want_first, cudaStream_t s)
        {
            if (smem > 48 * 1024)
                CUDA_CHECK(cudaFuncSetAttribute(motion_states_kernel<KIND, MAXS>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            motion_states_kernel<KIND, MAXS>
                <<<nblk(total, 128), 128, smem, s>>>(c, sp, s1, s2, n, nd, offs, total, code, want_first);
        }

        template <int KIND>
        void states_by_stride(const CheckerDesc &c, size_t smem, int stride, const double *st, size_t n, uint8_t *v,
                              cudaStream_t s)
        {
            if (stride <= 8)
                launch_states<KIND, 8>(c, smem, stride, st, n, v, s);
            else if (stride <= 32)
                launch_states<KIND, 32>(c, smem, stride, st, n, v, s);
            else
                launch_states<KIND, MAX_STRIDE>(c, smem, stride, st, n, v, s);
        }

        template <int KIND>
        void motion_by_stride(const CheckerDesc &c, size_t smem, const SpaceDesc &sp, const double *s1,
                              const double *s2, size_t n, const uint32_t *nd, const unsigned long long *offs,
                              unsigned long long total, unsigned *code, int want_first, cudaStream_t s)
        {
            if (sp.stride <= 8)
                launch_motion<KIND, 8>(c, smem, sp, s1, s2, n, nd, offs, total, code, want_first, s);
            else if (sp.stride <= 32)
                launch_motion<KIND, 32>(c, smem, sp, s1, s2, n, nd, offs, total, code, want_first, s);
            else
                launch_motion<KIND, MAX_STRIDE>(c, smem, sp, s1, s2, n, nd, offs, total, code, want_first, s);
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
                states_by_stride<OMPLCU_CHK_HYPERCUBE>(desc_, 0, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_CHAIN:
                states_by_stride<OMPLCU_CHK_CHAIN>(desc_, 0, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_SE3:
                states_by_stride<OMPLCU_CHK_SE3>(desc_, smem_bytes_, stride, d_states, n, d_valid, s);
                break;
            case OMPLCU_CHK_CIRCLES:
                states_by_stride<OMPLCU_CHK_CIRCLES>(desc_, 0, stride, d_states, n, d_valid, s);
                break;
            default:
                states_by_stride<OMPLCU_CHK_ALLVALID>(desc_, 0, stride, d_states, n, d_valid, s);
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
        offs_.reserve(n + 1);
        code_.reserve(n);
        segment_count_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd_.p, cnt_.p, code_.p);
        CUDA_CHECK(cudaGetLastError());
        size_t tmp = 0;
        CUDA_CHECK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, cnt_.p, offs_.p, (int)n, s));
        cub_tmp_.reserve(tmp);
        CUDA_CHECK(cub::DeviceScan::ExclusiveSum(cub_tmp_.p, tmp, cnt_.p, offs_.p, (int)n, s));
        unsigned long long last_off = 0, last_cnt = 0;
        CUDA_CHECK(cudaMemcpyAsync(&last_off, offs_.p + (n - 1), sizeof(last_off), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaMemcpyAsync(&last_cnt, cnt_.p + (n - 1), sizeof(last_cnt), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        const unsigned long long total = last_off + last_cnt;
        REQUIRE(total < (1ull << 40), "omplcu_check_motions: too many interpolated states in one batch");
        const int want_first = d_last_t ? 1 : 0;
        // at most 2^31-1 blocks of 128 threads per launch: split very large batches
        switch (desc_.kind)
        {
            case OMPLCU_CHK_HYPERCUBE:
                motion_by_stride<OMPLCU_CHK_HYPERCUBE>(desc_, 0, sp, d_s1, d_s2, n, nd_.p, offs_.p, total, code_.p,
                                                       want_first, s);
                break;
            case OMPLCU_CHK_CHAIN:
                motion_by_stride<OMPLCU_CHK_CHAIN>(desc_, 0, sp, d_s1, d_s2, n, nd_.p, offs_.p, total, code_.p,
                                                   want_first, s);
                break;
            case OMPLCU_CHK_SE3:
                motion_by_stride<OMPLCU_CHK_SE3>(desc_, smem_bytes_, sp, d_s1, d_s2, n, nd_.p, offs_.p, total,
                                                 code_.p, want_first, s);
                break;
            case OMPLCU_CHK_CIRCLES:
                motion_by_stride<OMPLCU_CHK_CIRCLES>(desc_, 0, sp, d_s1, d_s2, n, nd_.p, offs_.p, total, code_.p,
                                                     want_first, s);
                break;
            default:
                motion_by_stride<OMPLCU_CHK_ALLVALID>(desc_, 0, sp, d_s1, d_s2, n, nd_.p, offs_.p, total, code_.p,
                                                      want_first, s);
        }
        CUDA_CHECK(cudaGetLastError());
        motion_finalize_kernel<<<nblk(n, 128), 128, 0, s>>>(sp, d_s1, d_s2, n, nd_.p, code_.p, d_valid, d_last_t,
                                                            d_last_state);
        CUDA_CHECK(cudaGetLastError());
        launches_ += 4;
        last_states_ = (long)total;
        CUDA_CHECK(cudaStreamSynchronize(s));
    }

    void Checker::check_states_host(int stride, const double *states, size_t n, uint8_t *valid)
    {
        if (n == 0)
            return;
        REQUIRE(states && valid, "omplcu_check_states: NULL buffer");
        REQUIRE(stride >= 1 && stride <= MAX_STRIDE, "state stride out of range (1..128)");
        DeviceGuard g(device_);
        static std::mutex host_mu;
        std::lock_guard<std::mutex> hl(host_mu);
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
        static std::mutex host_mu;
        std::lock_guard<std::mutex> hl(host_mu);
        const size_t st = (size_t)sp.stride;
        st_a_.reserve(n * st);
        st_b_.reserve(n * st);
        st_v_.reserve(n);
        if (last_t)
            st_t_.reserve(n);
        if (last_state)
            st_ls_.reserve(n * st);
        CUDA_CHECK(cudaMemcpyAsync(st_a_.p, s1, n * st * sizeof(double), cudaMemcpyHostToDevice, stream_));
        CUDA_CHECK(cudaMemcpyAsync(st_b_.p, s2, n * st * sizeof(double), cudaMemcpyHostToDevice, stream_));
        if (last_state)  // rows of valid segments stay as the caller initialised them: send them over first
            CUDA_CHECK(
                cudaMemcpyAsync(st_ls_.p, last_state, n * st * sizeof(double), cudaMemcpyHostToDevice, stream_));
        check_motions_device(sp, st_a_.p, st_b_.p, n, st_v_.p, last_t ? st_t_.p : nullptr,
                             last_state ? st_ls_.p : nullptr, stream_);
        CUDA_CHECK(cudaMemcpyAsync(valid, st_v_.p, n, cudaMemcpyDeviceToHost, stream_));
        if (last_t)
            CUDA_CHECK(cudaMemcpyAsync(last_t, st_t_.p, n * sizeof(double), cudaMemcpyDeviceToHost, stream_));
        if (last_state)
            CUDA_CHECK(
                cudaMemcpyAsync(last_state, st_ls_.p, n * st * sizeof(double), cudaMemcpyDeviceToHost, stream_));
        CUDA_CHECK(cudaStreamSynchronize(stream_));
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
