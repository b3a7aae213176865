// knn_scan.cuh -- the brute-force FP32 filter scan for k-nearest queries on
// plain RealVector spaces (RealVectorStateSpace::distance,
// base/spaces/src/RealVectorStateSpace.cpp:230-242 is the metric; the exact
// FP64 re-rank that reproduces it bit for bit is knn_rerank_kernel in nn.cu).
//
// What is computed.  Points and queries are recentred by a fixed vector c and
// rounded to FP32: pt = fl32(p - c), qt = fl32(q - c).  Per (query, point)
// pair the kernel evaluates, in FP32 with one rounding per FMA,
//        s = pn + sum_c (-2 qt_c) * pt_c ,   pn = fl32(|pt|^2)
// which approximates |qt - pt|^2 - |qt|^2.  Each thread owns RQ queries and
// keeps, per query, the L smallest s seen in its segment of the point set (an
// unsorted list in shared memory; the running maximum, tagged with its slot, is
// cached in a register as the insertion threshold).  The re-rank recomputes the exact
// distance of every listed point and proves, per query, that nothing outside
// the lists can belong to the k best (the "band check"); queries for which the
// proof fails are redone by the exact sort path.  So the scan only has to be
// fast, never exact.
//
// Mapping to B200.  The (query tile x point tile) rectangle is linearised and
// cut into gridDim.x equal pieces, one persistent CTA each (grid = 148 SMs x
// resident CTAs), so every SM gets the same number of pairs regardless of the
// query count.  Point tiles (PT rows of ROWF floats, 8 KB at d = 7) stream from L2/HBM
// into a 3-stage shared-memory ring with cp.async.bulk (TMA bulk copy)
// completing on mbarriers; all lanes of a warp read the same point row
// (LDS.128 broadcast) and pair it with their own queries held in registers as
// packed float2, so the inner loop is FFMA2 (fma.rn.f32x2: two queries per
// instruction with the point coordinate as the broadcast scalar operand).
// FP32-pipe bound: D FMAs per pair; see DESIGN.md for the roofline.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace omplcu
{
    constexpr int SCAN_PT = 256;       // points per shared-memory tile
    constexpr int SCAN_STAGES = 3;     // TMA ring depth
    constexpr int SCAN_THREADS = 128;  // 4 warps per CTA
    constexpr int SCAN_GP = 4;         // points per group in the inner loop (SCAN_PT % SCAN_GP == 0)
    constexpr uint32_t SLOT_NONE = 0xFFFFFFFFu;

    __host__ __device__ constexpr int scan_rowf(int D)
    {
        return ((D + 1 + 3) / 4) * 4;  // floats per filter row: D coords + norm, padded to 16 B
    }

    struct ScanParams
    {
        const float *filt;      // [units*SCAN_PT][ROWF]
        const float *qp;        // [D][qpad]  (-2 * qt_c)
        float *cand_thr;        // [smax][qpad]   final list threshold per (segment, query)
        uint32_t *cand_ids;     // [smax*L][qpad] SLOT_NONE where the list did not fill
        unsigned long long units;        // point tiles
        unsigned long long total_units;  // qtiles * units
        uint32_t qpad;
        uint32_t smax;
        const float *init_thr;  // [qpad] starting bound per query, or nullptr
    };

    // ---- PTX helpers: mbarrier + TMA bulk copy (cp.async.bulk) ----
    __device__ __forceinline__ uint32_t smem_u32(const void *p)
    {
        return static_cast<uint32_t>(__cvta_generic_to_shared(p));
    }
    __device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    }
    __device__ __forceinline__ void mbar_fence_init()
    {
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                     : "memory");
    }
    __device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
    {
        asm volatile("{\n"
                     ".reg .pred p;\n"
                     "WAIT_LOOP:\n"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                     "@p bra WAIT_DONE;\n"
                     "bra WAIT_LOOP;\n"
                     "WAIT_DONE:\n"
                     "}\n" ::"r"(smem_u32(bar)),
                     "r"(parity)
                     : "memory");
    }
    // global -> shared bulk copy; bytes and both addresses are multiples of 16
    __device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(dst)),
                     "l"(src), "r"(bytes), "r"(smem_u32(bar))
                     : "memory");
    }

    // d = a * b + c on two packed FP32 lanes (FFMA2 on sm_100)
    __device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
    {
        unsigned long long ra, rb, rc, rd;
        ra = *reinterpret_cast<unsigned long long *>(&a);
        rb = *reinterpret_cast<unsigned long long *>(&b);
        rc = *reinterpret_cast<unsigned long long *>(&c);
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
        return *reinterpret_cast<float2 *>(&rd);
    }

    // ---- candidate lists ----
    // Each query keeps the L smallest scan values of its segment in shared
    // memory: LS = L rounded up to 4 floats, contiguous per query so the whole
    // list is read back with LS/4 LDS.128.  A stored value carries its own slot
    // index in the low 6 mantissa bits and is rounded UP to make room for it
    // (stored >= true value > stored - 128 ulp), so one FMNMX3 tree over the
    // list yields both the new threshold and the slot to overwrite next.  The
    // threshold is therefore never below the true L-th smallest value (nothing
    // is wrongly rejected); eviction may pick an element within 128 ulp of the
    // true maximum, which the re-rank's band check accounts for (nn.cu).
    constexpr uint32_t TAG_MASK = 63u;
    constexpr uint32_t EMPTY_BITS = 0x7f000000u;  // 1.7e38 | slot: larger than any real scan value

    __host__ __device__ constexpr int scan_ls(int L)
    {
        return (L + 3) & ~3;
    }

    __device__ __forceinline__ float tag_up(float v, uint32_t slot)
    {
        const uint32_t b = __float_as_uint(v);
        const uint32_t m = b & 0x7fffffffu;
        const uint32_t pos = ((b + 64u) & ~TAG_MASK) | slot;                     // v >= +0
        const uint32_t neg = 0x80000000u | (((m - 64u) & ~TAG_MASK) | slot);     // v < 0, |v| not tiny
        const uint32_t r = (b & 0x80000000u) ? (m >= 128u ? neg : slot) : pos;
        return __uint_as_float(r);
    }

    template <int N>
    __device__ __forceinline__ float max_tree(const float (&x)[N])
    {
        float y[(N + 2) / 3];
#pragma unroll
        for (int i = 0; i < (N + 2) / 3; ++i)
        {
            float m = x[3 * i];
            if (3 * i + 1 < N)
                m = fmaxf(m, x[3 * i + 1]);
            if (3 * i + 2 < N)
                m = fmaxf(m, x[3 * i + 2]);
            y[i] = m;
        }
        if constexpr ((N + 2) / 3 == 1)
            return y[0];
        else
            return max_tree<(N + 2) / 3>(y);
    }

    constexpr size_t scan_smem_bytes(int D, int RQ, int L)
    {
        return (size_t)SCAN_STAGES * SCAN_PT * scan_rowf(D) * sizeof(float)  // tile ring
               + (size_t)scan_ls(L) * SCAN_THREADS * RQ * 8                   // list values + ids
               + SCAN_STAGES * (sizeof(uint64_t) + sizeof(unsigned)) + 64;
    }

    template <int D, int RQ, int L>
    __global__ void __launch_bounds__(SCAN_THREADS) knn_scan_kernel(const ScanParams P)
    {
        static_assert(RQ % 2 == 0, "queries are processed as packed pairs");
        static_assert(L <= 64, "slot tags are 6 bits");
        static_assert(SCAN_PT % (2 * SCAN_GP) == 0, "tile must hold whole pairs of groups");
        constexpr int ROWF = scan_rowf(D);
        constexpr int RW = ROWF / 4;           // float4 per filter row
        constexpr int LS = scan_ls(L);
        constexpr int QT = SCAN_THREADS * RQ;  // queries per CTA
        constexpr uint32_t TILE_BYTES = SCAN_PT * ROWF * sizeof(float);

        extern __shared__ __align__(128) unsigned char smem_raw[];
        float *tiles = reinterpret_cast<float *>(smem_raw);
        float *lvals = tiles + SCAN_STAGES * SCAN_PT * ROWF;              // [QT][LS]
        uint32_t *lids = reinterpret_cast<uint32_t *>(lvals + LS * QT);    // [QT][LS]
        uint64_t *bars = reinterpret_cast<uint64_t *>(lids + LS * QT);
        unsigned *done = reinterpret_cast<unsigned *>(bars + SCAN_STAGES);  // warps finished with a stage

        const int tid = threadIdx.x;
        const unsigned long long TU = P.total_units, U = P.units, G = gridDim.x, g = blockIdx.x;
        const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;
        if (u0 >= u1)
            return;

        if (tid == 0)
        {
            for (int s = 0; s < SCAN_STAGES; ++s)
            {
                mbar_init(&bars[s], 1);
                done[s] = 0;
            }
            mbar_fence_init();
        }
        __syncthreads();
        if (tid == 0)
        {
            for (int s = 0; s < SCAN_STAGES && u0 + s < u1; ++s)
            {
                const unsigned long long pt = (u0 + s) % U;
                mbar_expect_tx(&bars[s], TILE_BYTES);
                tma_bulk_g2s(tiles + s * SCAN_PT * ROWF, P.filt + pt * SCAN_PT * ROWF, TILE_BYTES, &bars[s]);
            }
        }

        float2 qq[RQ / 2][D];
        float thr[RQ];
        unsigned long long cur_tile = ~0ull;

        // query j of this thread is column j*SCAN_THREADS + tid of the query tile
        auto list_of = [&](int j) { return (j * SCAN_THREADS + tid) * LS; };

        auto flush = [&](unsigned long long t) {
            // ordinal of this piece among the pieces that overlap query tile t
            const unsigned long long x = t * U;
            const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
            const unsigned seg = (unsigned)(g - gfirst);
#pragma unroll
            for (int j = 0; j < RQ; ++j)
            {
                const size_t q = (size_t)t * QT + j * SCAN_THREADS + tid;
                P.cand_thr[(size_t)seg * P.qpad + q] = thr[j];
                if (P.cand_ids == nullptr)  // threshold pre-pass: only the bound is wanted
                    continue;
                const uint32_t *li = lids + list_of(j);
                for (int i = 0; i < L; ++i)
                    P.cand_ids[((size_t)seg * L + i) * P.qpad + q] = li[i];
            }
        };

        // put v (scan value of point id) into query j's list; returns the new threshold
        auto insert = [&](int j, float v, uint32_t id) {
            float *lv = lvals + list_of(j);
            const uint32_t slot = __float_as_uint(thr[j]) & TAG_MASK;
            lv[slot] = tag_up(v, slot);
            lids[list_of(j) + slot] = id;
            float x[LS];
#pragma unroll
            for (int w = 0; w < LS / 4; ++w)
            {
                const float4 f = reinterpret_cast<const float4 *>(lv)[w];
                x[4 * w + 0] = f.x;
                x[4 * w + 1] = f.y;
                x[4 * w + 2] = f.z;
                x[4 * w + 3] = f.w;
            }
            thr[j] = max_tree<LS>(x);
        };

        for (unsigned long long u = u0; u < u1; ++u)
        {
            const unsigned long long t = u / U, pt = u % U;
            if (t != cur_tile)
            {
                if (cur_tile != ~0ull)
                    flush(cur_tile);
                cur_tile = t;
#pragma unroll
                for (int jj = 0; jj < RQ / 2; ++jj)
                {
                    const size_t qa = (size_t)t * QT + (2 * jj) * SCAN_THREADS + tid;
                    const size_t qb = qa + SCAN_THREADS;
#pragma unroll
                    for (int c = 0; c < D; ++c)
                        qq[jj][c] = make_float2(P.qp[(size_t)c * P.qpad + qa], P.qp[(size_t)c * P.qpad + qb]);
                }
#pragma unroll
                for (int j = 0; j < RQ; ++j)
                {
                    float *lv = lvals + list_of(j);
                    uint32_t *li = lids + list_of(j);
                    // Starting bound T of this query (threshold pre-pass over a sample of the
                    // points, nn.cu): at least L points have a scan value below T, so values
                    // >= T can never be among the L smallest.  The list starts full of
                    // "virtual" entries worth T (id SLOT_NONE); real points displace them.
                    const float T = P.init_thr ? P.init_thr[(size_t)t * QT + j * SCAN_THREADS + tid] : 3.0e38f;
                    const bool bounded = T < 1.0e38f;
                    float m = __int_as_float(0xff800000);
                    for (int i = 0; i < LS; ++i)
                    {
                        // real slots: empty marker (filled from slot L-1 downwards) or T;
                        // padding slots: never the maximum
                        float v = __int_as_float(0xff800000);
                        if (i < L)
                            v = bounded ? tag_up(T, (uint32_t)i) : __uint_as_float(EMPTY_BITS | (uint32_t)i);
                        lv[i] = v;
                        li[i] = SLOT_NONE;
                        m = fmaxf(m, v);
                    }
                    thr[j] = m;
                }
            }

            const unsigned it = (unsigned)(u - u0);
            const int stage = it % SCAN_STAGES;
            mbar_wait(&bars[stage], (it / SCAN_STAGES) & 1);

            const float4 *tp = reinterpret_cast<const float4 *>(tiles + stage * SCAN_PT * ROWF);
            const uint32_t base = (uint32_t)(pt * SCAN_PT);
            // Points are taken SCAN_GP at a time: the FMA chains of a group are
            // independent (SCAN_GP * RQ/2 packed chains in flight), the rows of
            // the NEXT group are already on their way from shared memory while
            // this one is computed, and one predicate decides whether any
            // (point, query) of the group beats its query's threshold.  Only then
            // (rare after the first few tiles) the group is replayed in point order.
            // Two register buffers of SCAN_GP rows each, used in turn (no copies between them).
            auto load_group = [&](float4 (&buf)[SCAN_GP][RW], int i) {
#pragma unroll
                for (int p = 0; p < SCAN_GP; ++p)
#pragma unroll
                    for (int w = 0; w < RW; ++w)
                        buf[p][w] = tp[(i + p) * RW + w];
            };
            auto coord = [](const float4 &f, int c) { return c == 0 ? f.x : (c == 1 ? f.y : (c == 2 ? f.z : f.w)); };
            auto process = [&](const float4 (&buf)[SCAN_GP][RW], int i) {
                float2 acc[SCAN_GP][RQ / 2];
#pragma unroll
                for (int p = 0; p < SCAN_GP; ++p)
                {
                    const float pn = coord(buf[p][D / 4], D % 4);
#pragma unroll
                    for (int jj = 0; jj < RQ / 2; ++jj)
                        acc[p][jj] = make_float2(pn, pn);
                }
                // (ptxas schedules these freely; source order does not reach the SASS)
#pragma unroll
                for (int c = 0; c < D; ++c)
#pragma unroll
                    for (int jj = 0; jj < RQ / 2; ++jj)
#pragma unroll
                        for (int p = 0; p < SCAN_GP; ++p)
                        {
                            const float x = coord(buf[p][c / 4], c % 4);
                            acc[p][jj] = ffma2(qq[jj][c], make_float2(x, x), acc[p][jj]);
                        }
                // per-query minimum over the group: one compare per query decides the group
                float gm[RQ];
#pragma unroll
                for (int jj = 0; jj < RQ / 2; ++jj)
                {
                    float mx = acc[0][jj].x, my = acc[0][jj].y;
#pragma unroll
                    for (int p = 1; p < SCAN_GP; ++p)
                    {
                        mx = fminf(mx, acc[p][jj].x);
                        my = fminf(my, acc[p][jj].y);
                    }
                    gm[2 * jj] = mx;
                    gm[2 * jj + 1] = my;
                }
                bool hit = false;
#pragma unroll
                for (int j = 0; j < RQ; ++j)
                    hit = hit | (gm[j] < thr[j]);
                if (hit)
                {
                    // slow path: only the queries whose group minimum passed look at the points
#pragma unroll
                    for (int j = 0; j < RQ; ++j)
                        if (gm[j] < thr[j])
                        {
#pragma unroll
                            for (int p = 0; p < SCAN_GP; ++p)
                            {
                                const float v = (j & 1) ? acc[p][j / 2].y : acc[p][j / 2].x;
                                if (v < thr[j])
                                    insert(j, v, base + i + p);
                            }
                        }
                }
            };
            float4 bufa[SCAN_GP][RW], bufb[SCAN_GP][RW];
            load_group(bufa, 0);
#pragma unroll 1
            for (int i = 0; i < SCAN_PT; i += 2 * SCAN_GP)
            {
                load_group(bufb, i + SCAN_GP);
                process(bufa, i);
                // wraps to the first group on the last iteration (harmless re-read)
                load_group(bufa, (i + 2 * SCAN_GP) & (SCAN_PT - 1));
                process(bufb, i + SCAN_GP);
            }
            // Warps run the ring independently (no block barrier): the last warp to
            // finish a stage refills it.  Every lane's reads of the stage are complete
            // (their values were consumed above) before lane 0 counts the warp in.
            __syncwarp();
            if ((tid & 31) == 0)
            {
                const unsigned prev = atomicAdd(&done[stage], 1u);
                if (prev == SCAN_THREADS / 32 - 1)
                {
                    done[stage] = 0;  // the other warps cannot pass the barrier wait before the refill lands
                    if (u + SCAN_STAGES < u1)
                    {
                        const unsigned long long npt = (u + SCAN_STAGES) % U;
                        mbar_expect_tx(&bars[stage], TILE_BYTES);
                        tma_bulk_g2s(tiles + stage * SCAN_PT * ROWF, P.filt + npt * SCAN_PT * ROWF, TILE_BYTES,
                                     &bars[stage]);
                    }
                }
            }
        }
        flush(cur_tile);
    }
}  // namespace omplcu
