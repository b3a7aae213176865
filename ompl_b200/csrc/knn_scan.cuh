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
// unsorted list in shared memory with the running maximum cached in a
// register as the insertion threshold).  The re-rank recomputes the exact
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
    constexpr uint32_t SLOT_NONE = 0xFFFFFFFFu;

    __host__ __device__ constexpr int scan_rowf(int D)
    {
        return ((D + 1 + 3) / 4) * 4;  // floats per filter row: D coords + norm, padded to 16 B
    }

    struct ScanParams
    {
        const float *filt;      // [units*SCAN_PT][ROWF]
        const float *qp;        // [D][qpad]  (-2 * qt_c)
        float *cand_vals;       // [smax*L][qpad]
        uint32_t *cand_ids;     // [smax*L][qpad]
        unsigned long long units;        // point tiles
        unsigned long long total_units;  // qtiles * units
        uint32_t qpad;
        uint32_t smax;
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

    constexpr size_t scan_smem_bytes(int D, int RQ, int L)
    {
        return (size_t)SCAN_STAGES * SCAN_PT * scan_rowf(D) * sizeof(float)  // tile ring
               + (size_t)L * SCAN_THREADS * RQ * 8                            // list values + ids
               + SCAN_STAGES * sizeof(uint64_t) + 64;
    }

    template <int D, int RQ, int L>
    __global__ void __launch_bounds__(SCAN_THREADS) knn_scan_kernel(const ScanParams P)
    {
        static_assert(RQ % 2 == 0, "queries are processed as packed pairs");
        constexpr int ROWF = scan_rowf(D);
        constexpr int QT = SCAN_THREADS * RQ;  // queries per CTA
        constexpr uint32_t TILE_BYTES = SCAN_PT * ROWF * sizeof(float);

        extern __shared__ __align__(128) unsigned char smem_raw[];
        float *tiles = reinterpret_cast<float *>(smem_raw);
        float *lvals = tiles + SCAN_STAGES * SCAN_PT * ROWF;
        uint32_t *lids = reinterpret_cast<uint32_t *>(lvals + L * QT);
        uint64_t *bars = reinterpret_cast<uint64_t *>(lids + L * QT);

        const int tid = threadIdx.x;
        const unsigned long long TU = P.total_units, U = P.units, G = gridDim.x, g = blockIdx.x;
        const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;
        if (u0 >= u1)
            return;

        if (tid == 0)
        {
            for (int s = 0; s < SCAN_STAGES; ++s)
                mbar_init(&bars[s], 1);
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
        int maxpos[RQ];
        unsigned long long cur_tile = ~0ull;

        auto flush = [&](unsigned long long t) {
            // ordinal of this piece among the pieces that overlap query tile t
            const unsigned long long x = t * U;
            const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
            const unsigned seg = (unsigned)(g - gfirst);
#pragma unroll
            for (int j = 0; j < RQ; ++j)
            {
                const int col = j * SCAN_THREADS + tid;
                const size_t q = (size_t)t * QT + col;
                for (int i = 0; i < L; ++i)
                {
                    const size_t o = ((size_t)seg * L + i) * P.qpad + q;
                    P.cand_vals[o] = lvals[i * QT + col];
                    P.cand_ids[o] = lids[i * QT + col];
                }
            }
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
                    thr[j] = __int_as_float(0x7f800000);
                    maxpos[j] = 0;
                    const int col = j * SCAN_THREADS + tid;
                    for (int i = 0; i < L; ++i)
                    {
                        lvals[i * QT + col] = __int_as_float(0x7f800000);
                        lids[i * QT + col] = SLOT_NONE;
                    }
                }
            }

            const unsigned it = (unsigned)(u - u0);
            const int stage = it % SCAN_STAGES;
            mbar_wait(&bars[stage], (it / SCAN_STAGES) & 1);

            const float4 *tp = reinterpret_cast<const float4 *>(tiles + stage * SCAN_PT * ROWF);
            const uint32_t base = (uint32_t)(pt * SCAN_PT);
#pragma unroll 4
            for (int i = 0; i < SCAN_PT; ++i)
            {
                float pv[ROWF];
#pragma unroll
                for (int w = 0; w < ROWF / 4; ++w)
                {
                    const float4 v = tp[i * (ROWF / 4) + w];
                    pv[4 * w + 0] = v.x;
                    pv[4 * w + 1] = v.y;
                    pv[4 * w + 2] = v.z;
                    pv[4 * w + 3] = v.w;
                }
                float2 acc[RQ / 2];
#pragma unroll
                for (int jj = 0; jj < RQ / 2; ++jj)
                    acc[jj] = make_float2(pv[D], pv[D]);
#pragma unroll
                for (int c = 0; c < D; ++c)
#pragma unroll
                    for (int jj = 0; jj < RQ / 2; ++jj)
                        acc[jj] = ffma2(qq[jj][c], make_float2(pv[c], pv[c]), acc[jj]);
                bool hit = false;
#pragma unroll
                for (int jj = 0; jj < RQ / 2; ++jj)
                    hit = hit | (acc[jj].x < thr[2 * jj]) | (acc[jj].y < thr[2 * jj + 1]);
                if (hit)
                {
#pragma unroll
                    for (int j = 0; j < RQ; ++j)
                    {
                        const float v = (j & 1) ? acc[j / 2].y : acc[j / 2].x;
                        if (v < thr[j])
                        {
                            const int col = j * SCAN_THREADS + tid;
                            lvals[maxpos[j] * QT + col] = v;
                            lids[maxpos[j] * QT + col] = base + i;
                            float m = lvals[col];
                            int mp = 0;
#pragma unroll
                            for (int l = 1; l < L; ++l)
                            {
                                const float x = lvals[l * QT + col];
                                if (x > m)
                                {
                                    m = x;
                                    mp = l;
                                }
                            }
                            thr[j] = m;
                            maxpos[j] = mp;
                        }
                    }
                }
            }
            __syncthreads();  // every warp is done with this stage before it is refilled
            if (tid == 0 && u + SCAN_STAGES < u1)
            {
                const unsigned long long npt = (u + SCAN_STAGES) % U;
                mbar_expect_tx(&bars[stage], TILE_BYTES);
                tma_bulk_g2s(tiles + stage * SCAN_PT * ROWF, P.filt + npt * SCAN_PT * ROWF, TILE_BYTES,
                             &bars[stage]);
            }
        }
        flush(cur_tile);
    }
}  // namespace omplcu
