// knn_scan_hd.cuh -- the FP32 filter scan for k-nearest queries on plain RealVector
// spaces of 17..128 dimensions (the d = 100 kinematic-chain L2 case of SURVEY 8(d), M4).
//
// Same contract as knn_scan.cuh: per (query, point) pair
//        s = pn + sum_c (-2 qt_c) * pt_c        (one FP32 accumulator, c ascending)
// and per (segment, query) the L smallest s with their slots, handed to the exact FP64
// re-rank and its band check (nn.cu), which do not care which kernel filled the lists.
//
// At these dimensions a thread cannot keep its queries in registers, so the mapping is
// the register-tiled one of a dense contraction: a CTA of 128 threads owns a 64-query x
// 128-point tile, every thread an 8 x 8 block of accumulators; the query tile (all
// coordinates, [c][q]) stays in shared memory for the CTA's whole segment, point
// coordinates stream through a double-buffered [16][128] chunk.  Per coordinate a thread
// issues 4 LDS.128 and 64 FFMA.  The finished 64 x 128 block of scan values goes through
// shared memory to the selection step: warp w owns queries 16w..16w+15 and their lists, so
// no two warps ever touch the same list.
// FP32-pipe bound (d FMAs per pair); tensor cores would need a split-precision scheme to
// keep the band check's error bound and are left for a later round.
#pragma once

#include "knn_scan.cuh"

namespace omplcu
{
    constexpr int HD_QT = 64;       // queries per CTA tile
    constexpr int HD_PT = 128;      // points per tile
    constexpr int HD_KC = 16;       // coordinates per chunk
    constexpr int HD_THREADS = 128;
    constexpr int HD_SS = HD_PT + 4;  // row stride of the scan-value tile (keeps 16 B alignment)

    __host__ __device__ constexpr int hd_dpad(int D)
    {
        return (D + HD_KC - 1) / HD_KC * HD_KC;
    }

    constexpr size_t hd_smem_bytes(int D, int L)
    {
        return ((size_t)hd_dpad(D) * HD_QT          // query tile [c][q]
                + 2 * HD_KC * HD_PT                 // point chunks
                + HD_PT                             // point norms
                + (size_t)HD_QT * HD_SS             // scan values
                + (size_t)HD_QT * scan_ls(L) * 2    // list values + ids
                + HD_QT)                            // thresholds
               * sizeof(float);
    }

    template <int L>
    __global__ void __launch_bounds__(HD_THREADS) knn_scan_hd_kernel(const ScanParams P, int D, int rowf)
    {
        constexpr int LS = scan_ls(L);
        extern __shared__ __align__(16) float hd_smem[];
        const int DP = hd_dpad(D);
        float *Qs = hd_smem;                               // [DP][HD_QT]
        float *Ps = Qs + (size_t)DP * HD_QT;               // [2][HD_KC][HD_PT]
        float *PN = Ps + 2 * HD_KC * HD_PT;                // [HD_PT]
        float *Ss = PN + HD_PT;                            // [HD_QT][HD_SS]
        float *lvals = Ss + HD_QT * HD_SS;                 // [HD_QT][LS]
        uint32_t *lids = reinterpret_cast<uint32_t *>(lvals + HD_QT * LS);
        float *thr = reinterpret_cast<float *>(lids + HD_QT * LS);  // [HD_QT]

        const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
        const int ty = tid >> 4, tx = tid & 15;  // 8 x 16 threads: queries ty*8.., points tx*8..
        const unsigned long long TU = P.total_units, U = P.units, G = gridDim.x, g = blockIdx.x;
        const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;
        if (u0 >= u1)
            return;
        const int nchunk = DP / HD_KC;
        unsigned long long cur_tile = ~0ull;

        auto flush = [&](unsigned long long t) {
            const unsigned long long x = t * U;
            const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
            const unsigned seg = (unsigned)(g - gfirst);
            if (tid < HD_QT)
            {
                const size_t q = (size_t)t * HD_QT + tid;
                P.cand_thr[(size_t)seg * P.qpad + q] = thr[tid];
                for (int i = 0; i < L; ++i)
                    P.cand_ids[((size_t)seg * L + i) * P.qpad + q] = lids[tid * LS + i];
            }
        };

        // one float4 of row `row`, columns col..col+3, coordinates >= D zeroed
        auto load_p = [&](size_t row, int col) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col < rowf)
                v = __ldg(reinterpret_cast<const float4 *>(P.filt + row * rowf + col));
            v.x = col + 0 < D ? v.x : 0.f;
            v.y = col + 1 < D ? v.y : 0.f;
            v.z = col + 2 < D ? v.z : 0.f;
            v.w = col + 3 < D ? v.w : 0.f;
            return v;
        };

        for (unsigned long long u = u0; u < u1; ++u)
        {
            const unsigned long long t = u / U, pt = u % U;
            if (t != cur_tile)
            {
                __syncthreads();
                if (cur_tile != ~0ull)
                    flush(cur_tile);
                __syncthreads();
                cur_tile = t;
                for (int i = tid; i < DP * HD_QT; i += HD_THREADS)
                {
                    const int c = i / HD_QT, q = i % HD_QT;
                    Qs[i] = c < D ? P.qp[(size_t)c * P.qpad + t * HD_QT + q] : 0.f;
                }
                for (int i = tid; i < HD_QT * LS; i += HD_THREADS)
                {
                    const int s = i % LS;
                    lvals[i] = s < L ? __uint_as_float(EMPTY_BITS | (uint32_t)s) : __int_as_float(0xff800000);
                    lids[i] = SLOT_NONE;
                }
                if (tid < HD_QT)
                    thr[tid] = __uint_as_float(EMPTY_BITS | (uint32_t)(L - 1));
            }
            const size_t row0 = (size_t)pt * HD_PT;
            // chunk 0 and the norms
            float4 pre[4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
                pre[r] = load_p(row0 + tid, r * 4);
            PN[tid] = P.filt[(row0 + tid) * rowf + D];
            __syncthreads();  // Qs / lists / PN visible; previous tile's selection finished with Ss
            float acc[8][8];
            {
                float pn[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    pn[j] = PN[tx * 8 + j];
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        acc[i][j] = pn[j];
            }
            for (int ch = 0; ch < nchunk; ++ch)
            {
                float *pb = Ps + (ch & 1) * HD_KC * HD_PT;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                {
                    pb[(r * 4 + 0) * HD_PT + tid] = pre[r].x;
                    pb[(r * 4 + 1) * HD_PT + tid] = pre[r].y;
                    pb[(r * 4 + 2) * HD_PT + tid] = pre[r].z;
                    pb[(r * 4 + 3) * HD_PT + tid] = pre[r].w;
                }
                __syncthreads();
                if (ch + 1 < nchunk)
                {
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        pre[r] = load_p(row0 + tid, (ch + 1) * HD_KC + r * 4);
                }
                const float *qrow = Qs + (size_t)ch * HD_KC * HD_QT + ty * 8;
                const float *prow = pb + tx * 8;
#pragma unroll
                for (int kk = 0; kk < HD_KC; ++kk)
                {
                    const float4 qa = *reinterpret_cast<const float4 *>(qrow + kk * HD_QT);
                    const float4 qb = *reinterpret_cast<const float4 *>(qrow + kk * HD_QT + 4);
                    const float4 pa = *reinterpret_cast<const float4 *>(prow + kk * HD_PT);
                    const float4 pc = *reinterpret_cast<const float4 *>(prow + kk * HD_PT + 4);
                    const float qv[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
                    const float pv[8] = {pa.x, pa.y, pa.z, pa.w, pc.x, pc.y, pc.z, pc.w};
#pragma unroll
                    for (int i = 0; i < 8; ++i)
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            acc[i][j] = fmaf(qv[i], pv[j], acc[i][j]);
                }
                // the other buffer is rewritten next iteration only after everyone passed this sync
            }
            // scan values to shared memory, then selection by the warp that owns the query
#pragma unroll
            for (int i = 0; i < 8; ++i)
            {
                float *d = Ss + (ty * 8 + i) * HD_SS + tx * 8;
                *reinterpret_cast<float4 *>(d) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                *reinterpret_cast<float4 *>(d + 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
            }
            __syncthreads();
            const uint32_t base = (uint32_t)row0;
            for (int qq = 0; qq < HD_QT / 4; ++qq)
            {
                const int q = w * (HD_QT / 4) + qq;
                const float4 v = *reinterpret_cast<const float4 *>(Ss + q * HD_SS + lane * 4);
                const float tq = thr[q];
                unsigned mask = __ballot_sync(0xffffffffu, (v.x < tq) | (v.y < tq) | (v.z < tq) | (v.w < tq));
                while (mask)
                {
                    const int l = __ffs(mask) - 1;
                    mask &= mask - 1;
                    if (lane == l)
                    {
                        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                        {
                            const float cur = thr[q];
                            if (vv[e] < cur)
                            {
                                const uint32_t slot = __float_as_uint(cur) & TAG_MASK;
                                lvals[q * LS + slot] = tag_up(vv[e], slot);
                                lids[q * LS + slot] = base + lane * 4 + e;
                                float m = lvals[q * LS];
                                for (int s = 1; s < LS; ++s)
                                    m = fmaxf(m, lvals[q * LS + s]);
                                thr[q] = m;
                            }
                        }
                    }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
        flush(cur_tile);
    }
}  // namespace omplcu
