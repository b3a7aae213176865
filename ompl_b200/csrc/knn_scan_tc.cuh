// knn_scan_tc.cuh -- interface of the tensor-core filter scan (knn_scan_tc.cu) for k-nearest
// queries on plain RealVector spaces of 64..128 dimensions: the "dense contraction" case
// of the hot path (d >= 64 plain-L2 distance matrix, a filter only, followed
// by the exact FP64 re-rank of nn.cu).
//
// Contract, as knn_scan.cuh / knn_scan_hd.cuh: per (query, point) pair a scan value
//        s ~= pn + sum_c (-2 qt_c) * pt_c ,     pn = fl32(|pt|^2)
// and per (segment, query) the L smallest s with their slots.  Here s comes out of
// tcgen05.mma (BF16 operands, FP32 accumulators in TMEM) on split operands:
//        x = b1(x) + b2(x) + eps,   b1 = bf16(x), b2 = bf16(x - b1),  |eps| <= 2^-17 |x|
//        s_tc = sum_c [ b1(q_c) b1(p_c) + b1(q_c) b2(p_c) + b2(q_c) b1(p_c) ] + n1 + n2 + n3
// with q_c = -2 qt_c and n1 + n2 + n3 = pn exactly (three BF16 pieces hold 24 bits).  That
// is ONE GEMM with K = 3 d + 3: the A row of a query is [b1(q) | b1(q) | b2(q) | 1 1 1],
// the B row of a point is [b1(p) | b2(p) | b1(p) | n1 n2 n3].  The re-rank's band check uses
// the error bound of THIS evaluation (tc_err_units below) instead of the FP32 kernels' (d + 3) u.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace omplcu
{
    constexpr int TC_QT = 128;          // queries per tile  (UMMA M)
    constexpr int TC_PT = 128;          // points per tile   (UMMA N)
    constexpr int TC_KC = 64;           // BF16 elements of K per shared-memory chunk (128 B per row)
    constexpr int TC_CHUNK_ELEMS = TC_PT * TC_KC;        // 8192 BF16 = 16 KB: one operand chunk
    constexpr int TC_MIN_DIM = 64, TC_MAX_DIM = 128;

    __host__ __device__ constexpr int tc_kpad(int D)
    {
        return (3 * D + 3 + TC_KC - 1) / TC_KC * TC_KC;
    }

    // Scan-value error bound in units of u = 2^-24 times (pmax2 + 2 |qt| pmax), see DESIGN.md 3.5:
    //   3      pn as stored (rounding of the FP32 norm, as in the FP32 kernels)
    //   192    split: dropped b2 b2 products and the two eps terms, 3 * 2^-18 = 192 u
    //   32/MMA accumulation: every tcgen05.mma (K = 16) is allowed 2^-19 of the largest partial sum
    //   x 1.02 partial sums of split operands exceed |q||p| by at most 2^-7
    __host__ __device__ constexpr double tc_err_units(int D)
    {
        return 1.02 * (3.0 + 192.0 + 32.0 * (tc_kpad(D) / 16));
    }

    struct TcScanParams
    {
        const uint16_t *pimg;   // [units][nkc][TC_CHUNK_ELEMS] BF16, UMMA core-matrix order
        const uint16_t *qimg;   // [qtiles][nkc][TC_CHUNK_ELEMS]
        float *cand_thr;        // [smax][qpad]
        uint32_t *cand_ids;     // [smax*L][qpad]
        unsigned long long units, total_units;
        uint32_t qpad, smax;
        int nkc, nstages;
        uint32_t lbo, sbo;      // byte strides of the image: K-adjacent core matrices / 8-row groups
        int swap_desc;          // debugging: exchange the two descriptor fields
        float *dbg;             // [qpad][units*TC_PT] raw scan values, or nullptr
    };

    size_t tc_smem_bytes(int nkc, int nstages, int L);
    int tc_pick_stages(int nkc, int L);  // 0: does not fit

    // images
    void tc_build_point_image(const float *filt, int rowf, int D, size_t units, uint32_t lbo, uint32_t sbo,
                              uint16_t *out, cudaStream_t s);
    void tc_build_query_image(const float *qp, uint32_t qpad, int D, uint32_t lbo, uint32_t sbo, uint16_t *out,
                              cudaStream_t s);
    // L in {16, 24, 48}
    void tc_launch_scan(int L, const TcScanParams &P, unsigned grid, cudaStream_t s);
}  // namespace omplcu
