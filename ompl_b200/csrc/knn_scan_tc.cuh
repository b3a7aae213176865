// knn_scan_tc.cuh -- interface of the tensor-core filter scan (knn_scan_tc.cu) for batched
// k-nearest queries on plain RealVector spaces of 1..128 dimensions: the distance matrix as a
// dense contraction (SURVEY 8(d) asks for it at d >= 64; measured faster at every dimension once
// queries come in batches), a filter only, followed by the exact FP64 re-rank of nn.cu.
//
// Contract, as knn_scan.cuh / knn_scan_hd.cuh: per (query, point) pair a scan value
//        s ~= pn + sum_c (-2 qt_c) * pt_c ,     pn = fl32(|pt|^2)
// and per (segment, query) the L smallest s with their slots.  Here s comes out of
// tcgen05.mma (BF16 operands, FP32 accumulators in TMEM) on split operands:
//        x = b1(x) + b2(x) + eps,   b1 = bf16(x), b2 = bf16(x - b1),  |eps| <= 2^-16 |x|
//        s_tc = sum_c [ b1(q_c) b1(p_c) + b1(q_c) b2(p_c) + b2(q_c) b1(p_c) ] + n1 + n2 + n3
// with q_c = -2 qt_c and n1 + n2 + n3 = pn exactly (three BF16 pieces hold 24 bits).  That
// is ONE GEMM with K = 3 d + 3: the A row of a query is [b1(q) | b1(q) | b2(q) | 1 1 1],
// the B row of a point is [b1(p) | b2(p) | b1(p) | n1 n2 n3].  The re-rank's band check uses
// the error bound of THIS evaluation (tc_err_units below) instead of the FP32 kernels' (d + 3) u.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>

namespace omplcu
{
    constexpr int TC_QT = 128;          // queries per A tile (UMMA M); a CTA keeps 1, 2 or 4 tiles resident
    constexpr int TC_PT = 128;          // points per tile   (UMMA N)
    constexpr int TC_MAX_DIM = 128;
    constexpr int TC_MIN_BATCH = 1;     // "auto": measured faster than the FP32 scans down to a single query (0.51 vs 0.97 ms at 1M x 7-D)

    // K = 3 d + 3 is covered by k16 = ceil(K / 16) tcgen05.mma instructions (K = 16 each).  The operand images are
    // cut into nkc chunks of kw 16-byte column groups (kw even: two groups per MMA; 2 KB per group and 128 rows), a
    // chunk being the unit of the TMA ring.  Every chunk costs a ring hand-shake and a tcgen05.commit on top of its
    // MMAs (measured ~200 clk, profiles/r2_experiments.md), so chunks hold up to tc_chunk_mmas() MMAs; columns
    // beyond 16 k16 are zero padding that is copied but never multiplied.
    struct TcShape
    {
        int kw, nkc, na;  // column groups per chunk, chunks, resident query tiles
        int k16;          // MMAs per (query tile, point tile) pair
        __host__ __device__ int kpad() const
        {
            return nkc * kw * 8;
        }
        __host__ __device__ size_t chunk_elems() const
        {
            return (size_t)kw * 8 * TC_PT;
        }
    };
    // The shape is a function of the dimension alone (the point image is built once and shared by all queries): fewest
    // padded bytes plus a per-chunk charge, under "4 ring stages fit beside a k <= 10 query's lists" (knn_scan_tc.cu).
    // OMPLCU_TC_CHUNK_MMAS caps the MMAs per chunk (tuning; default 10).
    TcShape tc_shape(int D);

    // Scan-value error bound in units of u = 2^-24 times (pmax2 + 2 |qt| pmax), see DESIGN.md 3.5:
    //   3      pn as stored (rounding of the FP32 norm, as in the FP32 kernels)
    //   768    split: BF16 keeps 8 significant bits, so x = b1 + b2 + eps with |b2| <= 2^-8 |x| and
    //          |eps| <= 2^-16 |x|; the dropped b2 b2 product and the two eps terms are each
    //          <= 2^-16 |q_c||p_c|, summed over c <= 3 * 2^-16 |q||p| = 768 u (Cauchy-Schwarz)
    //   32/MMA accumulation: every tcgen05.mma (K = 16) is allowed 2^-19 of the largest partial sum
    //          (products of BF16 values are exact in FP32; the accumulator rounding is not documented,
    //          the allowance is 16 ulp per instruction and the tests measure the real figure)
    //   x 1.02 second-order terms of the above
    inline double tc_err_units(int D)
    {
        return 1.02 * (3.0 + 768.0 + 32.0 * ((3 * D + 3 + 15) / 16));
    }

    struct TcScanParams
    {
        const uint16_t *pimg;   // [units][nkc][chunk] BF16, UMMA core-matrix order
        const uint16_t *qimg;   // [qtiles][nkc][chunk]; a CTA loads na consecutive tiles at once
        float *cand_thr;        // [smax][qpad]
        uint32_t *cand_ids;     // [smax*L][qpad]
        unsigned long long units, total_units;  // point tiles; (query super-tiles) x units
        uint32_t qpad, smax;
        TcShape sh;
        int nstages;
        int diag;               // diagnostics: 1 swap descriptor fields, 2 no selection, 4 no insertion
        float *dbg;             // [qpad][units*TC_PT] raw scan values, or nullptr
        int ts;                 // list kernel, one resident tile: query tile as the TMEM-resident A operand
        int pair;               // ts only: CTA pairs share every point chunk (grid = 2 x work pieces, super-tiles of 2 query tiles)
        // list kernel only: the list values themselves, [smax*L][qpad] (tagged, i.e. rounded up by < 128 ulp;
        // EMPTY markers > 1e38 where the list did not fill), or nullptr.  Read by the threshold selection.
        float *list_vals;
        // threshold kernel only (tc_launch_thr_scan): per-query bound and the append rows
        const float *tau;       // [qpad]  a point is kept iff its scan value < tau[query]
        uint32_t *app_ids;      // [smax][qpad][cap] slots kept, in scan order
        uint32_t *app_cnt;      // [smax][qpad] how many (> cap: the row overflowed, its content is incomplete)
        uint32_t cap;           // multiple of 32, >= 64
    };

    size_t tc_smem_bytes(const TcShape &sh, int nstages, int L, bool ts = false);
    int tc_pick_stages(const TcShape &sh, int L, bool ts = false);  // 0: does not fit
    bool tc_ts_usable(const TcShape &sh);  // query tile as TMEM-resident A operand possible for this shape

    // images
    void tc_build_point_image(const float *filt, int rowf, int D, size_t units, uint16_t *out, cudaStream_t s);
    void tc_build_query_image(const float *qp, uint32_t qpad, int D, uint16_t *out, cudaStream_t s);
    // every row_step-th filter row (rows >= src_rows count as dead): the sample the thresholds come from
    void tc_build_sample_image(const float *filt, int rowf, int D, size_t src_rows, size_t row_step, size_t units,
                               uint16_t *out, cudaStream_t s);
    // list scan: the L smallest scan values per (segment, query); L in {8, 16, 24, 48}
    void tc_launch_scan(int L, const TcScanParams &P, unsigned grid, cudaStream_t s);
    // threshold scan: every slot whose scan value is below the query's bound is appended to the
    // query's row of its segment; no lists, no shared-memory state per query
    void tc_launch_thr_scan(const TcScanParams &P, unsigned grid, cudaStream_t s);
    size_t tc_thr_smem_bytes(const TcShape &sh, int nstages);
    int tc_thr_pick_stages(const TcShape &sh);
}  // namespace omplcu
