// knn_scan_tc.cu -- tensor-core filter scan (tcgen05.mma, BF16 split operands, FP32
// accumulators in TMEM) for batched k-nearest queries on plain RealVector spaces of 1..128
// dimensions.  See knn_scan_tc.cuh for the contract and DESIGN.md 3.5 for the error bound,
// the measurements and what bounds each shape.
//
// One persistent CTA per SM, warp-specialised; NA = 4 / 2 / 1 resident query tiles of 128 queries:
//   warp 0, one lane      producer: cp.async.bulk (TMA bulk copy) of the query image (once per
//                         query super-tile, stays resident) and of 8 / 16 KB point chunks into a ring
//   warps 1..NA, one lane tcgen05.mma issuers, one per resident tile: per point tile Kpad/16 MMAs
//                         (M 128 x N 128 x K 16; N 64 x 2 at NA = 4) into a TMEM accumulator;
//                         tcgen05.commit frees the ring slot and hands the accumulator over
//   4 NA selection warps  thread = TMEM lane = query; tcgen05.ld 32 columns at a time, one FMNMX3
//                         tree against the query's threshold (a register), the rare hit goes
//                         into the query's private list in shared memory (out of line)
// Operand images are written by the prep kernels in exactly the order the UMMA shared-memory
// descriptor walks them (K-major, no swizzle: 8-row x 16-byte core matrices), so every stage is
// ONE contiguous bulk copy and no tensor map is needed.
#include "knn_scan_tc.cuh"

#include <cuda_bf16.h>

#include "common.cuh"
#include "knn_scan.cuh"

namespace omplcu
{
    namespace
    {
        // threads per CTA: TMA warp, then per resident query tile one MMA warp and four selection warps
        __host__ __device__ constexpr int tc_threads(int NA)
        {
            return 32 * (1 + NA + 4 * NA);
        }
        constexpr int TC_ACC = 4;  // accumulator buffers (4 x 128 = all 512 TMEM columns)
        constexpr int TC_MAX_STAGES = 12;

        // ---------------- operand images ----------------
        // Within a 128-row block the image is [chunk][column group j][row group g][row r][8 BF16]:
        // an 8-row x 16-byte core matrix is 128 contiguous bytes, row groups follow each other
        // (stride byte offset 128), K-adjacent core matrices are 2048 bytes apart (leading byte offset).
        constexpr uint32_t TC_LBO = 2048, TC_SBO = 128;

        __host__ __device__ inline size_t tc_img_off(int row, int k, int kw)
        {
            const int cw = kw * 8, kc = k / cw, j = (k % cw) >> 3, e = k & 7;
            return (size_t)kc * cw * TC_PT + (size_t)j * (TC_LBO / 2) + (size_t)(row >> 3) * (TC_SBO / 2) +
                   (size_t)(row & 7) * 8 + e;
        }

        __device__ __forceinline__ uint16_t bf16_bits(float x)
        {
            return __bfloat16_as_ushort(__float2bfloat16_rn(x));
        }
        __device__ __forceinline__ float bf16_val(float x)
        {
            return __bfloat162float(__float2bfloat16_rn(x));
        }
        // x = p[0] + p[1] + p[2] exactly (three BF16 pieces hold the 24 bits of an FP32 value)
        __device__ __forceinline__ void bf16_split3(float x, float p[3])
        {
            if (!(fabsf(x) < 3.0e38f))
            {
                p[0] = x;  // +inf marks a dead row (or NaN: poisons the row, the exact path decides)
                p[1] = p[2] = 0.f;
                return;
            }
            p[0] = bf16_val(x);
            const float r1 = x - p[0];
            p[1] = bf16_val(r1);
            p[2] = bf16_val(r1 - p[1]);
        }

        // role of column k for a point row (B operand)
        __device__ __forceinline__ float tc_point_elem(const float *row, int D, int k)
        {
            if (k < 3 * D)
            {
                const int blk = k / D, c = k - blk * D;
                float p[3];
                bf16_split3(row[c], p);
                return blk == 1 ? p[1] : p[0];  // [b1 | b2 | b1]
            }
            if (k < 3 * D + 3)
            {
                float p[3];
                bf16_split3(row[D], p);  // the stored squared norm
                return p[k - 3 * D];
            }
            return 0.f;
        }

        // one thread writes 8 consecutive K elements (16 bytes) of one row; consecutive threads take
        // consecutive rows of the same column group, so a warp writes 512 contiguous bytes
        // Image row i is filter row i * row_step (row_step = 1: the whole set; > 1: the strided sample);
        // rows at or beyond src_rows are dead (norm +inf).
        __global__ void tc_point_image_kernel(const float *filt, int rowf, int D, TcShape sh, size_t rows,
                                              size_t src_rows, size_t row_step, uint16_t *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            const int groups = sh.kpad() / 8;
            if (i >= rows * groups)
                return;
            const size_t tile = i / ((size_t)TC_PT * groups);
            const size_t rem = i % ((size_t)TC_PT * groups);
            const int grp = (int)(rem / TC_PT), r = (int)(rem % TC_PT);
            const size_t src = (tile * TC_PT + r) * row_step;
            const float *row = filt + src * rowf;
            uint16_t v[8];
            if (src >= src_rows)
            {
                // dead row: coordinates 0, n1 = +inf
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    v[e] = (grp * 8 + e == 3 * D) ? (uint16_t)0x7f80u : (uint16_t)0;
            }
            else
#pragma unroll
            for (int e = 0; e < 8; ++e)
                v[e] = bf16_bits(tc_point_elem(row, D, grp * 8 + e));
            uint16_t *dst = out + tile * (size_t)sh.nkc * sh.chunk_elems() + tc_img_off(r, grp * 8, sh.kw);
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(v);
        }

        // qp is [c][qpad] = -2 * qt_c (query_prep_kernel in nn.cu); padded queries are zero rows
        __global__ void tc_query_image_kernel(const float *qp, uint32_t qpad, int D, TcShape sh, uint16_t *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            const int groups = sh.kpad() / 8;
            if (i >= (size_t)qpad * groups)
                return;
            const size_t tile = i / ((size_t)TC_QT * groups);
            const size_t rem = i % ((size_t)TC_QT * groups);
            const int grp = (int)(rem / TC_QT), r = (int)(rem % TC_QT);
            const size_t q = tile * TC_QT + r;
            uint16_t v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
            {
                const int k = grp * 8 + e;
                float x = 0.f;
                if (k < 3 * D)
                {
                    const int blk = k / D, c = k - blk * D;
                    const float qv = qp[(size_t)c * qpad + q];
                    const float b1 = bf16_val(qv);
                    x = blk == 2 ? bf16_val(qv - b1) : b1;  // [b1 | b1 | b2]
                }
                else if (k < 3 * D + 3)
                    x = 1.f;
                v[e] = bf16_bits(x);
            }
            uint16_t *dst = out + tile * (size_t)sh.nkc * sh.chunk_elems() + tc_img_off(r, grp * 8, sh.kw);
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(v);
        }

        // ---------------- tcgen05 / TMEM helpers (PTX) ----------------
        __device__ __forceinline__ void mbar_arrive(uint64_t *bar)
        {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
        }
        __device__ __forceinline__ void tc_fence_before()
        {
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        }
        __device__ __forceinline__ void tc_fence_after()
        {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        // whole warp; the TMEM base address lands in *slot (shared memory)
        __device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols)
        {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)),
                         "r"(cols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        __device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols)
        {
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
        }
        // arrive on `bar` once every tcgen05.mma issued so far by this thread has completed
        __device__ __forceinline__ void umma_commit(uint64_t *bar)
        {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                             smem_u32(bar))
                         : "memory");
        }
        // ---- CTA pairs (cluster of 2): both CTAs scan the same point tiles for two different query tiles; each
        // fetches half of every chunk and the copy is multicast into both CTAs' rings, so the L2 serves every
        // chunk once per pair
        __device__ __forceinline__ uint32_t cluster_ctarank()
        {
            uint32_t r;
            asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
            return r;
        }
        __device__ __forceinline__ void cluster_sync_all()
        {
            asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
        __device__ __forceinline__ void mbar_fence_init_cluster()
        {
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        // the bytes land at the same offset of every CTA in `mask`, and complete_tx is signalled on the barrier at
        // the same offset of each of them
        __device__ __forceinline__ void tma_bulk_g2s_mc(void *dst, const void *src, uint32_t bytes, uint64_t *bar,
                                                        uint16_t mask)
        {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], "
                         "%2, [%3], %4;" ::"r"(smem_u32(dst)),
                         "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
                         : "memory");
        }
        // umma_commit arriving on the barrier at this offset in every CTA of `mask`
        __device__ __forceinline__ void umma_commit_mc(uint64_t *bar, uint16_t mask)
        {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                             smem_u32(bar)),
                         "h"(mask)
                         : "memory");
        }
        // D[tmem] (+)= A[smem] * B[smem]^T, BF16 x BF16 -> FP32, issued by one thread
        __device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                  uint32_t accumulate)
        {
            asm volatile("{\n"
                         ".reg .pred p;\n"
                         "setp.ne.b32 p, %4, 0;\n"
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
                         "}\n" ::"r"(tmem_d),
                         "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
                         : "memory");
        }
        // The same with the A operand read from tensor memory (lane = row, every 32-bit column holds two
        // consecutive K elements, low half first): D[tmem] (+)= A[tmem] * B[smem]^T
        __device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                                     uint32_t accumulate)
        {
            asm volatile("{\n"
                         ".reg .pred p;\n"
                         "setp.ne.b32 p, %4, 0;\n"
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
                         "}\n" ::"r"(tmem_d),
                         "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
                         : "memory");
        }
        // four consecutive columns of this thread's TMEM lane <- four registers
        __device__ __forceinline__ void tmem_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
        {
            asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(a), "r"(b),
                         "r"(c), "r"(d)
                         : "memory");
        }
        __device__ __forceinline__ void tmem_st_wait()
        {
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
        // mbarrier wait / arrive on a precomputed shared-memory address.  The wait carries a suspend-time hint:
        // the thread sleeps in hardware until the phase completes (or the hint expires) instead of re-issuing
        // try_wait -- in round 1 a quarter of the kernel's executed instructions were polls.
        __device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity)
        {
            asm volatile("{\n"
                         ".reg .pred p;\n"
                         "WAITA_LOOP:\n"
                         "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
                         "@p bra WAITA_DONE;\n"
                         "bra WAITA_LOOP;\n"
                         "WAITA_DONE:\n"
                         "}\n" ::"r"(bar),
                         "r"(parity), "r"(1000000u)
                         : "memory");
        }
        __device__ __forceinline__ void mbar_arrive_addr(uint32_t bar)
        {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
        }
        // one lane of a converged warp
        __device__ __forceinline__ bool elect_one()
        {
            uint32_t pred;
            asm volatile("{\n"
                         ".reg .pred p;\n"
                         "elect.sync _|p, 0xffffffff;\n"
                         "selp.u32 %0, 1, 0, p;\n"
                         "}\n"
                         : "=r"(pred));
            return pred != 0;
        }
        // shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor):
        //   [0,14) start >> 4, [16,30) leading byte offset >> 4, [32,46) stride byte offset >> 4,
        //   [46,48) version = 1 on sm_100, [61,64) layout type = 0 (no swizzle)
        __device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
        {
            uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
            d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
            d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
            d |= 1ull << 46;
            return d;
        }
        // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = BF16, both K-major
        __host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N)
        {
            return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        }
        // 32 consecutive columns of this thread's TMEM lane.  The load is asynchronous: the registers
        // are valid after tmem_ld_wait, which takes them as operands so that no use can be scheduled
        // ahead of it.
        __device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&r)[32])
        {
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
                  "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
                  "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
                  "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                : "r"(taddr)
                : "memory");
        }
        __device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[32])
        {
            asm volatile("tcgen05.wait::ld.sync.aligned;"
                         : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]),
                           "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]),
                           "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]),
                           "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]),
                           "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                         :
                         : "memory");
        }

        // ---------------- the scan ----------------
        struct TcSmem
        {
            uint64_t full[TC_MAX_STAGES], empty[TC_MAX_STAGES];  // point-chunk ring: TMA <-> MMA
            uint64_t a_full, a_empty;                            // query image:     TMA <-> MMA
            uint64_t t_full[2 * TC_ACC], t_empty[2 * TC_ACC];    // accumulators:    MMA <-> selection (8 of 64 columns at NA = 4)
            uint32_t tmem_base;
        };

        __device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16])
        {
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                         "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
                           "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
                           "=r"(r[14]), "=r"(r[15])
                         : "r"(taddr)
                         : "memory");
        }
        __device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[16])
        {
            asm volatile("tcgen05.wait::ld.sync.aligned;"
                         : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]),
                           "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]),
                           "+r"(r[14]), "+r"(r[15])
                         :
                         : "memory");
        }
        __device__ __forceinline__ void tmem_ld_issue(uint32_t taddr, uint32_t (&r)[32])
        {
            tmem_ld32_issue(taddr, r);
        }
        __device__ __forceinline__ void tmem_ld_issue(uint32_t taddr, uint32_t (&r)[16])
        {
            tmem_ld16_issue(taddr, r);
        }

        // diagnostics (DBG instantiation only): raw scan values of one W-column step
        template <int W>
        __device__ __noinline__ void tc_dump(const TcScanParams &P, const uint32_t (&v)[W], size_t q, uint32_t col0)
        {
            float *d = P.dbg + q * (P.units * TC_PT) + col0;
#pragma unroll
            for (int i = 0; i < W; ++i)
                d[i] = __uint_as_float(v[i]);
        }

        // The slow path of the selection, out of line so that the hot loop stays small in the
        // instruction cache: visit the values of this thread's staged W-column step that beat the
        // threshold, in column order, and put each into the list (it replaces the current maximum).
        // scr_q: the thread's column of the warp's [W][32] staging block.
        template <int L, int W>
        __device__ __noinline__ float tc_select_hits(float thr, uint32_t mask, const float *scr_q, float *lvq,
                                                     uint32_t *liq, uint32_t id0)
        {
            while (mask)
            {
                const int i = __ffs(mask) - 1;
                mask &= mask - 1;
                const float x = scr_q[i * 32];
                if (x < thr)
                {
                    const uint32_t slot = __float_as_uint(thr) & TAG_MASK;
                    lvq[slot] = tag_up(x, slot);
                    liq[slot * TC_QT] = id0 + i;
                    float4 a = *reinterpret_cast<const float4 *>(lvq);
                    float m0 = fmaxf(a.x, a.y), m1 = fmaxf(a.z, a.w);
#pragma unroll
                    for (int k = 1; k < L / 4; ++k)
                    {
                        a = *reinterpret_cast<const float4 *>(lvq + 4 * k);
                        m0 = fmaxf(m0, fmaxf(a.x, a.y));
                        m1 = fmaxf(m1, fmaxf(a.z, a.w));
                    }
                    thr = fmaxf(m0, m1);
                }
            }
            return thr;
        }

        // The common case late in a segment: exactly one column of the step beats the threshold, and its value
        // is the step's minimum, which the caller already holds.  No staging, one insertion.  Inlined: the call
        // and its register shuffling cost more than the 40 instructions' worth of code (d = 7: 5.92 -> 5.88 ms,
        // d = 32: 1.61 -> 1.49 ms).
        template <int L>
        __device__ __forceinline__ float tc_insert_one(float thr, float x, float *lvq, uint32_t *liq, uint32_t id)
        {
            const uint32_t slot = __float_as_uint(thr) & TAG_MASK;
            lvq[slot] = tag_up(x, slot);
            liq[slot * TC_QT] = id;
            float4 a = *reinterpret_cast<const float4 *>(lvq);
            float m0 = fmaxf(a.x, a.y), m1 = fmaxf(a.z, a.w);
#pragma unroll
            for (int k = 1; k < L / 4; ++k)
            {
                a = *reinterpret_cast<const float4 *>(lvq + 4 * k);
                m0 = fmaxf(m0, fmaxf(a.x, a.y));
                m1 = fmaxf(m1, fmaxf(a.z, a.w));
            }
            return fmaxf(m0, m1);
        }

        // Work: the (query super-tile x point tile) rectangle, linearised (super-tile major) and cut
        // into gridDim.x equal pieces as in knn_scan.cuh.  A super-tile is NA query tiles of 128; per
        // point tile the CTA runs NA "jobs" (one accumulator each, job j -> TMEM buffer j % 4).
        // TS (NA == 1 only, the d >= 53 shapes): the query tile is the TMEM-resident A operand.  An SS-mode
        // tcgen05.mma of M 128 x N 128 x K 16 reads 4 KB of A and 4 KB of B from shared memory in its 64 clk,
        // the SM's whole shared-memory bandwidth, and at K = 320 that is what bounds the kernel (ncu round 1:
        // tensor pipe 39 %, tc_wavefronts_mem_shared 38 %).  With A in tensor memory (kpad / 2 columns, written by
        // the selection warps with tcgen05.st at the start of a query super-tile: thread = lane = query) an MMA
        // reads only B.  Columns [0, 256): A; [256, 512): two 128-column accumulators (job j -> buffer j % 2).
        // PAIR (TS only): CTAs 2c and 2c + 1 are a cluster and together own the work piece of "CTA" c of a grid
        // half the size, on super-tiles of TWO query tiles: rank r takes tile 2 t + r.  Both walk the same point
        // tiles in step; rank r fetches half r of every chunk and multicasts it into both rings, a stage is free
        // once both CTAs' MMAs have read it (multicast commit, empty barriers count 2).  The L2 -> SM feed is what
        // bounds the d >= 53 shapes (feed alone 2.75 of 3.3 ms at d = 100, at the chip's L2 output rate).
        template <int L, int NA, bool DBG, bool TS = false, bool PAIR = false>
        __global__ void __launch_bounds__(tc_threads(NA), 1) knn_scan_tc_kernel(const TcScanParams P)
        {
            static_assert(!TS || NA == 1, "the TMEM-resident query tile is built for one resident tile");
            static_assert(!PAIR || TS, "CTA pairs are built on the TMEM-resident query tile");
            constexpr int TPS = PAIR ? 2 : NA;  // query tiles per super-tile
            const uint32_t crank = PAIR ? cluster_ctarank() : 0u;
            constexpr uint32_t NBUF = TS ? 2u : (uint32_t)TC_ACC;      // accumulator buffers of the general path
            constexpr uint32_t ACC0 = TS ? 256u : 0u;                  // first accumulator column
            extern __shared__ __align__(1024) uint8_t tc_smem[];
            constexpr int LS = L + 4;
            const int nkc = P.sh.nkc, nst = P.nstages, kw = P.sh.kw, k16 = P.sh.k16;
            const uint32_t chunk_bytes = (uint32_t)kw * 2048u;
            const uint32_t a_tile_bytes = (uint32_t)nkc * chunk_bytes;
            uint8_t *sA = tc_smem;                                    // [NA][nkc][chunk] (TS: not there, the tile lives in TMEM)
            uint8_t *sB = sA + (TS ? 0 : (size_t)NA * a_tile_bytes);  // [nst][chunk]
            float *lv = reinterpret_cast<float *>(sB + (size_t)nst * chunk_bytes);  // [NA][128][LS]
            uint32_t *li = reinterpret_cast<uint32_t *>(lv + NA * TC_QT * LS);      // [NA][L][128]
            float *scr = reinterpret_cast<float *>(li + NA * L * TC_QT);            // [4 NA warps][W][32] slow-path staging
            constexpr int W = 32;  // columns per selection step
            TcSmem *S = reinterpret_cast<TcSmem *>(scr + 4 * NA * W * 32);

            // the warp index through a shuffle: the compiler then knows that it is warp-uniform and keeps the
            // issuer's descriptors and loop counters in uniform registers
            const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
            const unsigned long long TU = P.total_units, U = P.units, G = PAIR ? gridDim.x / 2 : gridDim.x,
                                     g = PAIR ? blockIdx.x / 2 : blockIdx.x;
            const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;

            if (tid == 0)
            {
                for (int i = 0; i < TC_MAX_STAGES; ++i)
                {
                    mbar_init(&S->full[i], 1);
                    mbar_init(&S->empty[i], TPS);
                }
                mbar_init(&S->a_full, TS ? TC_QT : 1);
                mbar_init(&S->a_empty, NA);
                for (int i = 0; i < 2 * TC_ACC; ++i)
                {
                    mbar_init(&S->t_full[i], 1);
                    mbar_init(&S->t_empty[i], TC_QT);
                }
                mbar_fence_init();
                if (PAIR)
                    mbar_fence_init_cluster();
            }
            if (warp == 1)
                tmem_alloc(&S->tmem_base, 512);
            tc_fence_before();
            __syncthreads();
            if (PAIR)
                cluster_sync_all();  // the peer's barriers exist before anything is multicast at them
            tc_fence_after();
            const uint32_t tmem = S->tmem_base;

            if (warp == 0)
            {
                if (lane == 0)
                {
                    int st = 0;
                    uint32_t ph = 0, aph = 0;
                    unsigned long long cur = ~0ull, t = 0, pt = 0;
                    (void)aph;
                    (void)cur;
                    for (unsigned long long u = u0; u < u1; ++u, ++pt)
                    {
                        if (u == u0 || pt == U)
                        {
                            t = u / U;
                            pt = u % U;
                        }
                        if (!TS && t != cur)
                        {
                            mbar_wait_addr(smem_u32(&S->a_empty), aph ^ 1);  // the MMAs of the previous super-tile are done with sA
                            mbar_expect_tx(&S->a_full, NA * a_tile_bytes);
                            tma_bulk_g2s(sA, P.qimg + (size_t)t * NA * (a_tile_bytes / 2), NA * a_tile_bytes,
                                         &S->a_full);
                            aph ^= 1;
                            cur = t;
                        }
                        const uint16_t *src = P.pimg + (size_t)pt * (a_tile_bytes / 2);
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait_addr(smem_u32(&S->empty[st]), ph ^ 1);
                            mbar_expect_tx(&S->full[st], chunk_bytes);
                            if (PAIR)  // my half, into both rings; the peer sends the other half
                                tma_bulk_g2s_mc(sB + (size_t)st * chunk_bytes + crank * (chunk_bytes / 2),
                                                src + (size_t)kc * (chunk_bytes / 2) + crank * (chunk_bytes / 4),
                                                chunk_bytes / 2, &S->full[st], (uint16_t)3);
                            else
                            tma_bulk_g2s(sB + (size_t)st * chunk_bytes, src + (size_t)kc * (chunk_bytes / 2),
                                         chunk_bytes, &S->full[st]);
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
            else if (warp <= NA)
            {
                // one issuer per resident tile: the four accumulator chains (MMA -> selection -> MMA) advance
                // independently instead of queueing behind one thread's barrier waits.  The WHOLE warp runs the
                // loop (uniform control flow, uniform operands); one elected lane issues the tcgen05 instructions.
                // With the loop under `if (lane == 0)` every MMA cost ~16 instructions of register -> uniform
                // register broadcasts and the issuer, not the tensor pipe, paced the d >= 53 shapes (ncu round 2:
                // 110 instructions per 4-MMA chunk, 166 clk per MMA against 64 clk of tensor time).
                {
                    const bool leader = elect_one();
                    const int a = warp - 1;
                    constexpr uint32_t idesc = umma_idesc_bf16(TC_QT, TC_PT);
                    const uint32_t dl = (P.diag & 1) ? TC_SBO : TC_LBO, ds = (P.diag & 1) ? TC_LBO : TC_SBO;
                    const uint32_t a0 = (smem_u32(sA) & 0x3FFFFu) + (uint32_t)a * a_tile_bytes, b0 = smem_u32(sB) & 0x3FFFFu;
                    const uint64_t desc0 = umma_desc(0, dl, ds);  // one MMA covers K = 16 = two column groups
                    int st = 0;
                    uint32_t ph = 0, aph = 0, job = (uint32_t)a;  // job = (unit - u0) * NA + tile
                    unsigned long long cur = ~0ull, t = 0, pt = 0;
                    for (unsigned long long u = u0; u < u1; ++u, job += NA, ++pt)
                    {
                        if (u == u0 || pt == U)
                        {
                            t = u / U;
                            pt = u % U;
                        }
                        if (t != cur)
                        {
                            if (cur != ~0ull)
                                if (leader)
                                    umma_commit(&S->a_empty);
                            mbar_wait_addr(smem_u32(&S->a_full), aph);
                            aph ^= 1;
                            cur = t;
                        }
                        if constexpr (NA == 4)
                        {
                            // K fits one chunk (d <= 20).  The tile owns two 64-column accumulators and fills them
                            // alternately with the two halves of the point tile, so that the selection of one half
                            // overlaps the MMAs and the hand-over of the other.
                            constexpr uint32_t idesc64 = umma_idesc_bf16(TC_QT, TC_PT / 2);
                            mbar_wait_addr(smem_u32(&S->full[st]), ph);
                            const uint64_t ad = desc0 + (a0 >> 4);
                            const uint64_t bd = desc0 + ((b0 + (uint32_t)st * chunk_bytes) >> 4);
#pragma unroll
                            for (uint32_t h = 0; h < 2; ++h)
                            {
                                const uint32_t jt = 2 * (job >> 2) + h, b = 2 * (uint32_t)a + (jt & 1);
                                mbar_wait_addr(smem_u32(&S->t_empty[b]), ((jt >> 1) & 1) ^ 1);
                                tc_fence_after();
                                const uint64_t bh = bd + h * ((TC_PT / 2 / 8) * TC_SBO >> 4);  // rows 64 h .. 64 h + 63
                                if (leader && !(DBG && (P.diag & 8)))
                                {
                                    // k16 <= 4 here; spelled out: a counted loop costs the issuer ~2 % of the kernel
                                    umma_bf16(tmem + b * (TC_PT / 2), ad, bh, idesc64, 0u);
                                    if (k16 > 1)
                                        umma_bf16(tmem + b * (TC_PT / 2), ad + (2 * TC_LBO >> 4), bh + (2 * TC_LBO >> 4), idesc64, 1u);
                                    if (k16 > 2)
                                    {
                                        umma_bf16(tmem + b * (TC_PT / 2), ad + 2 * (2 * TC_LBO >> 4), bh + 2 * (2 * TC_LBO >> 4), idesc64, 1u);
                                        if (k16 > 3)
                                            umma_bf16(tmem + b * (TC_PT / 2), ad + 3 * (2 * TC_LBO >> 4), bh + 3 * (2 * TC_LBO >> 4), idesc64, 1u);
                                    }
                                }
                                if (leader)
                                    umma_commit(&S->t_full[b]);
                                __syncwarp();
                            }
                            if (leader)
                                umma_commit(&S->empty[st]);
                            __syncwarp();
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                            continue;
                        }
                        const uint32_t buf = job & (NBUF - 1);
                        const uint32_t dcol = tmem + ACC0 + buf * TC_PT;
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait_addr(smem_u32(&S->full[st]), ph);
                            if (kc == 0)
                                mbar_wait_addr(smem_u32(&S->t_empty[buf]), ((job / NBUF) & 1) ^ 1);
                            tc_fence_after();
                            // MMAs of this chunk: the last chunk stops where K ends (the rest is zero padding)
                            const uint32_t nm = (uint32_t)min(kw / 2, k16 - kc * (kw / 2));
                            // descriptors differ in the start address only (low word, units of 16 bytes)
                            const uint64_t ad = desc0 + ((a0 + (uint32_t)kc * chunk_bytes) >> 4);
                            const uint64_t bd = desc0 + ((b0 + (uint32_t)st * chunk_bytes) >> 4);
                            if (TS)
                            {
                                // A: 8 columns per K = 16 step, chunk kc starts at column kc * kw * 4
                                const uint32_t ac = tmem + (uint32_t)kc * (uint32_t)kw * 4u;
                                if (leader && !(DBG && (P.diag & 8)))
                                {
                                    umma_bf16_ts(dcol, ac, bd, idesc, kc ? 1u : 0u);
                                    for (uint32_t i = 1; i < nm; ++i)
                                        umma_bf16_ts(dcol, ac + 8 * i, bd + i * (2 * TC_LBO >> 4), idesc, 1u);
                                }
                            }
                            else if (leader && !(DBG && (P.diag & 8)))  // diagnostics: barriers and commits only
                            {
                                umma_bf16(dcol, ad, bd, idesc, kc ? 1u : 0u);
                                for (uint32_t i = 1; i < nm; ++i)
                                    umma_bf16(dcol, ad + i * (2 * TC_LBO >> 4), bd + i * (2 * TC_LBO >> 4), idesc, 1u);
                            }
                            if (leader)
                            {
                                // the stage is free once all NA issuers' (PAIR: both CTAs') MMAs have read it
                                if (PAIR)
                                    umma_commit_mc(&S->empty[st], (uint16_t)3);
                                else
                                    umma_commit(&S->empty[st]);
                                if (kc == nkc - 1)
                                    umma_commit(&S->t_full[buf]);
                            }
                            __syncwarp();
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
            else
            {
                // selection: warp (a, wq) owns TMEM lane quadrant wq of the accumulators of resident
                // tile a; a thread's lane is its query.  The query's list lives in shared memory,
                // LS = L + 4 floats apart (conflict-free LDS.128 across lanes); values carry their list
                // slot in the low mantissa bits (knn_scan.cuh: tag_up), so the running maximum `thr`
                // names the entry to overwrite and ONE max-tree over the list yields the next threshold
                // and the next victim.  The hot loop is two copies of the W-column step (the two register
                // buffers of the TMEM load pipeline) and stays resident in the instruction cache.
                const int a = (warp - 1 - NA) >> 2, wq = warp & 3, qrow = wq * 32 + lane;
                const int qrow_idx = qrow;
                uint32_t aph_sel = 0;
                (void)qrow_idx;
                (void)aph_sel;
                float *lvq = lv + (a * TC_QT + qrow) * LS;
                uint32_t *liq = li + a * L * TC_QT + qrow;
                float *scr_q = scr + (warp - 1 - NA) * W * 32 + lane;
                float thr = 0.f;
                auto reset = [&]() {
                    for (int i = 0; i < L; ++i)
                    {
                        lvq[i] = __uint_as_float(EMPTY_BITS | (uint32_t)i);
                        liq[i * TC_QT] = SLOT_NONE;
                    }
                    thr = __uint_as_float(EMPTY_BITS | (uint32_t)(L - 1));
                };
                auto flush = [&](unsigned long long t) {
                    const unsigned long long x = t * U;
                    const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
                    const unsigned seg = (unsigned)(g - gfirst);
                    const size_t q = ((size_t)t * TPS + (PAIR ? (int)crank : a)) * TC_QT + qrow;
                    P.cand_thr[(size_t)seg * P.qpad + q] = thr;
                    for (int i = 0; i < L; ++i)
                        P.cand_ids[((size_t)seg * L + i) * P.qpad + q] = liq[i * TC_QT];
                    if (P.list_vals)
                        for (int i = 0; i < L; ++i)
                            P.list_vals[((size_t)seg * L + i) * P.qpad + q] = lvq[i];
                };
                // one W-column step: min-tree against the threshold; the rare hit stages the values in
                // shared memory and calls the out-of-line routine
                auto step = [&](uint32_t(&v)[W], uint32_t id0) {
                    float m0 = fminf(__uint_as_float(v[0]), __uint_as_float(v[1]));
                    float m1 = fminf(__uint_as_float(v[2]), __uint_as_float(v[3]));
#pragma unroll
                    for (int i = 4; i < W; i += 4)
                    {
                        m0 = fminf(m0, fminf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
                        m1 = fminf(m1, fminf(__uint_as_float(v[i + 2]), __uint_as_float(v[i + 3])));
                    }
                    const float mn = fminf(m0, m1);
                    if (mn < thr && !(DBG && (P.diag & 4)))
                    {
                        // the columns that beat the threshold at entry; m0 covers columns 0,1 mod 4, m1 columns 2,3
                        // mod 4, so usually only one of the two halves has to be looked at
                        uint32_t mask = 0;
                        if (m0 < thr)
                        {
#pragma unroll
                            for (int i = 0; i < W; i += 4)
                            {
                                mask |= __uint_as_float(v[i]) < thr ? 1u << i : 0u;
                                mask |= __uint_as_float(v[i + 1]) < thr ? 2u << i : 0u;
                            }
                        }
                        if (m1 < thr)
                        {
#pragma unroll
                            for (int i = 0; i < W; i += 4)
                            {
                                mask |= __uint_as_float(v[i + 2]) < thr ? 4u << i : 0u;
                                mask |= __uint_as_float(v[i + 3]) < thr ? 8u << i : 0u;
                            }
                        }
                        if ((mask & (mask - 1)) == 0)
                            thr = tc_insert_one<L>(thr, mn, lvq, liq, id0 + (uint32_t)__ffs(mask) - 1);
                        else
                        {
#pragma unroll
                            for (int i = 0; i < W; ++i)
                                scr_q[i * 32] = __uint_as_float(v[i]);
                            thr = tc_select_hits<L, W>(thr, mask, scr_q, lvq, liq, id0);
                        }
                    }
                };
                // loop invariants the compiler does not hoist by itself (they hang off run-time shapes): the
                // shared addresses of this tile's two accumulator barriers and the TMEM address of its lane quadrant
                const uint32_t tfull0 = smem_u32(&S->t_full[2 * a]), tempty0 = smem_u32(&S->t_empty[2 * a]);
                const uint32_t tq = tmem + ((uint32_t)(wq * 32) << 16);
                (void)tfull0;
                (void)tempty0;
                unsigned long long cur = ~0ull, job = a, t = 0, pt = 0;
                for (unsigned long long u = u0; u < u1; ++u, job += NA, ++pt)
                {
                    if (u == u0 || pt == U)
                    {
                        t = u / U;  // once per super-tile: no 64-bit division in the unit loop
                        pt = u % U;
                        if (cur != ~0ull)
                            flush(cur);
                        reset();
                        cur = t;
                        if constexpr (TS)
                        {
                            // this thread's query row of the operand image -> its TMEM lane, columns [0, kpad / 2)
                            mbar_wait_addr(smem_u32(&S->a_empty), aph_sel ^ 1);  // the MMAs of the previous super-tile have read A
                            aph_sel ^= 1;
                            tc_fence_after();
                            const uint16_t *qrow = P.qimg + (size_t)(t * TPS + crank) * (a_tile_bytes / 2) +
                                                   (size_t)(qrow_idx >> 3) * (TC_SBO / 2) + (size_t)(qrow_idx & 7) * 8;
                            const uint32_t ta = tmem + ((uint32_t)(wq * 32) << 16);
                            for (int kc = 0; kc < nkc; ++kc)
                                for (int j = 0; j < kw; ++j)
                                {
                                    const uint4 v = *reinterpret_cast<const uint4 *>(
                                        qrow + (size_t)kc * (chunk_bytes / 2) + (size_t)j * (TC_LBO / 2));
                                    __syncwarp();
                                    tmem_st4(ta + (uint32_t)(kc * kw + j) * 4u, v.x, v.y, v.z, v.w);
                                }
                            tmem_st_wait();
                            tc_fence_before();
                            mbar_arrive(&S->a_full);
                        }
                    }
                    const uint32_t slot0 = (uint32_t)(pt * TC_PT);
                    if constexpr (NA == 4)
                    {
                        // two 64-column half-jobs per point tile (see the issuer); 21 warps per CTA hide the
                        // TMEM load latency, one register buffer keeps the thread within 96 registers
                        // half-job h of unit job / 4 uses buffer h of this tile; both halves share the unit's parity
                        const uint32_t par = (uint32_t)(job >> 2) & 1u;
#pragma unroll
                        for (uint32_t h = 0; h < 2; ++h)
                        {
                            const uint32_t bb = h, b = 2 * (uint32_t)a + bb;
                            mbar_wait_addr(tfull0 + 8 * bb, par);
                            tc_fence_after();
                            const uint32_t taddr = tq + b * (TC_PT / 2);
                            if (!(DBG && (P.diag & 2)))
                            {
                                // (both steps spelled out: the compiler schedules across them, 6.70 -> 5.92 ms at 1M states;
                                // two register buffers with both loads in flight spill at the 80 registers 21 warps leave: 6.56 ms)
                                uint32_t ra[W];
#pragma unroll
                                for (int c = 0; c < TC_PT / 2 / W; ++c)
                                {
                                    tmem_ld_issue(taddr + c * W, ra);
                                    tmem_ld_wait(ra);
                                    if (DBG && P.dbg)
                                        tc_dump<W>(P, ra, ((size_t)t * TPS + (PAIR ? (int)crank : a)) * TC_QT + qrow, slot0 + h * (TC_PT / 2) + c * W);
                                    step(ra, slot0 + h * (TC_PT / 2) + c * W);
                                }
                            }
                            tc_fence_before();
                            mbar_arrive_addr(tempty0 + 8 * bb);
                        }
                        continue;
                    }
                    const int buf = (int)(job & (NBUF - 1));
                    mbar_wait_addr(smem_u32(&S->t_full[buf]), (uint32_t)((job / NBUF) & 1));
                    tc_fence_after();
                    const uint32_t taddr = tmem + ((uint32_t)(wq * 32) << 16) + ACC0 + (uint32_t)buf * TC_PT;
                    if (!(DBG && (P.diag & 2)))
                    {
                        {
                            uint32_t ra[W], rb[W];
                            tmem_ld_issue(taddr, ra);
#pragma unroll
                            for (int c = 0; c < TC_PT / W; c += 2)
                            {
                                tmem_ld_wait(ra);
                                tmem_ld_issue(taddr + (c + 1) * W, rb);  // in flight during the tree below
                                if (DBG && P.dbg)
                                    tc_dump<W>(P, ra, ((size_t)t * TPS + (PAIR ? (int)crank : a)) * TC_QT + qrow, slot0 + c * W);
                                step(ra, slot0 + c * W);
                                tmem_ld_wait(rb);
                                if (c + 2 < TC_PT / W)
                                    tmem_ld_issue(taddr + (c + 2) * W, ra);
                                if (DBG && P.dbg)
                                    tc_dump<W>(P, rb, ((size_t)t * TPS + (PAIR ? (int)crank : a)) * TC_QT + qrow, slot0 + (c + 1) * W);
                                step(rb, slot0 + (c + 1) * W);
                            }
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(&S->t_empty[buf]);
                }
                if (cur != ~0ull)
                    flush(cur);
            }
            tc_fence_before();
            __syncthreads();
            if (PAIR)
                cluster_sync_all();  // no CTA leaves while the peer may still multicast into it
            if (warp == 1)
            {
                __syncwarp();
                tc_fence_after();
                tmem_dealloc(tmem, 512);
            }
        }

        // ---------------- the threshold scan ----------------
        // Same operand images, producer and issuer roles as knn_scan_tc_kernel, but the selection keeps no
        // list: every query has a fixed bound tau (nn.cu: the r-th smallest scan value of a strided sample
        // of the points, an exact distance bound, or a radius, all mapped to a scan value), and every slot
        // whose scan value is below it is appended to the query's row of this CTA's segment -- one
        // predicated store and one add per hit, a register counter, no shared-memory state per query, no
        // convergence phase.  The exact epilogue (nn.cu) evaluates every kept slot in FP64 and proves
        // that nothing dropped can matter (any point within the k-th exact distance has a scan value
        // below tau, or the query is redone exactly).  A job is a whole 128-column accumulator
        // (job j -> TMEM buffer j % 4), read W columns at a time through two register buffers.
        constexpr int TC_THR_MAX_STAGES = 12;
        struct TcThrSmem
        {
            uint64_t full[TC_THR_MAX_STAGES], empty[TC_THR_MAX_STAGES];
            uint64_t a_full, a_empty;
            uint64_t t_full[TC_ACC], t_empty[TC_ACC];
            uint32_t tmem_base;
        };

        __device__ __forceinline__ float min8(const uint32_t *v)
        {
            float x = fminf(__uint_as_float(v[0]), fminf(__uint_as_float(v[1]), __uint_as_float(v[2])));
            x = fminf(x, fminf(__uint_as_float(v[3]), __uint_as_float(v[4])));
            x = fminf(x, fminf(__uint_as_float(v[5]), __uint_as_float(v[6])));
            return fminf(x, __uint_as_float(v[7]));
        }

        template <int NA, int W, bool DBG>
        __global__ void __launch_bounds__(tc_threads(NA), 1) knn_thr_scan_tc_kernel(const TcScanParams P)
        {
            extern __shared__ __align__(1024) uint8_t tc_smem[];
            static_assert(W == 16 || W == 32, "columns per selection step");
            const int nkc = P.sh.nkc, nst = P.nstages, kw = P.sh.kw, k16 = P.sh.k16;
            const uint32_t chunk_bytes = (uint32_t)kw * 2048u;
            const uint32_t a_tile_bytes = (uint32_t)nkc * chunk_bytes;
            uint8_t *sA = tc_smem;                         // [NA][nkc][chunk]
            uint8_t *sB = sA + (size_t)NA * a_tile_bytes;  // [nst][chunk]
            TcThrSmem *S = reinterpret_cast<TcThrSmem *>(sB + (size_t)nst * chunk_bytes);

            const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
            const unsigned long long TU = P.total_units, U = P.units, G = gridDim.x, g = blockIdx.x;
            const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;

            if (tid == 0)
            {
                for (int i = 0; i < TC_THR_MAX_STAGES; ++i)
                {
                    mbar_init(&S->full[i], 1);
                    mbar_init(&S->empty[i], NA);
                }
                mbar_init(&S->a_full, 1);
                mbar_init(&S->a_empty, NA);
                for (int i = 0; i < TC_ACC; ++i)
                {
                    mbar_init(&S->t_full[i], 1);
                    mbar_init(&S->t_empty[i], TC_QT);
                }
                mbar_fence_init();
            }
            if (warp == 1)
                tmem_alloc(&S->tmem_base, 512);
            tc_fence_before();
            __syncthreads();
            tc_fence_after();
            const uint32_t tmem = S->tmem_base;

            if (warp == 0)
            {
                if (lane == 0)
                {
                    int st = 0;
                    uint32_t ph = 0, aph = 0;
                    unsigned long long cur = ~0ull, t = 0, pt = 0;
                    for (unsigned long long u = u0; u < u1; ++u, ++pt)
                    {
                        if (u == u0 || pt == U)
                        {
                            t = u / U;
                            pt = u % U;
                        }
                        if (t != cur)
                        {
                            mbar_wait_addr(smem_u32(&S->a_empty), aph ^ 1);
                            mbar_expect_tx(&S->a_full, NA * a_tile_bytes);
                            tma_bulk_g2s(sA, P.qimg + (size_t)t * NA * (a_tile_bytes / 2), NA * a_tile_bytes,
                                         &S->a_full);
                            aph ^= 1;
                            cur = t;
                        }
                        const uint16_t *src = P.pimg + (size_t)pt * (a_tile_bytes / 2);
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait_addr(smem_u32(&S->empty[st]), ph ^ 1);
                            mbar_expect_tx(&S->full[st], chunk_bytes);
                            tma_bulk_g2s(sB + (size_t)st * chunk_bytes, src + (size_t)kc * (chunk_bytes / 2),
                                         chunk_bytes, &S->full[st]);
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
            else if (warp <= NA)
            {
                // whole warp, one elected lane issues (see knn_scan_tc_kernel)
                {
                    const bool leader = elect_one();
                    const int a = warp - 1;
                    constexpr uint32_t idesc = umma_idesc_bf16(TC_QT, TC_PT);
                    const uint32_t a0 = (smem_u32(sA) & 0x3FFFFu) + (uint32_t)a * a_tile_bytes, b0 = smem_u32(sB) & 0x3FFFFu;
                    const uint64_t desc0 = umma_desc(0, TC_LBO, TC_SBO);
                    const bool no_mma = (P.diag & 8) != 0;
                    int st = 0;
                    uint32_t ph = 0, aph = 0, job = (uint32_t)a;  // job = (unit - u0) * NA + tile
                    unsigned long long cur = ~0ull, t = 0, pt = 0;
                    for (unsigned long long u = u0; u < u1; ++u, job += NA, ++pt)
                    {
                        if (u == u0 || pt == U)
                        {
                            t = u / U;
                            pt = u % U;
                        }
                        if (t != cur)
                        {
                            if (cur != ~0ull && leader)
                                umma_commit(&S->a_empty);
                            mbar_wait_addr(smem_u32(&S->a_full), aph);
                            aph ^= 1;
                            cur = t;
                        }
                        const uint32_t buf = job & (TC_ACC - 1);
                        const uint32_t dcol = tmem + buf * TC_PT;
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait_addr(smem_u32(&S->full[st]), ph);
                            if (kc == 0)
                                mbar_wait_addr(smem_u32(&S->t_empty[buf]), ((job / TC_ACC) & 1) ^ 1);
                            tc_fence_after();
                            const uint32_t nm = (uint32_t)min(kw / 2, k16 - kc * (kw / 2));
                            const uint64_t ad = desc0 + ((a0 + (uint32_t)kc * chunk_bytes) >> 4);
                            const uint64_t bd = desc0 + ((b0 + (uint32_t)st * chunk_bytes) >> 4);
                            if (leader && !no_mma)
                            {
                                umma_bf16(dcol, ad, bd, idesc, kc ? 1u : 0u);
                                for (uint32_t i = 1; i < nm; ++i)
                                    umma_bf16(dcol, ad + i * (2 * TC_LBO >> 4), bd + i * (2 * TC_LBO >> 4), idesc, 1u);
                            }
                            if (leader)
                            {
                                umma_commit(&S->empty[st]);
                                if (kc == nkc - 1)
                                    umma_commit(&S->t_full[buf]);
                            }
                            __syncwarp();
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
            else
            {
                // selection: warp (a, wq) owns TMEM lane quadrant wq of the accumulators of resident tile a;
                // a thread's lane is its query
                const int a = (warp - 1 - NA) >> 2, wq = warp & 3, qrow = wq * 32 + lane;
                const bool no_sel = (P.diag & 2) != 0, no_app = (P.diag & 4) != 0;
                const uint32_t cap_w = P.cap - W;
                float thr = 0.f;
                uint32_t cnt = 0;
                uint32_t *row = P.app_ids;
                size_t qslot = 0;  // seg * qpad + q of the current (segment, query)
                auto begin = [&](unsigned long long t) {
                    const unsigned long long x = t * U;
                    const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
                    const unsigned seg = (unsigned)(g - gfirst);
                    const size_t q = ((size_t)t * NA + a) * TC_QT + qrow;
                    qslot = (size_t)seg * P.qpad + q;
                    thr = P.tau[q];
                    cnt = 0;
                    row = P.app_ids + qslot * P.cap;
                };
                // one W-column step: a min-tree against the bound; on a hit the columns below it are
                // appended in column order.  Eight columns share a partial minimum, so a hit usually
                // looks at one group only.
                auto step = [&](uint32_t(&v)[W], uint32_t id0) {
                    float p[W / 8];
#pragma unroll
                    for (int j = 0; j < W / 8; ++j)
                        p[j] = min8(&v[8 * j]);
                    float mn = fminf(p[0], p[1]);
                    if constexpr (W == 32)
                        mn = fminf(mn, fminf(p[2], p[3]));
                    if (mn < thr && !no_app)
                    {
                        if (cnt <= cap_w)
                        {
#pragma unroll
                            for (int j = 0; j < W / 8; ++j)
                                if (p[j] < thr)
                                {
#pragma unroll
                                    for (int i = 0; i < 8; ++i)
                                        if (__uint_as_float(v[8 * j + i]) < thr)
                                        {
                                            row[cnt] = id0 + 8 * j + i;
                                            ++cnt;
                                        }
                                }
                        }
                        else
                            cnt = 0xFFFFFFFFu;  // the row is full: the count says so, the epilogue redoes the query exactly
                    }
                };
                unsigned long long cur = ~0ull, job = a, t = 0, pt = 0;
                for (unsigned long long u = u0; u < u1; ++u, job += NA, ++pt)
                {
                    if (u == u0 || pt == U)
                    {
                        t = u / U;  // once per super-tile: no 64-bit division in the unit loop
                        pt = u % U;
                        if (cur != ~0ull)
                            P.app_cnt[qslot] = cnt;
                        begin(t);
                        cur = t;
                    }
                    const uint32_t slot0 = (uint32_t)(pt * TC_PT);
                    const int buf = (int)(job & (TC_ACC - 1));
                    mbar_wait_addr(smem_u32(&S->t_full[buf]), (uint32_t)((job / TC_ACC) & 1));
                    tc_fence_after();
                    const uint32_t taddr = tmem + ((uint32_t)(wq * 32) << 16) + (uint32_t)buf * TC_PT;
                    if (!no_sel)
                    {
                        uint32_t ra[W], rb[W];
                        __syncwarp();
                        tmem_ld_issue(taddr, ra);
#pragma unroll
                        for (int c = 0; c < TC_PT / W; c += 2)
                        {
                            __syncwarp();
                            tmem_ld_wait(ra);
                            tmem_ld_issue(taddr + (c + 1) * W, rb);  // in flight during the tree below
                            if (DBG)
                                tc_dump<W>(P, ra, ((size_t)t * NA + a) * TC_QT + qrow, slot0 + c * W);
                            step(ra, slot0 + c * W);
                            __syncwarp();
                            tmem_ld_wait(rb);
                            if (c + 2 < TC_PT / W)
                                tmem_ld_issue(taddr + (c + 2) * W, ra);
                            if (DBG)
                                tc_dump<W>(P, rb, ((size_t)t * NA + a) * TC_QT + qrow, slot0 + (c + 1) * W);
                            step(rb, slot0 + (c + 1) * W);
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(&S->t_empty[buf]);
                }
                if (cur != ~0ull)
                    P.app_cnt[qslot] = cnt;
            }
            tc_fence_before();
            __syncthreads();
            if (warp == 1)
            {
                __syncwarp();
                tc_fence_after();
                tmem_dealloc(tmem, 512);
            }
        }
    }  // namespace

    size_t tc_smem_bytes(const TcShape &sh, int nstages, int L, bool ts)
    {
        const size_t chunk = (size_t)sh.kw * 2048;
        return (ts ? 0 : (size_t)sh.na * sh.nkc * chunk) + (size_t)nstages * chunk +
               (size_t)sh.na * TC_QT * (2 * L + 4) * 4 + (size_t)4 * sh.na * 32 * 32 * 4 + sizeof(TcSmem) + 64;
    }

    int tc_pick_stages(const TcShape &sh, int L, bool ts)
    {
        for (int st = TC_MAX_STAGES; st >= 2; --st)
            if (tc_smem_bytes(sh, st, L, ts) <= 227 * 1024)
                return st;
        return 0;
    }

    TcShape tc_shape(int D)
    {
        static const int mc = [] {
            const char *e = std::getenv("OMPLCU_TC_CHUNK_MMAS");
            const int x = e ? std::atoi(e) : 0;
            return x > 0 && x <= 32 ? x : 10;
        }();
        TcShape s;
        s.k16 = (3 * D + 3 + 15) / 16;
        s.na = s.k16 <= 4 ? 4 : (s.k16 <= 10 ? 2 : 1);  // d <= 20: 4 tiles (<= 32 KB of query operand), d <= 52: 2, else 1
        s.nkc = 1;
        s.kw = 2 * s.k16;
        if (s.na == 4)
            return s;  // the 4-tile kernel wants K in one chunk (<= 16 KB)
        // measured at d = 100 / 64 / 32 (profiles/r2_experiments.md): a padded column group costs its bytes, one more
        // chunk about as much as 0.6 MMA slots
        long best = -1;
        for (int nkc = (s.k16 + mc - 1) / mc; nkc <= s.k16; ++nkc)
        {
            TcShape c = s;
            c.nkc = nkc;
            c.kw = 2 * ((s.k16 + nkc - 1) / nkc);
            // a k <= 10 query must get 3 stages and 48 KB in flight (query tile in tensor memory where the shape
            // allows it), and 2 stages with the tile in shared memory (option tc_ts = 0, diagnostics)
            const int st = tc_pick_stages(c, 16, tc_ts_usable(c));
            if (st < 3 || (size_t)st * c.kw * 2048 < 48 * 1024 || tc_pick_stages(c, 16, false) < 2)
                continue;
            const long cost = 10L * c.nkc * (c.kw / 2) + 6L * c.nkc;
            if (best < 0 || cost < best)
            {
                best = cost;
                s = c;
            }
        }
        return s;
    }

    bool tc_ts_usable(const TcShape &sh)
    {
        return sh.na == 1 && sh.kpad() <= 512;  // the tile takes kpad / 2 of the 256 TMEM columns below the accumulators
    }

    void tc_build_point_image(const float *filt, int rowf, int D, size_t units, uint16_t *out, cudaStream_t s)
    {
        const TcShape sh = tc_shape(D);
        const size_t rows = units * TC_PT, n = rows * (sh.kpad() / 8);
        if (n)
            tc_point_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(filt, rowf, D, sh, rows, rows, 1, out);
    }

    void tc_build_sample_image(const float *filt, int rowf, int D, size_t src_rows, size_t row_step, size_t units,
                               uint16_t *out, cudaStream_t s)
    {
        const TcShape sh = tc_shape(D);
        const size_t rows = units * TC_PT, n = rows * (sh.kpad() / 8);
        if (n)
            tc_point_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(filt, rowf, D, sh, rows, src_rows,
                                                                              row_step, out);
    }

    void tc_build_query_image(const float *qp, uint32_t qpad, int D, uint16_t *out, cudaStream_t s)
    {
        const TcShape sh = tc_shape(D);
        const size_t n = (size_t)qpad * (sh.kpad() / 8);
        if (n)
            tc_query_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(qp, qpad, D, sh, out);
    }

    namespace
    {
        template <int L, int NA>
        void tc_launch(const TcScanParams &P, unsigned grid, cudaStream_t s)
        {
            const size_t smem = tc_smem_bytes(P.sh, P.nstages, L);
            if (P.dbg || (P.diag & ~1))  // raw-value dump or a diagnostic variant: the instrumented instantiation
            {
                CUDA_CHECK(cudaFuncSetAttribute(knn_scan_tc_kernel<L, NA, true>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                knn_scan_tc_kernel<L, NA, true><<<grid, tc_threads(NA), smem, s>>>(P);
                return;
            }
            CUDA_CHECK(cudaFuncSetAttribute(knn_scan_tc_kernel<L, NA, false>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            knn_scan_tc_kernel<L, NA, false><<<grid, tc_threads(NA), smem, s>>>(P);
        }
        template <int L>
        void tc_launch_ts(const TcScanParams &P, unsigned grid, cudaStream_t s)
        {
            const size_t smem = tc_smem_bytes(P.sh, P.nstages, L, true);
            CUDA_CHECK(cudaFuncSetAttribute(knn_scan_tc_kernel<L, 1, false, true>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            if (P.diag & ~1)  // diagnostic switches live in the instrumented instantiation
            {
                CUDA_CHECK(cudaFuncSetAttribute(knn_scan_tc_kernel<L, 1, true, true>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                knn_scan_tc_kernel<L, 1, true, true><<<grid, tc_threads(1), smem, s>>>(P);
                return;
            }
            knn_scan_tc_kernel<L, 1, false, true><<<grid, tc_threads(1), smem, s>>>(P);
        }
        template <int L>
        void tc_launch_pair(const TcScanParams &P, unsigned grid, cudaStream_t s)  // grid: CTAs, even
        {
            const size_t smem = tc_smem_bytes(P.sh, P.nstages, L, true);
            auto kern = (P.diag & ~1) ? knn_scan_tc_kernel<L, 1, true, true, true> : knn_scan_tc_kernel<L, 1, false, true, true>;
            CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(grid);
            cfg.blockDim = dim3(tc_threads(1));
            cfg.dynamicSmemBytes = smem;
            cfg.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, P));
        }
        template <int L>
        void tc_launch_na(const TcScanParams &P, unsigned grid, cudaStream_t s)
        {
            if (P.ts && P.pair)
                tc_launch_pair<L>(P, grid, s);
            else if (P.ts)  // decided by the caller (tc_ts_usable, no diagnostics): the ring depth was sized for it
                tc_launch_ts<L>(P, grid, s);
            else if (P.sh.na == 4)
                tc_launch<L, 4>(P, grid, s);
            else if (P.sh.na == 2)
                tc_launch<L, 2>(P, grid, s);
            else
                tc_launch<L, 1>(P, grid, s);
        }
    }  // namespace

    size_t tc_thr_smem_bytes(const TcShape &sh, int nstages)
    {
        const size_t chunk = (size_t)sh.kw * 2048;
        return (size_t)sh.na * sh.nkc * chunk + (size_t)nstages * chunk + sizeof(TcThrSmem) + 64;
    }

    int tc_thr_pick_stages(const TcShape &sh)
    {
        for (int st = TC_THR_MAX_STAGES; st >= 3; --st)
            if (tc_thr_smem_bytes(sh, st) <= 227 * 1024)
                return st;
        return 0;
    }

    namespace
    {
        template <int NA, int W>
        void tc_thr_launch(const TcScanParams &P, unsigned grid, cudaStream_t s)
        {
            const size_t smem = tc_thr_smem_bytes(P.sh, P.nstages);
            if (P.dbg)
            {
                CUDA_CHECK(cudaFuncSetAttribute(knn_thr_scan_tc_kernel<NA, W, true>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                knn_thr_scan_tc_kernel<NA, W, true><<<grid, tc_threads(NA), smem, s>>>(P);
                return;
            }
            CUDA_CHECK(cudaFuncSetAttribute(knn_thr_scan_tc_kernel<NA, W, false>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            knn_thr_scan_tc_kernel<NA, W, false><<<grid, tc_threads(NA), smem, s>>>(P);
        }
    }  // namespace

    void tc_launch_thr_scan(const TcScanParams &P, unsigned grid, cudaStream_t s)
    {
        REQUIRE(P.cap >= 64 && P.cap % 32 == 0, "threshold scan: bad row capacity");
        const bool w16 = (P.diag & 16) != 0;
        if (P.sh.na == 4)
            w16 ? tc_thr_launch<4, 16>(P, grid, s) : tc_thr_launch<4, 32>(P, grid, s);
        else if (P.sh.na == 2)
            w16 ? tc_thr_launch<2, 16>(P, grid, s) : tc_thr_launch<2, 32>(P, grid, s);
        else
            w16 ? tc_thr_launch<1, 16>(P, grid, s) : tc_thr_launch<1, 32>(P, grid, s);
    }

    void tc_launch_scan(int L, const TcScanParams &P, unsigned grid, cudaStream_t s)
    {
        if (L == 8)
            tc_launch_na<8>(P, grid, s);
        else if (L == 16)
            tc_launch_na<16>(P, grid, s);
        else if (L == 24)
            tc_launch_na<24>(P, grid, s);
        else if (L == 48)
            tc_launch_na<48>(P, grid, s);
        else
            throw Error(OMPLCU_ERR_INTERNAL, "tensor-core scan: unsupported list length");
    }
}  // namespace omplcu
