// knn_scan_tc.cu -- tensor-core filter scan (tcgen05.mma, BF16 split operands, FP32
// accumulators in TMEM) for k-nearest queries in 64..128 dimensions.  See knn_scan_tc.cuh
// for the contract and DESIGN.md 3.5 for the error bound and the roofline.
//
// One persistent CTA per SM (192 threads), warp-specialised:
//   warp 0, one lane   producer: cp.async.bulk (TMA bulk copy) of the query image (once per
//                      query tile, stays resident) and of 16 KB point chunks into a ring
//   warp 1, one lane   tcgen05.mma issuer: per point tile nkc x 4 MMAs (M 128 x N 128 x K 16)
//                      into one of four 128-column TMEM accumulators; tcgen05.commit frees the
//                      ring slot and hands the accumulator to the selection warps
//   warps 2..5         selection: thread = TMEM lane = query; tcgen05.ld 32 columns at a time,
//                      one FMNMX tree against the query's threshold (a register), the rare hit
//                      goes into the query's private list in shared memory
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
        constexpr int TC_THREADS = 192;
        constexpr int TC_ACC = 4;  // accumulator buffers (4 x 128 = all 512 TMEM columns)
        constexpr int TC_MAX_STAGES = 8;

        // ---------------- operand images ----------------
        // BF16 element offset of (row, k) inside one 128-row image block
        __host__ __device__ inline size_t tc_img_off(int row, int k, uint32_t lbo, uint32_t sbo)
        {
            const int kc = k / TC_KC, kk = k % TC_KC, j = kk >> 3, e = kk & 7;
            return (size_t)kc * TC_CHUNK_ELEMS + (size_t)j * (lbo / 2) + (size_t)(row >> 3) * (sbo / 2) +
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
        // x = p[0] + p[1] + p[2] (+ residual <= 2^-26 |x|; zero for every FP32 x in range)
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

        // role of column k for a point row (B operand) / a query row (A operand)
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

        // one thread writes 8 consecutive K elements (16 bytes) of one row
        __global__ void tc_point_image_kernel(const float *filt, int rowf, int D, int Kp, size_t rows, uint32_t lbo,
                                              uint32_t sbo, uint16_t *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            const int groups = Kp / 8;
            if (i >= rows * groups)
                return;
            // consecutive threads take consecutive rows of the same K group: 16-byte stores that
            // are contiguous across the 8 rows of a core matrix
            const size_t tile = i / ((size_t)TC_PT * groups);
            const size_t rem = i % ((size_t)TC_PT * groups);
            const int grp = (int)(rem / TC_PT), r = (int)(rem % TC_PT);
            const float *row = filt + (tile * TC_PT + r) * rowf;
            uint16_t v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
                v[e] = bf16_bits(tc_point_elem(row, D, grp * 8 + e));
            uint16_t *dst = out + tile * (size_t)(Kp / TC_KC) * TC_CHUNK_ELEMS + tc_img_off(r, grp * 8, lbo, sbo);
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(v);
        }

        // qp is [c][qpad] = -2 * qt_c (query_prep_kernel in nn.cu); padded queries are zero rows
        __global__ void tc_query_image_kernel(const float *qp, uint32_t qpad, int D, int Kp, uint32_t lbo,
                                              uint32_t sbo, uint16_t *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            const int groups = Kp / 8;
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
            uint16_t *dst = out + tile * (size_t)(Kp / TC_KC) * TC_CHUNK_ELEMS + tc_img_off(r, grp * 8, lbo, sbo);
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
        // 32 consecutive columns of this thread's TMEM lane
        __device__ __forceinline__ void tmem_ld32(uint32_t taddr, float v[32])
        {
            uint32_t r[32];
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
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 32; ++i)
                v[i] = __uint_as_float(r[i]);
        }

        // ---------------- the scan ----------------
        struct TcSmem
        {
            uint64_t full[TC_MAX_STAGES], empty[TC_MAX_STAGES];  // point-chunk ring: TMA <-> MMA
            uint64_t a_full, a_empty;                            // query image:     TMA <-> MMA
            uint64_t t_full[TC_ACC], t_empty[TC_ACC];            // accumulators:    MMA <-> selection
            uint32_t tmem_base;
        };

        template <int L>
        __global__ void __launch_bounds__(TC_THREADS, 1) knn_scan_tc_kernel(const TcScanParams P)
        {
            extern __shared__ __align__(1024) uint8_t tc_smem[];
            const int nkc = P.nkc, nst = P.nstages;
            uint8_t *sA = tc_smem;                                  // [nkc][16 KB]
            uint8_t *sB = sA + (size_t)nkc * TC_CHUNK_ELEMS * 2;    // [nst][16 KB]
            float *lv = reinterpret_cast<float *>(sB + (size_t)nst * TC_CHUNK_ELEMS * 2);  // [128][L + 4]
            uint32_t *li = reinterpret_cast<uint32_t *>(lv + (L + 4) * TC_QT);             // [L][128]
            float *scr = reinterpret_cast<float *>(li + L * TC_QT);                        // [32][128] slow-path staging
            TcSmem *S = reinterpret_cast<TcSmem *>(scr + 32 * TC_QT);

            const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
            const unsigned long long TU = P.total_units, U = P.units, G = gridDim.x, g = blockIdx.x;
            const unsigned long long u0 = TU * g / G, u1 = TU * (g + 1) / G;
            constexpr uint32_t CHUNK_BYTES = TC_CHUNK_ELEMS * 2;

            if (tid == 0)
            {
                for (int i = 0; i < TC_MAX_STAGES; ++i)
                {
                    mbar_init(&S->full[i], 1);
                    mbar_init(&S->empty[i], 1);
                }
                mbar_init(&S->a_full, 1);
                mbar_init(&S->a_empty, 1);
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
                    unsigned long long cur = ~0ull;
                    for (unsigned long long u = u0; u < u1; ++u)
                    {
                        const unsigned long long t = u / U, pt = u % U;
                        if (t != cur)
                        {
                            mbar_wait(&S->a_empty, aph ^ 1);  // the MMAs of the previous tile are done with sA
                            mbar_expect_tx(&S->a_full, (uint32_t)nkc * CHUNK_BYTES);
                            tma_bulk_g2s(sA, P.qimg + (size_t)t * nkc * TC_CHUNK_ELEMS, (uint32_t)nkc * CHUNK_BYTES,
                                         &S->a_full);
                            aph ^= 1;
                            cur = t;
                        }
                        const uint16_t *src = P.pimg + (size_t)pt * nkc * TC_CHUNK_ELEMS;
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait(&S->empty[st], ph ^ 1);
                            mbar_expect_tx(&S->full[st], CHUNK_BYTES);
                            tma_bulk_g2s(sB + (size_t)st * CHUNK_BYTES, src + (size_t)kc * TC_CHUNK_ELEMS, CHUNK_BYTES,
                                         &S->full[st]);
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
            else if (warp == 1)
            {
                if (lane == 0)
                {
                    constexpr uint32_t idesc = umma_idesc_bf16(TC_QT, TC_PT);
                    const uint32_t dl = (P.swap_desc & 1) ? P.sbo : P.lbo, ds = (P.swap_desc & 1) ? P.lbo : P.sbo;
                    const uint32_t kstep = 2 * P.lbo;  // one MMA covers K = 16 = two 16-byte column groups
                    const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
                    int st = 0, buf = 0;
                    uint32_t ph = 0, aph = 0, bph = 0;
                    unsigned long long cur = ~0ull;
                    for (unsigned long long u = u0; u < u1; ++u)
                    {
                        const unsigned long long t = u / U;
                        if (t != cur)
                        {
                            if (cur != ~0ull)
                                umma_commit(&S->a_empty);
                            mbar_wait(&S->a_full, aph);
                            aph ^= 1;
                            cur = t;
                        }
                        mbar_wait(&S->t_empty[buf], bph ^ 1);
                        tc_fence_after();
                        const uint32_t dcol = tmem + (uint32_t)buf * TC_PT;
                        for (int kc = 0; kc < nkc; ++kc)
                        {
                            mbar_wait(&S->full[st], ph);
                            tc_fence_after();
#pragma unroll
                            for (int s = 0; s < TC_KC / 16; ++s)
                            {
                                const uint64_t ad = umma_desc(a0 + (uint32_t)kc * CHUNK_BYTES + s * kstep, dl, ds);
                                const uint64_t bd = umma_desc(b0 + (uint32_t)st * CHUNK_BYTES + s * kstep, dl, ds);
                                umma_bf16(dcol, ad, bd, idesc, (kc | s) ? 1u : 0u);
                            }
                            umma_commit(&S->empty[st]);
                            if (++st == nst)
                            {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                        umma_commit(&S->t_full[buf]);
                        if (++buf == TC_ACC)
                        {
                            buf = 0;
                            bph ^= 1;
                        }
                    }
                }
            }
            else
            {
                // selection: this thread's TMEM lane is its query.  The query's list lives in shared
                // memory, LS = L + 4 floats apart (conflict-free LDS.128 across lanes); values carry
                // their list slot in the low mantissa bits (knn_scan.cuh: tag_up), so the running
                // maximum `thr` names the entry to overwrite and ONE max-tree over the list yields the
                // next threshold and the next victim.
                const int wq = warp & 3, qrow = wq * 32 + lane;
                constexpr int LS = L + 4;
                float *lvq = lv + qrow * LS;
                float thr;
                auto reset = [&]() {
#pragma unroll
                    for (int i = 0; i < L; ++i)
                    {
                        lvq[i] = __uint_as_float(EMPTY_BITS | (uint32_t)i);
                        li[i * TC_QT + qrow] = SLOT_NONE;
                    }
                    thr = __uint_as_float(EMPTY_BITS | (uint32_t)(L - 1));
                };
                auto flush = [&](unsigned long long t) {
                    const unsigned long long x = t * U;
                    const unsigned long long gfirst = ((x + 1) * G + TU - 1) / TU - 1;
                    const unsigned seg = (unsigned)(g - gfirst);
                    const size_t q = (size_t)t * TC_QT + qrow;
                    P.cand_thr[(size_t)seg * P.qpad + q] = thr;
#pragma unroll
                    for (int i = 0; i < L; ++i)
                        P.cand_ids[((size_t)seg * L + i) * P.qpad + q] = li[i * TC_QT + qrow];
                };
                auto insert = [&](float v, uint32_t id) {
                    const uint32_t slot = __float_as_uint(thr) & TAG_MASK;
                    lvq[slot] = tag_up(v, slot);
                    li[slot * TC_QT + qrow] = id;
                    float4 a = *reinterpret_cast<const float4 *>(lvq);
                    float m0 = fmaxf(a.x, a.y), m1 = fmaxf(a.z, a.w);
#pragma unroll
                    for (int i = 1; i < L / 4; ++i)
                    {
                        a = *reinterpret_cast<const float4 *>(lvq + 4 * i);
                        m0 = fmaxf(m0, fmaxf(a.x, a.y));
                        m1 = fmaxf(m1, fmaxf(a.z, a.w));
                    }
                    thr = fmaxf(m0, m1);
                };
                int buf = 0;
                uint32_t bph = 0;
                unsigned long long cur = ~0ull;
                for (unsigned long long u = u0; u < u1; ++u)
                {
                    const unsigned long long t = u / U, pt = u % U;
                    if (t != cur)
                    {
                        if (cur != ~0ull)
                            flush(cur);
                        reset();
                        cur = t;
                    }
                    mbar_wait(&S->t_full[buf], bph);
                    tc_fence_after();
                    const uint32_t taddr = tmem + ((uint32_t)(wq * 32) << 16) + (uint32_t)buf * TC_PT;
                    const uint32_t slot0 = (uint32_t)(pt * TC_PT);
#pragma unroll 1
                    for (int c4 = 0; c4 < TC_PT / 32; ++c4)
                    {
                        float v[32];
                        if (P.swap_desc & 2)
                            break;  // diagnostics: no selection at all
                        __syncwarp();
                        tmem_ld32(taddr + c4 * 32, v);
                        if (P.dbg)
                        {
                            float *d = P.dbg + ((size_t)t * TC_QT + qrow) * (U * TC_PT) + slot0 + c4 * 32;
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                d[i] = v[i];
                        }
                        float m = v[0];
#pragma unroll
                        for (int i = 1; i < 32; ++i)
                            m = fminf(m, v[i]);
                        // Rare after the first few thousand points of a segment.  The values go through
                        // shared memory so that the loop can index them; it visits the hits only.
                        if (m < thr && !(P.swap_desc & 4))
                        {
                            uint32_t mask = 0;
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                            {
                                scr[i * TC_QT + qrow] = v[i];
                                mask |= v[i] < thr ? 1u << i : 0u;
                            }
                            while (mask)
                            {
                                const int i = __ffs(mask) - 1;
                                mask &= mask - 1;
                                const float x = scr[i * TC_QT + qrow];
                                if (x < thr)
                                    insert(x, slot0 + c4 * 32 + i);
                            }
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(&S->t_empty[buf]);
                    if (++buf == TC_ACC)
                    {
                        buf = 0;
                        bph ^= 1;
                    }
                }
                if (cur != ~0ull)
                    flush(cur);
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

    size_t tc_smem_bytes(int nkc, int nstages, int L)
    {
        return (size_t)(nkc + nstages) * TC_CHUNK_ELEMS * 2 + (size_t)(2 * L + 4) * TC_QT * 4 + 32 * TC_QT * 4 + sizeof(TcSmem) + 64;
    }

    int tc_pick_stages(int nkc, int L)
    {
        for (int st = 6; st >= 3; --st)
            if (tc_smem_bytes(nkc, st, L) <= 227 * 1024)
                return st;
        return 0;
    }

    void tc_build_point_image(const float *filt, int rowf, int D, size_t units, uint32_t lbo, uint32_t sbo,
                              uint16_t *out, cudaStream_t s)
    {
        const int Kp = tc_kpad(D);
        const size_t rows = units * TC_PT, n = rows * (Kp / 8);
        if (n)
            tc_point_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(filt, rowf, D, Kp, rows, lbo, sbo, out);
    }

    void tc_build_query_image(const float *qp, uint32_t qpad, int D, uint32_t lbo, uint32_t sbo, uint16_t *out,
                              cudaStream_t s)
    {
        const int Kp = tc_kpad(D);
        const size_t n = (size_t)qpad * (Kp / 8);
        if (n)
            tc_query_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(qp, qpad, D, Kp, lbo, sbo, out);
    }

    namespace
    {
        template <int L>
        void tc_launch(const TcScanParams &P, unsigned grid, cudaStream_t s)
        {
            const size_t smem = tc_smem_bytes(P.nkc, P.nstages, L);
            CUDA_CHECK(cudaFuncSetAttribute(knn_scan_tc_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
            knn_scan_tc_kernel<L><<<grid, TC_THREADS, smem, s>>>(P);
        }
    }  // namespace

    void tc_launch_scan(int L, const TcScanParams &P, unsigned grid, cudaStream_t s)
    {
        if (L == 16)
            tc_launch<16>(P, grid, s);
        else if (L == 24)
            tc_launch<24>(P, grid, s);
        else if (L == 48)
            tc_launch<48>(P, grid, s);
        else
            throw Error(OMPLCU_ERR_INTERNAL, "tensor-core scan: unsupported list length");
    }
}  // namespace omplcu
