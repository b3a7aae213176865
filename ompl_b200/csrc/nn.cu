// nn.cu -- device-resident nearest-neighbour index behind omplcu_nn_*.
//
// Replaces NearestNeighborsLinear (datastructures/NearestNeighborsLinear.h:56-173)
// for batched queries; results equal Linear's with ties broken by insertion
// index (total order (distance, id)).
//
//   fast path   plain RealVector space, dim <= 16, k <= 40:
//                 query_prep -> knn_scan (FP32 filter, knn_scan.cuh) -> knn_rerank (exact FP64)
//                 + exact sort path for the (rare) queries whose band check fails
//   exact path  any space / any k: FP64 distance rows + stable segmented radix
//                 sort by (distance, slot); also serves radius queries.
#include "nn.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <cub/cub.cuh>

#include "knn_scan.cuh"
#include "knn_scan_hd.cuh"
#include "knn_scan_tc.cuh"

namespace omplcu
{
    namespace
    {
        constexpr int KNN_MAX_FAST_DIM = 128;  // 1..16: knn_scan.cuh, 17..128: knn_scan_hd.cuh
        constexpr int KNN_MAX_REG_DIM = 16;
        constexpr size_t EXACT_PAIR_BUDGET = (size_t)1 << 25;
        constexpr unsigned RESCAN_MIN = 64;  // unproven queries of a tensor-core pass worth a batched FP32-pipe pass
        constexpr size_t RADIUS_PAIR_BUDGET = (size_t)1 << 27;  // 8 B per pair: 1 GB of distances per batch
        constexpr size_t THR_MIN_SLOTS = 32768;  // below: the list scan (its warm-up is cheaper than a sample pass)

        __device__ __forceinline__ float inff()
        {
            return __int_as_float(0x7f800000);
        }

        // ---- Euclidean lower-bound embedding of SO(2) / SO(3) / compound spaces ----
        // The tensor-core filter works on squared Euclidean distances.  For a space whose components are
        // RealVector, SO(2) and SO(3) (SE(2), SE(3), any weighted compound of them) there is a map E into R^D with
        //        || E(x) - E(y) ||  <=  distance(x, y) + slack            for all states x, y,
        // so "every slot within B of the query" (what the exact paths need, B being their exact distance bound) is a
        // subset of "every slot whose embedded point is within B + slack of the embedded query", which is what
        // filter_scan returns; the exact distance of the kept slots then decides as before.  Per component, with
        // weight w (CompoundStateSpace::distance is the weighted SUM, StateSpace.cpp:1068-1076, and
        // a + b >= sqrt(a^2 + b^2) for non-negative terms):
        //   RealVector   w x                                   the distance itself
        //   SO(2)        w (cos a, sin a)                      chord 2 sin(t/2) <= t = SO2StateSpace::distance
        //                                                      (angles in [-pi, pi], SO2StateSpace.cpp:108-115)
        //   SO(3)        w (q_i^2 / sqrt 2, q_i q_j (i < j))   = vec(q q^T) / sqrt 2: ||.||^2 = 1 - <q1,q2>^2 = sin^2 t
        //                                                      <= t^2, t = acos|<q1,q2>| (SO3StateSpace.cpp:254-262)
        // SO(3) slack: the reference returns 0 once |<q1,q2>| > 1 - 1e-9, where sin t is still up to 4.5e-5, and
        // quaternions whose squared norm is off by eta add 2 eta under the root: 7e-5 per unit of weight covers
        // eta <= 1e-9.  A state outside these assumptions (|angle| > pi, | |q|^2 - 1 | > 1e-9) raises a flag:
        // a point switches the filter off for the structure, a query is sent to the unfiltered path.
        __host__ __device__ inline int embed_dim(const omplcu_component &c)
        {
            return c.type == OMPLCU_RV ? c.dim : c.type == OMPLCU_SO2 ? 2 : c.type == OMPLCU_SO3 ? 10 : -1;
        }
        constexpr double EMBED_SO3_SLACK = 7e-5;
        template <typename PS>
        __host__ __device__ inline bool embed_state(const SpaceDesc &sp, PS st, double *out)
        {
            bool ok = true;
            int off = 0, o = 0;
            for (int i = 0; i < sp.ncomp; ++i)
            {
                const omplcu_component &c = sp.comp[i];
                const double w = sp.compound ? c.weight : 1.0;
                if (c.type == OMPLCU_RV)
                {
                    for (int k = 0; k < c.dim; ++k)
                        out[o++] = w * st[off + k];
                }
                else if (c.type == OMPLCU_SO2)
                {
                    const double a = st[off];
                    ok = ok && fabs(a) <= 3.14159265358979323846 * (1.0 + 1e-15);
                    out[o++] = w * cos(a);
                    out[o++] = w * sin(a);
                }
                else
                {
                    const double q0 = st[off], q1 = st[off + 1], q2 = st[off + 2], q3 = st[off + 3];
                    const double n2 = q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3;
                    ok = ok && fabs(n2 - 1.0) <= 1e-9;
                    const double h = w * 0.70710678118654752440;
                    out[o++] = h * q0 * q0;
                    out[o++] = h * q1 * q1;
                    out[o++] = h * q2 * q2;
                    out[o++] = h * q3 * q3;
                    out[o++] = w * q0 * q1;
                    out[o++] = w * q0 * q2;
                    out[o++] = w * q0 * q3;
                    out[o++] = w * q1 * q2;
                    out[o++] = w * q1 * q3;
                    out[o++] = w * q2 * q3;
                }
                off += c.dim;
            }
            return ok;
        }

        // filt_append_kernel for an embedded space: row = fl32(E(state) - c), norm as there
        __global__ void filt_append_embed_kernel(SpaceDesc sp, const double *pts, const double *center, float *filt,
                                                 unsigned *pmax2, int D, int rowf, size_t first, size_t n, unsigned *bad)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            float pn = 0.f;
            if (i < n)
            {
                double e[KNN_MAX_FAST_DIM];
                if (!embed_state(sp, pts + (first + i) * sp.stride, e))
                    atomicOr(bad, 1u);
                float *f = filt + (first + i) * rowf;
                double acc = 0.0;
                for (int c = 0; c < D; ++c)
                {
                    float v = __double2float_rn(e[c] - center[c]);
                    f[c] = v;
                    acc += (double)v * (double)v;
                }
                pn = __double2float_rn(acc);
                if (!(pn < 3.0e38f))
                    atomicOr(bad, 1u);  // non-finite coordinates: let the exact path deal with them
                f[D] = pn;
                for (int c = D + 1; c < rowf; ++c)
                    f[c] = 0.f;
            }
            unsigned b = __float_as_uint(pn < 3.0e38f ? pn : 0.f);
            for (int o = 16; o > 0; o >>= 1)
                b = max(b, __shfl_xor_sync(0xffffffffu, b, o));
            if ((threadIdx.x & 31) == 0 && b)
                atomicMax(pmax2, b);
        }

        // queries of an exact path -> embedded queries and the bound the filter has to honour
        // (bound_in: exact distance bound per query, or per-query radius)
        __global__ void embed_queries_kernel(SpaceDesc sp, const double *q, size_t nq, int D, double slack,
                                             const double *bound_in, double *out, double *bound_out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nq)
                return;
            double e[KNN_MAX_FAST_DIM];
            const bool ok = embed_state(sp, q + i * sp.stride, e);
            for (int c = 0; c < D; ++c)
                out[i * D + c] = e[c];
            const double b = bound_in[i];
            // not ok / NaN / negative: an infinite bound keeps every slot, the row overflows, the query is redone unfiltered
            bound_out[i] = (ok && b >= 0.0) ? b * (1.0 + 1e-9) + slack + 1e-12 : __longlong_as_double(0x7ff0000000000000ll);
        }

        // the unproven queries of a batch, compacted for the threshold scan
        __global__ void gather_unproven_kernel(const uint32_t *list, size_t n, int D, const double *q_emb, const double *bound,
                                               double slack, double *q_out, double *ebound_out, double *bound_out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            const size_t src = list[i];
            for (int c = 0; c < D; ++c)
                q_out[i * D + c] = q_emb[src * D + c];
            const double b = bound[src];
            bound_out[i] = b;
            // no bound (fewer than kk candidates) -> infinite: the row overflows and the query is redone unfiltered
            ebound_out[i] = (b >= 0.0 && b < 1e300) ? b * (1.0 + 1e-9) + slack + 1e-12 : __longlong_as_double(0x7ff0000000000000ll);
        }

        // ---- filter-copy maintenance ----
        // rows [first, first+n): coords 0, norm +inf  ("dead": never passes the filter)
        __global__ void filt_fill_dead_kernel(float *filt, int rowf, int D, size_t first, size_t n)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n * rowf)
                return;
            int c = (int)(i % rowf);
            filt[first * rowf + i] = (c == D) ? inff() : 0.f;
        }

        // pt = fl32(p - c); pn = fl32(sum pt^2) (sum in FP64 of exact squares)
        __global__ void filt_append_kernel(const double *pts, const double *center, float *filt, unsigned *pmax2,
                                           int D, int rowf, size_t first, size_t n)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            float pn = 0.f;
            if (i < n)
            {
                const double *p = pts + (first + i) * D;
                float *f = filt + (first + i) * rowf;
                double acc = 0.0;
                for (int c = 0; c < D; ++c)
                {
                    float v = __double2float_rn(p[c] - center[c]);
                    f[c] = v;
                    acc += (double)v * (double)v;
                }
                pn = __double2float_rn(acc);
                f[D] = pn;
                for (int c = D + 1; c < rowf; ++c)
                    f[c] = 0.f;
            }
            // block max of pn (non-negative floats order like their bit patterns)
            unsigned b = __float_as_uint(pn);
            for (int o = 16; o > 0; o >>= 1)
                b = max(b, __shfl_xor_sync(0xffffffffu, b, o));
            if ((threadIdx.x & 31) == 0 && b)
                atomicMax(pmax2, b);
        }

        __global__ void mark_dead_kernel(uint8_t *alive, float *filt, int rowf, int D, size_t slot)
        {
            alive[slot] = 0;
            if (filt)
                filt[slot * rowf + D] = inff();
        }

        __global__ void mark_dead_batch_kernel(uint8_t *alive, float *filt, int rowf, int D, const uint32_t *slots,
                                               size_t n)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= n)
                return;
            alive[slots[i]] = 0;
            if (filt)
                filt[(size_t)slots[i] * rowf + D] = inff();
        }

        __global__ void fill_ids_kernel(uint64_t *ids, uint8_t *alive, size_t first, size_t n, uint64_t id0)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i < n)
            {
                ids[first + i] = id0 + i;
                alive[first + i] = 1;
            }
        }

        __global__ void gather_rows_kernel(const double *src, double *dst, const uint64_t *src_ids, uint64_t *dst_ids,
                                           const uint32_t *live_slots, size_t nlive, int stride)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nlive * stride)
                return;
            size_t r = i / stride;
            int c = (int)(i % stride);
            dst[i] = src[(size_t)live_slots[r] * stride + c];
            if (c == 0)
                dst_ids[r] = src_ids[live_slots[r]];
        }

        // ---- filter scan for a handful of queries (the literal nearestK of PRM::addMilestone, prm/src/PRM.cpp:575) ----
        // A tensor-core tile has 128 query rows; with 1 - 8 queries nearly all of them are padding and the call is
        // bound by the MMA -> selection hand-over per point tile (66 us at 1M states).  Here a lane owns a point: the
        // warp reads 32 consecutive rows of the FP32 filter copy (coalesced), every lane evaluates the scan value of
        // its point against each query exactly as knn_scan_kernel does (s = pn, then one FMA per coordinate in
        // order: same value, same (d + 3) u error bound), and the warp keeps, per query, the FEW_L smallest values
        // of the rows it has seen: list entry i lives in lane i's registers, the minimum over the batch and the
        // list maximum come from redux.sync on an order-preserving integer image of the floats.  Every warp is one
        // segment of the candidate arrays (cand_thr = its list maximum once full), which knn_rerank_select_kernel
        // re-ranks exactly and proves like any other scan's.  HBM/L2-bound: one pass over 4 (d + 1) bytes per point.
        constexpr int FEW_MAXQ = 8, FEW_L = 4, FEW_WARPS = 4, FEW_UB = 4;
        __device__ __forceinline__ uint32_t few_key(float v)
        {
            // monotone float -> unsigned; +inf (dead rows), NaN and the padding sort last and are never kept
            const uint32_t b = __float_as_uint(v);
            return v < 3.0e38f ? ((b & 0x80000000u) ? ~b : (b | 0x80000000u)) : 0xFFFFFFFFu;
        }
        __device__ __forceinline__ float few_unkey(uint32_t k)
        {
            return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
        }
        template <int NQ>
        __global__ void __launch_bounds__(FEW_WARPS * 32) knn_few_kernel(const float *filt, int rowf, int D, size_t nslots,
                                                                        const float *qp, uint32_t qpad, float *cand_thr,
                                                                        uint32_t *cand_ids)
        {
            extern __shared__ float few_q[];  // [D][NQ]
            for (int i = threadIdx.x; i < D * NQ; i += blockDim.x)
                few_q[i] = qp[(size_t)(i / NQ) * qpad + (i % NQ)];
            __syncthreads();
            const unsigned full = 0xffffffffu;
            const int lane = threadIdx.x & 31;
            const size_t gw = blockIdx.x * (size_t)FEW_WARPS + (threadIdx.x >> 5), GW = gridDim.x * (size_t)FEW_WARPS;
            const size_t nbatch = (nslots + 31) / 32;
            uint32_t lk[NQ], lid[NQ], thr[NQ];  // list entry `lane` (lanes < FEW_L) and the list maximum, per query
#pragma unroll
            for (int j = 0; j < NQ; ++j)
            {
                lk[j] = lane < FEW_L ? 0xFFFFFFFFu : 0u;
                lid[j] = SLOT_NONE;
                thr[j] = 0xFFFFFFFFu;
            }
            const int nc4 = (D + 3) / 4;
            for (size_t b0 = gw; b0 < nbatch; b0 += GW * FEW_UB)
            {
                float acc[FEW_UB][NQ];
                uint32_t row[FEW_UB];
#pragma unroll
                for (int u = 0; u < FEW_UB; ++u)
                {
                    const size_t b = b0 + (size_t)u * GW, r = b * 32 + lane;
                    const bool live = b < nbatch && r < nslots;
                    row[u] = live ? (uint32_t)r : SLOT_NONE;
                    const float pn = live ? filt[r * rowf + D] : __uint_as_float(0x7f800000u);
#pragma unroll
                    for (int j = 0; j < NQ; ++j)
                        acc[u][j] = pn;
                }
                for (int c4 = 0; c4 < nc4; ++c4)
                {
                    float4 x[FEW_UB];
#pragma unroll
                    for (int u = 0; u < FEW_UB; ++u)
                        x[u] = row[u] != SLOT_NONE
                                   ? *reinterpret_cast<const float4 *>(filt + (size_t)row[u] * rowf + 4 * c4)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
                    const int c = 4 * c4;
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (c + e < D)
                        {
#pragma unroll
                            for (int j = 0; j < NQ; ++j)
                            {
                                const float qv = few_q[(c + e) * NQ + j];
#pragma unroll
                                for (int u = 0; u < FEW_UB; ++u)
                                    acc[u][j] = fmaf(qv, e == 0 ? x[u].x : e == 1 ? x[u].y : e == 2 ? x[u].z : x[u].w, acc[u][j]);
                            }
                        }
                }
#pragma unroll
                for (int u = 0; u < FEW_UB; ++u)
#pragma unroll
                    for (int j = 0; j < NQ; ++j)
                    {
                        uint32_t key = few_key(acc[u][j]);
                        uint32_t m = __reduce_min_sync(full, key);
                        while (m < thr[j])
                        {
                            const int src = __ffs(__ballot_sync(full, key == m)) - 1;   // lowest row first
                            const uint32_t id = __shfl_sync(full, row[u], src);
                            const int dst = __ffs(__ballot_sync(full, lane < FEW_L && lk[j] == thr[j])) - 1;
                            if (lane == dst)
                            {
                                lk[j] = m;
                                lid[j] = id;
                            }
                            if (lane == src)
                                key = 0xFFFFFFFFu;
                            thr[j] = __reduce_max_sync(full, lk[j]);
                            m = __reduce_min_sync(full, key);
                        }
                    }
            }
            if (lane < FEW_L)
            {
#pragma unroll
                for (int j = 0; j < NQ; ++j)
                    cand_ids[(gw * FEW_L + lane) * qpad + j] = lid[j];
            }
            if (lane == 0)
            {
#pragma unroll
                for (int j = 0; j < NQ; ++j)
                    cand_thr[gw * qpad + j] = thr[j] != 0xFFFFFFFFu ? few_unkey(thr[j]) : __uint_as_float(0x7F7F7F7Fu);
            }
        }

        // ---- Morton order for large query batches (knn_fast) ----
        // The selection's slow path runs whenever ANY of a warp's 32 queries sees a value below its threshold.  When
        // the lanes hold nearby queries the same points are near all of them and the hits coincide: 5 % off the scan
        // at 1M states, 16 % at 125K (profiles/r2_experiments.md).  Two bits per dimension (the first 16 dimensions)
        // are all that helps; the sort is stable, so the order within a cell is the caller's.
        constexpr int QSORT_DIMS = 16;
        __global__ void qsort_bbox_kernel(const double *q, size_t nq, int D, int nd, uint32_t *bbox)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            for (int c = 0; c < nd; ++c)
            {
                const uint32_t key = i < nq ? few_key((float)q[i * D + c]) : 0xFFFFFFFFu;
                const uint32_t lo = __reduce_min_sync(0xffffffffu, key);
                const uint32_t hi = __reduce_max_sync(0xffffffffu, key == 0xFFFFFFFFu ? 0u : key);
                if ((threadIdx.x & 31) == 0)
                {
                    atomicMin(&bbox[c], lo);
                    atomicMax(&bbox[QSORT_DIMS + c], hi);
                }
            }
        }
        __global__ void qsort_key_kernel(const double *q, size_t nq, int D, int nd, const uint32_t *bbox, uint32_t *keys,
                                         uint32_t *idx)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nq)
                return;
            uint32_t cell[QSORT_DIMS];
            for (int c = 0; c < nd; ++c)
            {
                const float lo = few_unkey(bbox[c]), hi = few_unkey(bbox[QSORT_DIMS + c]);
                const float t = ((float)q[i * D + c] - lo) / fmaxf(hi - lo, 1e-30f) * 4.0f;
                cell[c] = t >= 0.f ? min((uint32_t)t, 3u) : 0u;  // NaN -> 0
            }
            uint32_t key = 0;
            for (int b = 1; b >= 0; --b)
                for (int c = 0; c < nd; ++c)
                    key = (key << 1) | ((cell[c] >> b) & 1u);
            keys[i] = key;
            idx[i] = (uint32_t)i;
        }

        // ---- second pass over the queries the tensor-core filter could not prove (knn_fast) ----
        __global__ void gather_rows_kernel(const double *q, const uint32_t *sel, size_t nsel, int D, double *out)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i < nsel * (size_t)D)
                out[i] = q[(size_t)sel[i / D] * D + i % D];
        }
        __global__ void scatter_results_kernel(const uint64_t *ids, const double *dists, const uint32_t *sel, size_t nsel,
                                               uint32_t kk, uint32_t kstride, uint64_t *out_ids, double *out_dists)
        {
            const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nsel * (size_t)kk)
                return;
            const size_t o = (size_t)sel[i / kk] * kstride + i % kk;
            out_ids[o] = ids[i];
            if (out_dists)
                out_dists[o] = dists[i];
        }

        // ---- query preparation for the scan ----
        // qp[c][q] = -2 * fl32(q_c - center_c); qaux[q] = (|qt|^2, |qt|) in FP64
        // perm (optional): position i of the scan holds query perm[i] (knn_fast sorts large batches along a Morton
        // curve so that the lanes of a selection warp hold nearby queries)
        __global__ void query_prep_kernel(const double *q, size_t nq, const double *center, int D, float *qp,
                                          double *qaux, uint32_t qpad, unsigned *zero = nullptr,
                                          const uint32_t *perm = nullptr)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i == 0 && zero)
                *zero = 0;  // the re-rank's fallback counter (saves a memset launch per call)
            if (i >= qpad)
                return;
            const size_t src = (perm && i < nq) ? perm[i] : i;
            double qn = 0.0;
            for (int c = 0; c < D; ++c)
            {
                float v = 0.f;
                if (i < nq)
                    v = __double2float_rn(q[src * D + c] - center[c]);
                qp[(size_t)c * qpad + i] = -2.f * v;
                qn += (double)v * (double)v;
            }
            qaux[2 * i + 0] = qn;
            qaux[2 * i + 1] = sqrt(qn);
        }

        // ---- exact re-rank of the scan's candidate lists (one warp per query) ----
        // Exact distance = RealVectorStateSpace::distance (RealVectorStateSpace.cpp:230-242)
        // with argument order (element, query) as in NearestNeighborsLinear.h:159-172.
        // Selection = the kk smallest by (distance, slot), written in that order.
        // Band check: let Dk be the kk-th selected exact distance.  Any true
        // member p* of the kk best has dist(p*) <= Dk, hence ("filter
        // bound") its scan value s(p*) <= tau(Dk).  A point the scan dropped from
        // a full list has s >= that list's maximum vL; if vL > tau for every full
        // list, no dropped point can be p*.  Otherwise the query is flagged and
        // redone exactly.
        template <int MAXC>
        __global__ void __launch_bounds__(128) knn_rerank_kernel(
            const float *cand_thr, const uint32_t *cand_ids, uint32_t smax, uint32_t L, uint32_t qpad,
            const double *pts, const uint64_t *slot_ids, int D, const double *q, const double *qaux,
            const unsigned *pmax2_bits, size_t nq, uint32_t kk, uint32_t kstride, uint64_t *out_ids,
            double *out_dists, uint32_t *ovf_list, unsigned *ovf_count, double err_units,
            const uint32_t *perm = nullptr)
        {
            __shared__ double s_key[4][MAXC];
            __shared__ uint32_t s_slot[4][MAXC];
            __shared__ double s_dk[4];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t qi = blockIdx.x * 4ull + w;  // position in the scan's order
            if (qi >= nq)
                return;
            const size_t qo = perm ? perm[qi] : qi;   // the caller's query index
            const uint32_t n = smax * L;
            const double *qrow = q + qo * D;
            for (uint32_t c = lane; c < n; c += 32)
            {
                const uint32_t slot = cand_ids[(size_t)c * qpad + qi];
                double key = __longlong_as_double(0x7ff0000000000000ll);
                if (slot != SLOT_NONE)
                    key = rv_distance(pts + (size_t)slot * D, qrow, D);
                s_key[w][c] = key;
                s_slot[w][c] = slot;
            }
            if (lane == 0)
                s_dk[w] = -1.0;  // stays negative if the lists hold fewer than kk candidates
            __syncwarp();
            for (uint32_t c = lane; c < n; c += 32)
            {
                const uint32_t slot = s_slot[w][c];
                if (slot == SLOT_NONE)
                    continue;
                const double key = s_key[w][c];
                uint32_t rank = 0;
                for (uint32_t o = 0; o < n; ++o)
                {
                    const double ko = s_key[w][o];
                    const uint32_t so = s_slot[w][o];
                    rank += (ko < key || (ko == key && so < slot)) ? 1u : 0u;
                }
                if (rank < kk)
                {
                    out_ids[qo * kstride + rank] = slot_ids[slot];
                    if (out_dists)
                        out_dists[qo * kstride + rank] = key;
                    if (rank == kk - 1)
                        s_dk[w] = key;
                }
            }
            __syncwarp();
            // ---- band check ----
            const double u = 5.9604644775390625e-08;  // 2^-24
            const double dk = s_dk[w];
            const double qn = qaux[2 * qi], qnorm = qaux[2 * qi + 1];
            const double pmax2 = (double)__uint_as_float(*pmax2_bits) * (1.0 + 2.0 * u);
            const double pmax = sqrt(pmax2) * (1.0 + 1e-9);
            const double delta = (double)(D + 4) * 1.1102230246251565e-16;
            const double rho = 1.0000002 * u * (qnorm + pmax) + 1e-44;
            const double A = dk * (1.0 + delta) + rho + 1e-150;
            const double e1 = err_units * u * 1.000001 * (pmax2 + 2.0 * qnorm * pmax) + 1e-43;
            const double tau64 = A * A - qn + e1 + (double)D * 2.3e-16 * qn + 2e-15 * (A * A + qn);
            bool bad = !(tau64 == tau64) || !(tau64 < 1e37) || dk < 0.0;  // NaN / overflow / too few -> redo exactly
            for (uint32_t s = lane; s < smax; s += 32)
            {
                // final threshold of the segment's list; an EMPTY marker (> 1e38) means
                // the list never filled, i.e. nothing was dropped from that segment.
                // Stored values are rounded up by < 128 ulp (knn_scan.cuh): step back 2^-16.
                const double vmax = (double)cand_thr[(size_t)s * qpad + qi];
                if (vmax < 1e38 && !(vmax - fabs(vmax) * 1.52587890625e-05 - 1e-37 > tau64))
                    bad = true;
            }
            bad = __any_sync(0xffffffffu, bad);
            if (bad && lane == 0)
            {
                unsigned pos = atomicAdd(ovf_count, 1u);
                ovf_list[pos] = (uint32_t)qo;
            }
        }

        // ---- threshold path (knn_thr_scan_tc_kernel): bound selection and exact epilogue ----
        // floats ordered as unsigned integers
        __device__ __forceinline__ uint32_t f32_key(float x)
        {
            const uint32_t b = __float_as_uint(x);
            return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
        }

        // tau[q] = the r-th smallest (1-based) of the nvals list values the sample scan left for query q
        // (vals[c][q], c < nvals; EMPTY markers > 1e38 where a list did not fill), found by bisection on
        // the bits.  Fewer than r real values: no bound (3e38, everything is kept).  LPQ lanes serve one
        // query: 1 when a query has a few dozen values, 32 when a lone query was cut into many segments.
        template <int LPQ>
        __global__ void __launch_bounds__(128) tau_select_kernel(const float *vals, uint32_t nvals, uint32_t qpad,
                                                                 uint32_t r, float *tau)
        {
            const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
            const uint32_t qi = t / LPQ, l = t % LPQ;
            if (qi >= qpad)
                return;
            uint32_t lo = 0, hi = 0xFFFFFFFFu;
            while (lo < hi)
            {
                const uint32_t mid = lo + ((hi - lo) >> 1);
                uint32_t c = 0;
                for (uint32_t j = l; j < nvals; j += LPQ)
                    c += f32_key(vals[(size_t)j * qpad + qi]) <= mid ? 1u : 0u;
                if (LPQ > 1)
                    c = __reduce_add_sync(0xffffffffu, c);
                if (c >= r)
                    hi = mid;
                else
                    lo = mid + 1;
            }
            if (l == 0)
            {
                const uint32_t b = (lo & 0x80000000u) ? (lo & 0x7FFFFFFFu) : ~lo;
                const float v = __uint_as_float(b);
                tau[qi] = (v < 1e38f) ? v : 3.0e38f;  // (NaN never enters a list)
            }
        }

        // Exact epilogue of the threshold scan, one warp per query.  Every kept slot gets its exact
        // distance (RealVectorStateSpace::distance, argument order (element, query) as in
        // NearestNeighborsLinear.h:159-172); the kk smallest by (distance, slot) are written in that order.
        // Proof obligation (the band check of knn_rerank_kernel with one list whose final threshold is
        // tau): a point within the kk-th selected distance Dk has a scan value <= bound(Dk); the scan
        // kept everything below tau; so tau > bound(Dk) proves the answer.  Otherwise -- or when a row
        // overflowed, or fewer than kk slots were kept -- the query goes to the exact path.
        // Selection: a 256-bucket histogram of the distances (monotone map) finds the bucket of the kk-th
        // smallest, the slots up to that bucket are compacted in place and ranked among themselves.
        template <int MAXC>
        __global__ void __launch_bounds__(128) knn_rerank_thr_kernel(
            const uint32_t *app_ids, const uint32_t *app_cnt, uint32_t smax, uint32_t cap, uint32_t qpad,
            const float *tau, const double *pts, const uint64_t *slot_ids, int D, const double *q,
            const double *qaux, const unsigned *pmax2_bits, size_t nq, uint32_t kk, uint32_t kstride,
            uint64_t *out_ids, double *out_dists, uint32_t *ovf_list, unsigned *ovf_count, double err_units,
            unsigned long long *kept_total)
        {
            extern __shared__ __align__(16) unsigned char rr_smem[];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            double *s_key = reinterpret_cast<double *>(rr_smem) + (size_t)w * MAXC;
            uint32_t *s_slot = reinterpret_cast<uint32_t *>(reinterpret_cast<double *>(rr_smem) + 4 * MAXC) + (size_t)w * MAXC;
            uint32_t *s_hist = reinterpret_cast<uint32_t *>(reinterpret_cast<double *>(rr_smem) + 4 * MAXC) + 4 * MAXC + w * 256;
            const size_t qi = blockIdx.x * 4ull + w;
            if (qi >= nq)
                return;
            const double *qrow = q + qi * D;
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            bool bad = false;
            uint32_t n = 0;
            double kmin = inf, kmax = 0.0;
            for (uint32_t s = 0; s < smax; ++s)
            {
                const uint32_t c = app_cnt[(size_t)s * qpad + qi];
                if (c > cap || n + c > MAXC)
                {
                    bad = true;  // the row overflowed, or more slots than the epilogue holds
                    break;
                }
                const uint32_t *row = app_ids + ((size_t)s * qpad + qi) * cap;
                for (uint32_t i = lane; i < c; i += 32)
                {
                    const uint32_t slot = row[i];
                    const double key = rv_distance(pts + (size_t)slot * D, qrow, D);
                    s_key[n + i] = key;
                    s_slot[n + i] = slot;
                    kmin = fmin(kmin, key);  // fmin / fmax drop a NaN: a NaN distance fails the band check below
                    kmax = fmax(kmax, key);
                }
                n += c;
            }
            if (lane == 0 && kept_total)
                atomicAdd(kept_total, (unsigned long long)n);
            double dk = -1.0;
            if (!bad && n >= kk)
            {
                for (int o = 16; o > 0; o >>= 1)
                {
                    kmin = fmin(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
                    kmax = fmax(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
                }
                for (int i = lane; i < 256; i += 32)
                    s_hist[i] = 0;
                __syncwarp();
                const double scale = (kmax > kmin) ? 255.0 / (kmax - kmin) : 0.0;
                auto bucket = [&](double key) -> uint32_t {
                    const double b = (key - kmin) * scale;  // monotone in key; NaN -> bucket 255
                    return (b >= 0.0 && b < 255.0) ? (uint32_t)b : 255u;
                };
                for (uint32_t i = lane; i < n; i += 32)
                    atomicAdd(&s_hist[bucket(s_key[i])], 1u);
                __syncwarp();
                // bucket of the kk-th smallest: lane l owns buckets 8 l .. 8 l + 7
                uint32_t mine = 0;
                for (int i = 0; i < 8; ++i)
                    mine += s_hist[8 * lane + i];
                uint32_t incl = mine;
                for (int o = 1; o < 32; o <<= 1)
                {
                    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o)
                        incl += v;
                }
                const uint32_t first = __ffs(__ballot_sync(0xffffffffu, incl >= kk)) - 1;  // n >= kk: some lane reaches it
                uint32_t bsel = 0;
                if ((uint32_t)lane == first)
                {
                    uint32_t acc = incl - mine;
                    for (int i = 0; i < 8; ++i)
                    {
                        acc += s_hist[8 * lane + i];
                        if (acc >= kk)
                        {
                            bsel = 8 * lane + i;
                            break;
                        }
                    }
                }
                bsel = __shfl_sync(0xffffffffu, bsel, first);
                // compact the slots of buckets <= bsel in place (a write never passes the chunk being read)
                uint32_t m = 0;
                for (uint32_t i0 = 0; i0 < n; i0 += 32)
                {
                    const uint32_t i = i0 + lane;
                    double key = 0.0;
                    uint32_t slot = 0;
                    bool keep = false;
                    if (i < n)
                    {
                        key = s_key[i];
                        slot = s_slot[i];
                        keep = bucket(key) <= bsel;
                    }
                    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                    __syncwarp();
                    if (keep)
                    {
                        const uint32_t pos = m + __popc(bal & ((1u << lane) - 1u));
                        s_key[pos] = key;
                        s_slot[pos] = slot;
                    }
                    m += __popc(bal);
                    __syncwarp();
                }
                // rank by (distance, slot) among the m compacted entries (m >= kk)
                for (uint32_t c = lane; c < m; c += 32)
                {
                    const double key = s_key[c];
                    const uint32_t slot = s_slot[c];
                    uint32_t rank = 0;
                    for (uint32_t o = 0; o < m; ++o)
                    {
                        const double ko = s_key[o];
                        const uint32_t so = s_slot[o];
                        rank += (ko < key || (ko == key && so < slot)) ? 1u : 0u;
                    }
                    if (rank < kk)
                    {
                        out_ids[qi * kstride + rank] = slot_ids[slot];
                        if (out_dists)
                            out_dists[qi * kstride + rank] = key;
                        if (rank == kk - 1)
                            dk = key;
                    }
                }
                for (int o = 16; o > 0; o >>= 1)
                    dk = fmax(dk, __shfl_xor_sync(0xffffffffu, dk, o));
            }
            // ---- band check (same bound as knn_rerank_kernel) ----
            const double u = 5.9604644775390625e-08;  // 2^-24
            const double qn = qaux[2 * qi], qnorm = qaux[2 * qi + 1];
            const double pmax2 = (double)__uint_as_float(*pmax2_bits) * (1.0 + 2.0 * u);
            const double pmax = sqrt(pmax2) * (1.0 + 1e-9);
            const double delta = (double)(D + 4) * 1.1102230246251565e-16;
            const double rho = 1.0000002 * u * (qnorm + pmax) + 1e-44;
            const double A = dk * (1.0 + delta) + rho + 1e-150;
            const double e1 = err_units * u * 1.000001 * (pmax2 + 2.0 * qnorm * pmax) + 1e-43;
            const double tau64 = A * A - qn + e1 + (double)D * 2.3e-16 * qn + 2e-15 * (A * A + qn);
            bad = bad || !(tau64 == tau64) || !(tau64 < 1e37) || dk < 0.0;  // NaN / overflow / too few -> redo exactly
            const double tq = (double)tau[qi];
            if (tq < 1e38 && !(tq > tau64))
                bad = true;  // tau >= 1e38: nothing was dropped
            if (bad && lane == 0)
            {
                const unsigned pos = atomicAdd(ovf_count, 1u);
                ovf_list[pos] = (uint32_t)qi;
            }
        }

        // ---- filter in front of the exact large-k / radius paths ----
        // tau[q] = an upper bound on the scan value of ANY point within distance bound[q] of query q (the band
        // formula of knn_rerank_kernel with Dk = bound), rounded up, one ulp added: the threshold scan keeps
        // every slot with s < tau, hence every point within the bound.  Padded queries keep nothing.
        __global__ void scan_bound_kernel(const double *bound, size_t nq, uint32_t qpad, const double *qaux,
                                          const unsigned *pmax2_bits, int D, double err_units, float *tau)
        {
            const size_t qi = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (qi >= qpad)
                return;
            if (qi >= nq)
            {
                tau[qi] = -3.0e38f;
                return;
            }
            const double u = 5.9604644775390625e-08;  // 2^-24
            const double dk = bound[qi];
            const double qn = qaux[2 * qi], qnorm = qaux[2 * qi + 1];
            const double pmax2 = (double)__uint_as_float(*pmax2_bits) * (1.0 + 2.0 * u);
            const double pmax = sqrt(pmax2) * (1.0 + 1e-9);
            const double delta = (double)(D + 4) * 1.1102230246251565e-16;
            const double rho = 1.0000002 * u * (qnorm + pmax) + 1e-44;
            const double A = dk * (1.0 + delta) + rho + 1e-150;
            const double e1 = err_units * u * 1.000001 * (pmax2 + 2.0 * qnorm * pmax) + 1e-43;
            const double tau64 = A * A - qn + e1 + (double)D * 2.3e-16 * qn + 2e-15 * (A * A + qn);
            float t = 3.0e38f;  // no usable bound (NaN, infinite radius, overflow): keep everything
            if (tau64 == tau64 && tau64 < 1e37 && dk >= 0.0)
            {
                t = __double2float_ru(tau64);
                t = __uint_as_float(t >= 0.f ? __float_as_uint(t) + 1u : (__float_as_uint(t) == 0x80000000u ? 1u : __float_as_uint(t) - 1u));
            }
            tau[qi] = t;
        }

        // One warp per query: the slots the threshold scan kept (rows in segment order, each in scan order,
        // i.e. increasing slot) get their exact distance; those within bound[b] are written to the query's
        // segment [b * cap2, ...) IN SLOT ORDER (ballot compaction), so one stable sort by distance finishes
        // the (distance, slot) order.  A row that overflowed, or more than cap2 survivors, marks the query
        // (cnt = cap2 + 1): the caller redoes it exactly.
        __global__ void __launch_bounds__(128) kept_collect_kernel(const uint32_t *app_ids, const uint32_t *app_cnt,
                                                                   uint32_t smax, uint32_t cap, uint32_t qpad,
                                                                   const double *pts, const uint8_t *alive, int D,
                                                                   const double *q, const double *bound, size_t nb,
                                                                   uint32_t cap2, double *keys, uint32_t *vals,
                                                                   uint32_t *cnt)
        {
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const double *qrow = q + b * D;
            const double bd = bound[b];
            uint32_t n = 0;
            bool over = false;
            for (uint32_t s = 0; s < smax && !over; ++s)
            {
                const uint32_t c = app_cnt[(size_t)s * qpad + b];
                if (c > cap)
                {
                    over = true;
                    break;
                }
                const uint32_t *row = app_ids + ((size_t)s * qpad + b) * cap;
                for (uint32_t i0 = 0; i0 < c; i0 += 32)
                {
                    const uint32_t i = i0 + lane;
                    bool keep = false;
                    uint32_t slot = 0;
                    double d = 0.0;
                    if (i < c)
                    {
                        slot = row[i];
                        if (alive[slot])
                        {
                            d = rv_distance(pts + (size_t)slot * D, qrow, D);
                            keep = d <= bd;
                        }
                    }
                    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                    const uint32_t pos = n + __popc(bal & ((1u << lane) - 1u));
                    if (keep && pos < cap2)
                    {
                        keys[b * cap2 + pos] = d;
                        vals[b * cap2 + pos] = slot;
                    }
                    n += __popc(bal);
                }
            }
            if (lane == 0)
                cnt[b] = (over || n > cap2) ? cap2 + 1u : n;
        }

        // fixed-capacity per-query segments [b * cap2, b * cap2 + cnt[b]) -> contiguous at offs[b], order kept
        __global__ void __launch_bounds__(128) segment_pack_kernel(const double *keys, const uint32_t *vals,
                                                                   const uint32_t *cnt, uint32_t cap2,
                                                                   const unsigned long long *offs, size_t nb,
                                                                   double *out_keys, uint32_t *out_vals)
        {
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const uint32_t n = cnt[b];
            const unsigned long long o = offs[b];
            for (uint32_t i = lane; i < n; i += 32)
            {
                out_keys[o + i] = keys[b * cap2 + i];
                out_vals[o + i] = vals[b * cap2 + i];
            }
        }

        // Small result lists (every query of the batch has at most MAXN survivors): pack, order by
        // (distance, slot) and emit in one kernel, one warp per query -- no segmented radix sort over
        // tens of thousands of tiny segments.
        template <int MAXN>
        __global__ void __launch_bounds__(128) radius_sort_emit_kernel(const double *keys, const uint32_t *vals,
                                                                       const uint32_t *cnt, uint32_t cap2,
                                                                       const unsigned long long *offs, size_t nb,
                                                                       const uint64_t *slot_ids, uint64_t *out_ids,
                                                                       double *out_dists)
        {
            __shared__ double s_key[4][MAXN];
            __shared__ uint32_t s_slot[4][MAXN];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const uint32_t n = cnt[b];
            const unsigned long long o = offs[b];
            for (uint32_t i = lane; i < n; i += 32)
            {
                s_key[w][i] = keys[b * cap2 + i];
                s_slot[w][i] = vals[b * cap2 + i];
            }
            __syncwarp();
            for (uint32_t i = lane; i < n; i += 32)
            {
                const double key = s_key[w][i];
                const uint32_t slot = s_slot[w][i];
                uint32_t rank = 0;
                for (uint32_t j = 0; j < n; ++j)
                {
                    const double kj = s_key[w][j];
                    rank += (kj < key || (kj == key && s_slot[w][j] < slot)) ? 1u : 0u;
                }
                out_ids[o + rank] = slot_ids[slot];
                out_dists[o + rank] = key;
            }
        }

        // Re-rank for small batches whose scan was cut into many segments (up to one per SM, so that a lone
        // query streams the point image on all of them): one warp per query, up to a few thousand candidates in
        // dynamic shared memory.  Same exact distances, same (distance, slot) order and the same band check as
        // knn_rerank_kernel; the kk best are extracted one by one (every lane caches the best of its own
        // candidates, a shuffle tree picks the winner, only the winning lane rescans) instead of ranking all pairs.
        constexpr int SEL_THREADS = 1024;

        // RealVectorStateSpace::distance with the loads of eight coordinates issued before their (sequential)
        // accumulation: the same operations in the same order as rv_distance, hence the same bits, but the
        // memory latency of a scattered row is paid once per eight coordinates instead of once per coordinate.
        __device__ __forceinline__ double rv_distance_mlp(const double *a, const double *b, int d)
        {
            double dist = 0.0;
            int i = 0;
            for (; i + 8 <= d; i += 8)
            {
                double v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    v[j] = a[i + j];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                {
                    const double diff = dsub(v[j], b[i + j]);
                    dist = dadd(dist, dmul(diff, diff));
                }
            }
            double v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                v[j] = (i + j < d) ? a[i + j] : 0.0;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (i + j < d)
                {
                    const double diff = dsub(v[j], b[i + j]);
                    dist = dadd(dist, dmul(diff, diff));
                }
            return dsqrt(dist);
        }

        __global__ void __launch_bounds__(SEL_THREADS) knn_rerank_select_kernel(
            const float *cand_thr, const uint32_t *cand_ids, uint32_t smax, uint32_t L, uint32_t qpad,
            const double *pts, const uint64_t *slot_ids, int D, const double *q, const double *qaux,
            const unsigned *pmax2_bits, size_t nq, uint32_t kk, uint32_t kstride, uint64_t *out_ids,
            double *out_dists, uint32_t *ovf_list, unsigned *ovf_count, double err_units,
            const uint32_t *perm = nullptr)
        {
            extern __shared__ __align__(16) unsigned char sel_smem[];
            const int lane = threadIdx.x & 31;
            const size_t qi = blockIdx.x;  // position in the scan's order
            if (qi >= nq)
                return;
            const size_t qo = perm ? perm[qi] : qi;  // the caller's query index
            const uint32_t n = smax * L;
            double *s_key = reinterpret_cast<double *>(sel_smem);
            double *s_q = s_key + n;
            uint32_t *s_slot = reinterpret_cast<uint32_t *>(s_q + D);
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            for (int i = threadIdx.x; i < D; i += SEL_THREADS)
                s_q[i] = q[qo * D + i];
            __syncthreads();
            // exact distances: the whole block (the candidates' rows are scattered loads); the selection below
            // is one warp's work
            for (uint32_t c = threadIdx.x; c < n; c += SEL_THREADS)
            {
                const uint32_t slot = cand_ids[(size_t)c * qpad + qi];
                s_key[c] = slot != SLOT_NONE ? rv_distance_mlp(pts + (size_t)slot * D, s_q, D) : inf;
                s_slot[c] = slot;
            }
            __syncthreads();
            // Selection in two levels: every warp extracts the kk best of its own 1/32 of the candidates (2 - 3 per
            // lane), then warp 0 extracts the kk best of those 32 kk survivors.  The kk best overall are among the
            // survivors, and both levels order by (distance, slot), so the result is the one-level result; a round
            // costs a lane-local scan of a few entries and a shuffle tree instead of a scan of n / 32 entries.
            double *s_key2 = reinterpret_cast<double *>(s_slot + n + (n & 1u));
            uint32_t *s_slot2 = reinterpret_cast<uint32_t *>(s_key2 + (size_t)(SEL_THREADS / 32) * kk);
            const uint32_t warp = threadIdx.x >> 5;
            double bk = inf;
            uint32_t bs = SLOT_NONE, bi = 0;
            // best live entry among this lane's share of [lo, hi) of (key, slot)
            auto rescan = [&](const double *key, const uint32_t *slot, uint32_t lo, uint32_t hi) {
                bk = inf;
                bs = SLOT_NONE;
                bi = 0;
                for (uint32_t c = lo + lane; c < hi; c += 32)
                {
                    const uint32_t sl = slot[c];
                    if (sl == SLOT_NONE)
                        continue;
                    const double k = key[c];
                    // a NaN distance never wins a comparison: it stays behind and the band check below fails on it
                    if (bs == SLOT_NONE || k < bk || (k == bk && sl < bs))
                    {
                        bk = k;
                        bs = sl;
                        bi = c;
                    }
                }
            };
            // one extraction round over the lanes' cached bests; returns false when nothing is left
            auto pick = [&](double &wk, uint32_t &ws, uint32_t &wi) {
                wk = bk;
                ws = bs;
                wi = bi;
                for (int o = 16; o > 0; o >>= 1)
                {
                    const double ok = __shfl_xor_sync(0xffffffffu, wk, o);
                    const uint32_t os = __shfl_xor_sync(0xffffffffu, ws, o);
                    const uint32_t oi = __shfl_xor_sync(0xffffffffu, wi, o);
                    if (os != SLOT_NONE && (ws == SLOT_NONE || ok < wk || (ok == wk && os < ws)))
                    {
                        wk = ok;
                        ws = os;
                        wi = oi;
                    }
                }
                return ws != SLOT_NONE;
            };
            {
                const uint32_t per = (n + SEL_THREADS / 32 - 1) / (SEL_THREADS / 32);
                const uint32_t lo = min(n, warp * per), hi = min(n, lo + per);
                rescan(s_key, s_slot, lo, hi);
                for (uint32_t r = 0; r < kk; ++r)
                {
                    double wk;
                    uint32_t ws, wi;
                    const bool any = pick(wk, ws, wi);
                    if (lane == 0)
                    {
                        s_key2[warp * kk + r] = any ? wk : inf;
                        s_slot2[warp * kk + r] = any ? ws : SLOT_NONE;
                    }
                    if (any && ((wi - lo) & 31u) == (uint32_t)lane)
                    {
                        s_slot[wi] = SLOT_NONE;
                        rescan(s_key, s_slot, lo, hi);
                    }
                    __syncwarp();
                }
            }
            __syncthreads();
            if (threadIdx.x >= 32)
                return;
            const uint32_t n2 = (SEL_THREADS / 32) * kk;
            rescan(s_key2, s_slot2, 0, n2);
            double dk = -1.0;
            for (uint32_t r = 0; r < kk; ++r)
            {
                double wk;
                uint32_t ws, wi;
                if (!pick(wk, ws, wi))
                    break;  // fewer than kk candidates: dk stays negative, the query is redone exactly
                if (lane == 0)
                {
                    out_ids[qo * kstride + r] = slot_ids[ws];
                    if (out_dists)
                        out_dists[qo * kstride + r] = wk;
                }
                if (r == kk - 1)
                    dk = wk;
                if ((wi & 31u) == (uint32_t)lane)
                {
                    s_slot2[wi] = SLOT_NONE;
                    rescan(s_key2, s_slot2, 0, n2);
                }
                __syncwarp();
            }
            // ---- band check (knn_rerank_kernel) ----
            const double u = 5.9604644775390625e-08;  // 2^-24
            const double qn = qaux[2 * qi], qnorm = qaux[2 * qi + 1];
            const double pmax2 = (double)__uint_as_float(*pmax2_bits) * (1.0 + 2.0 * u);
            const double pmax = sqrt(pmax2) * (1.0 + 1e-9);
            const double delta = (double)(D + 4) * 1.1102230246251565e-16;
            const double rho = 1.0000002 * u * (qnorm + pmax) + 1e-44;
            const double A = dk * (1.0 + delta) + rho + 1e-150;
            const double e1 = err_units * u * 1.000001 * (pmax2 + 2.0 * qnorm * pmax) + 1e-43;
            const double tau64 = A * A - qn + e1 + (double)D * 2.3e-16 * qn + 2e-15 * (A * A + qn);
            bool bad = !(tau64 == tau64) || !(tau64 < 1e37) || dk < 0.0 || !(dk == dk);
            for (uint32_t s = lane; s < smax; s += 32)
            {
                const double vmax = (double)cand_thr[(size_t)s * qpad + qi];
                if (vmax < 1e38 && !(vmax - fabs(vmax) * 1.52587890625e-05 - 1e-37 > tau64))
                    bad = true;
            }
            bad = __any_sync(0xffffffffu, bad);
            if (bad && lane == 0)
            {
                const unsigned pos = atomicAdd(ovf_count, 1u);
                ovf_list[pos] = (uint32_t)qo;
            }
        }

        // compact copy of every step-th filter row (the sample thr_sample_kernel reads)
        __global__ void sample_gather_kernel(const float4 *filt, int rw, uint32_t step, uint32_t nsample, float4 *out)
        {
            const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
            if (i < nsample * (uint32_t)rw)
                out[i] = filt[(size_t)(i / rw) * step * rw + (i % rw)];
        }

        // Starting bound per query for the scan lists.  One warp serves SQ queries (held in registers); lane l
        // evaluates the scan value (same FMA order as knn_scan_kernel, so bit-identical) of
        // sample rows l, l+32, ... and keeps two running minima over alternating rows:
        // 64 minima of 64 disjoint groups, i.e. of 64 distinct points.  T = the L-th
        // smallest of them, so at least L points have a scan value <= T.
        template <int D, int SQ>
        __global__ void __launch_bounds__(128) thr_sample_kernel(const float4 *sample, uint32_t nsample,
                                                                 const float *qp, uint32_t qpad, uint32_t L,
                                                                 float *out)
        {
            constexpr int RW = scan_rowf(D) / 4;
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const uint32_t q0 = (blockIdx.x * 4u + w) * SQ;
            if (q0 >= qpad)
                return;
            float qv[D][SQ];
#pragma unroll
            for (int c = 0; c < D; ++c)
#pragma unroll
                for (int j = 0; j < SQ; ++j)
                    qv[c][j] = qp[(size_t)c * qpad + q0 + j];
            const float inf = __int_as_float(0x7f800000);
            float m0[SQ], m1[SQ];
#pragma unroll
            for (int j = 0; j < SQ; ++j)
                m0[j] = m1[j] = inf;
            for (uint32_t i = lane; i < nsample; i += 64)
            {
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    const uint32_t r = i + 32u * h;
                    float x[4 * RW];
#pragma unroll
                    for (int v = 0; v < RW; ++v)
                    {
                        const float4 f = r < nsample ? sample[(size_t)r * RW + v] : make_float4(inf, inf, inf, inf);
                        x[4 * v + 0] = f.x;
                        x[4 * v + 1] = f.y;
                        x[4 * v + 2] = f.z;
                        x[4 * v + 3] = f.w;
                    }
                    // row layout: D coordinates, then the squared norm (+inf for dead rows)
                    float sv[SQ];
#pragma unroll
                    for (int j = 0; j < SQ; ++j)
                        sv[j] = x[D];
                    if (r < nsample)
                    {
#pragma unroll
                        for (int c = 0; c < D; ++c)
#pragma unroll
                            for (int j = 0; j < SQ; ++j)
                                sv[j] = fmaf(qv[c][j], x[c], sv[j]);
                    }
#pragma unroll
                    for (int j = 0; j < SQ; ++j)
                        if (h)
                            m1[j] = fminf(m1[j], sv[j]);
                        else
                            m0[j] = fminf(m0[j], sv[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < SQ; ++j)
            {
                // rank of each of the 64 minima (ties by position); the one of rank L-1 is T
                uint32_t r0 = 0, r1 = 0;
                for (int o = 0; o < 32; ++o)
                {
                    const float a0 = __shfl_sync(0xffffffffu, m0[j], o);
                    const float a1 = __shfl_sync(0xffffffffu, m1[j], o);
                    r0 += (a0 < m0[j] || (a0 == m0[j] && o < lane)) ? 1u : 0u;
                    r0 += (a1 < m0[j]) ? 1u : 0u;
                    r1 += (a0 < m1[j] || a0 == m1[j]) ? 1u : 0u;
                    r1 += (a1 < m1[j] || (a1 == m1[j] && o < lane)) ? 1u : 0u;
                }
                float T = inf;
                if (r0 == L - 1)
                    T = m0[j];
                if (r1 == L - 1)
                    T = m1[j];
                T = fminf(T, __shfl_xor_sync(0xffffffffu, T, 16));
                T = fminf(T, __shfl_xor_sync(0xffffffffu, T, 8));
                T = fminf(T, __shfl_xor_sync(0xffffffffu, T, 4));
                T = fminf(T, __shfl_xor_sync(0xffffffffu, T, 2));
                T = fminf(T, __shfl_xor_sync(0xffffffffu, T, 1));
                if (lane == 0)
                    out[q0 + j] = T;  // +inf (no bound) if fewer than L live sample points
            }
        }

        // ---- exact path ----
        // keys[b][slot] = exact distance(point slot, query sel[b]) (+inf for dead slots)
        // Distance between slot `slot` and query `qi` for the exact-path kernels.  For the kinematic
        // chain metric (demos/KinematicChain.h:105-121) the cos / sin of the cumulative joint angles
        // depend on one state only, so they are tabulated per state (chain_trig_kernel) and the
        // per-pair work is the reference's running sums on tabulated values: same operations on the
        // same operands in the same order, hence the same bits, without 4 * links libm calls per pair.
        struct PairMetric
        {
            SpaceDesc sp;
            const double *pts, *q;
            const double *ptrig, *qtrig;  // [slot][2*links] and [query][2*links], or nullptr
            __device__ __forceinline__ double operator()(size_t slot, size_t qi) const
            {
                if (ptrig == nullptr)
                    return space_distance(sp, pts + slot * sp.stride, q + qi * sp.stride);
                const int d = sp.comp[0].dim;
                const double *a = ptrig + slot * 2 * d, *b = qtrig + qi * 2 * d;
                double dx = 0., dy = 0., dist = 0.;
                for (int i = 0; i < d; ++i)
                {
                    dx = dadd(dx, dsub(a[2 * i], b[2 * i]));
                    dy = dadd(dy, dsub(a[2 * i + 1], b[2 * i + 1]));
                    dist = dadd(dist, dsqrt(dadd(dmul(dx, dx), dmul(dy, dy))));
                }
                return dmul(dist, sp.comp[0].param);
            }
        };

        // kept_collect_kernel for any metric: exact distance of the slots the (embedded) filter kept, survivors within
        // the query's exact bound written in slot order.  Query b of this batch is query q0 + b of pm.
        __global__ void __launch_bounds__(128) kept_collect_metric_kernel(const uint32_t *app_ids, const uint32_t *app_cnt,
                                                                          uint32_t smax, uint32_t cap, uint32_t qpad,
                                                                          PairMetric pm, size_t q0, const uint8_t *alive,
                                                                          const double *bound, size_t nb, uint32_t cap2,
                                                                          double *keys, uint32_t *vals, uint32_t *cnt,
                                                                          const uint32_t *sel = nullptr)
        {
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const size_t qsrc = q0 + (sel ? sel[b] : b);  // row b of this scan is query qsrc of pm
            const double bd = bound[b];
            uint32_t n = 0;
            bool over = false;
            for (uint32_t s = 0; s < smax && !over; ++s)
            {
                const uint32_t c = app_cnt[(size_t)s * qpad + b];
                if (c > cap)
                {
                    over = true;
                    break;
                }
                const uint32_t *row = app_ids + ((size_t)s * qpad + b) * cap;
                for (uint32_t i0 = 0; i0 < c; i0 += 32)
                {
                    const uint32_t i = i0 + lane;
                    bool keep = false;
                    uint32_t slot = 0;
                    double d = 0.0;
                    if (i < c)
                    {
                        slot = row[i];
                        if (alive[slot])
                        {
                            d = pm(slot, qsrc);
                            keep = d <= bd;
                        }
                    }
                    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                    const uint32_t pos = n + __popc(bal & ((1u << lane) - 1u));
                    if (keep && pos < cap2)
                    {
                        keys[b * cap2 + pos] = d;
                        vals[b * cap2 + pos] = slot;
                    }
                    n += __popc(bal);
                }
            }
            if (lane == 0)
                cnt[b] = (over || n > cap2) ? cap2 + 1u : n;
        }

        // Embedded spaces, bound for the filter: the kk-th smallest EXACT distance among the (<= 512) slots a list scan
        // on the embedding found for the query.  Any kk real points bound the kk-th nearest distance from above, and
        // the embedding's nearest are nearly the metric's nearest, so the bound is tight (a few times kk slots lie
        // within it, against thousands for a bound from a 1024-point sample).  Fewer than kk candidates: +inf.
        __global__ void __launch_bounds__(128) cand_bound_kernel(const uint32_t *cand_ids, const float *cand_thr, uint32_t smax,
                                                                 uint32_t L, uint32_t qpad, PairMetric pm, size_t q0,
                                                                 const uint8_t *alive, size_t nb, uint32_t kk, double *bound,
                                                                 const double *qaux, const unsigned *pmax2_bits, int D,
                                                                 double err_units, double slack, const uint64_t *slot_ids,
                                                                 uint32_t kstride, uint64_t *out_ids, double *out_dists,
                                                                 unsigned *unproven, uint32_t *unproven_list,
                                                                 uint32_t step, size_t nslots)
        {
            __shared__ double s_key[4][512];
            __shared__ uint32_t s_slot[4][512];
            __shared__ double s_wd[4][40];
            __shared__ uint32_t s_ws[4][40];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const uint32_t n = smax * L;
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            for (uint32_t c = lane; c < n; c += 32)
            {
                // step > 1: the scan ran on the image of every step-th state; its rows are slots / step
                const uint32_t row = cand_ids[(size_t)c * qpad + b];
                const size_t slot = row != SLOT_NONE ? (size_t)row * step : (size_t)SLOT_NONE;
                double d = inf;
                if (row != SLOT_NONE && slot < nslots && alive[slot])
                {
                    d = pm(slot, q0 + b);
                    if (!(d == d))
                        d = inf;
                }
                s_key[w][c] = d;
                s_slot[w][c] = d < inf ? (uint32_t)slot : SLOT_NONE;
            }
            __syncwarp();
            // the kk best by (distance, slot): every lane caches the best of its own candidates, a shuffle tree picks
            // the winner of a round, only the winning lane rescans (as knn_rerank_select_kernel)
            double bk = inf;
            uint32_t bs = SLOT_NONE, bi = 0;
            auto rescan = [&]() {
                bk = inf;
                bs = SLOT_NONE;
                bi = 0;
                for (uint32_t c = lane; c < n; c += 32)
                {
                    const uint32_t sl = s_slot[w][c];
                    if (sl == SLOT_NONE)
                        continue;
                    const double k = s_key[w][c];
                    if (bs == SLOT_NONE || k < bk || (k == bk && sl < bs))
                    {
                        bk = k;
                        bs = sl;
                        bi = c;
                    }
                }
            };
            rescan();
            double dk = inf;
            for (uint32_t r = 0; r < kk; ++r)
            {
                double wk = bk;
                uint32_t ws = bs, wi = bi;
                for (int o = 16; o > 0; o >>= 1)
                {
                    const double ok = __shfl_xor_sync(0xffffffffu, wk, o);
                    const uint32_t os = __shfl_xor_sync(0xffffffffu, ws, o);
                    const uint32_t oi = __shfl_xor_sync(0xffffffffu, wi, o);
                    if (os != SLOT_NONE && (ws == SLOT_NONE || ok < wk || (ok == wk && os < ws)))
                    {
                        wk = ok;
                        ws = os;
                        wi = oi;
                    }
                }
                if (ws == SLOT_NONE)
                    break;  // fewer than kk candidates: no bound
                if (lane == 0)
                {
                    s_wd[w][r] = wk;
                    s_ws[w][r] = ws;
                }
                if (r == kk - 1)
                    dk = wk;
                if ((wi & 31u) == (uint32_t)lane)
                {
                    s_slot[w][wi] = SLOT_NONE;
                    rescan();
                }
                __syncwarp();
            }
            // Can anything the lists dropped be within dk?  A slot within dk in the space's metric is within
            // dk' = dk (1 + 1e-9) + slack of the query in the embedding (embed_state), and the band formula of
            // knn_rerank_kernel bounds the scan value of such a slot by tau(dk'); every full list's maximum must lie above it.
            bool proven = dk < inf && step == 1;  // a sample's lists say nothing about the states outside the sample
            if (proven)
            {
                const double u = 5.9604644775390625e-08;  // 2^-24
                const double de = dk * (1.0 + 1e-9) + slack + 1e-12;
                const double qn = qaux[2 * b], qnorm = qaux[2 * b + 1];
                const double pmax2 = (double)__uint_as_float(*pmax2_bits) * (1.0 + 2.0 * u);
                const double pmax = sqrt(pmax2) * (1.0 + 1e-9);
                const double delta = (double)(D + 4) * 1.1102230246251565e-16;
                const double rho = 1.0000002 * u * (qnorm + pmax) + 1e-44;
                const double A = de * (1.0 + delta) + rho + 1e-150;
                const double e1 = err_units * u * 1.000001 * (pmax2 + 2.0 * qnorm * pmax) + 1e-43;
                const double tau64 = A * A - qn + e1 + (double)D * 2.3e-16 * qn + 2e-15 * (A * A + qn);
                bool bad = !(tau64 == tau64) || !(tau64 < 1e37);
                for (uint32_t sg = lane; sg < smax; sg += 32)
                {
                    const double vmax = (double)cand_thr[(size_t)sg * qpad + b];
                    if (vmax < 1e38 && !(vmax - fabs(vmax) * 1.52587890625e-05 - 1e-37 > tau64))
                        bad = true;
                }
                proven = !__any_sync(0xffffffffu, bad);
            }
            if (proven)
            {
                // the listed slots contain every slot within dk: the answer is their kk best by (distance, slot)
                for (uint32_t r = lane; r < kk; r += 32)
                {
                    out_ids[(q0 + b) * kstride + r] = slot_ids[s_ws[w][r]];
                    if (out_dists)
                        out_dists[(q0 + b) * kstride + r] = s_wd[w][r];
                }
            }
            if (lane == 0)
            {
                bound[b] = dk;
                if (!proven)
                    unproven_list[atomicAdd(unproven, 1u)] = (uint32_t)b;
            }
        }

        // ---- a handful of queries on ANY space (RRT's nn_->nearest on SE(3), rrt/src/RRT.cpp:170): exact brute force ----
        // One pass, a lane per slot: every lane evaluates the space's exact FP64 distance of its slot to each query and
        // the warp keeps, per query, the kk smallest (distance, slot) pairs it has seen (entry i in lane i).  A warp
        // walks its slots in increasing order, so among equal distances the earlier slot is kept: the union of the
        // warps' lists contains the exact answer, which exact_few_select_kernel extracts.  No filter, no proof.
        constexpr int XFEW_MAXQ = 8, XFEW_MAXK = 32, XFEW_WARPS = 4;
        __device__ __forceinline__ bool xfew_less(double da, uint32_t sa, double db, uint32_t sb)
        {
            return da < db || (da == db && sa < sb);
        }
        template <int NQ>
        __global__ void __launch_bounds__(XFEW_WARPS * 32) exact_few_kernel(PairMetric pm, const uint8_t *alive, size_t nslots,
                                                                           uint32_t nq, uint32_t kk, uint32_t qpad,
                                                                           double *cand_d, uint32_t *cand_slot)
        {
            const unsigned full = 0xffffffffu;
            const int lane = threadIdx.x & 31;
            const size_t gw = blockIdx.x * (size_t)XFEW_WARPS + (threadIdx.x >> 5), GW = gridDim.x * (size_t)XFEW_WARPS;
            const size_t nbatch = (nslots + 31) / 32;
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            double ld[NQ], maxd[NQ];       // list entry `lane` (lanes < kk) and the list maximum, per query
            uint32_t ls[NQ], maxs[NQ];
#pragma unroll
            for (int j = 0; j < NQ; ++j)
            {
                ld[j] = maxd[j] = inf;
                ls[j] = maxs[j] = SLOT_NONE;
            }
            for (size_t b = gw; b < nbatch; b += GW)
            {
                const size_t slot = b * 32 + lane;
                const bool live = slot < nslots && alive[slot];
#pragma unroll
                for (int j = 0; j < NQ; ++j)
                {
                    if (j >= (int)nq)
                        break;
                    double d = live ? pm(slot, (size_t)j) : inf;
                    if (!(d == d))
                        d = inf;  // a NaN distance never enters a list (NearestNeighborsLinear's comparisons are false too)
                    uint32_t sl = live ? (uint32_t)slot : SLOT_NONE;
                    while (true)
                    {
                        // smallest (distance, slot) still on offer in this batch
                        double md = d;
                        uint32_t ms = sl;
                        for (int o = 16; o > 0; o >>= 1)
                        {
                            const double od = __shfl_xor_sync(full, md, o);
                            const uint32_t os = __shfl_xor_sync(full, ms, o);
                            if (xfew_less(od, os, md, ms))
                            {
                                md = od;
                                ms = os;
                            }
                        }
                        if (!(md < inf) || !xfew_less(md, ms, maxd[j], maxs[j]))
                            break;
                        // replace the list's maximum
                        const int dst = __ffs(__ballot_sync(full, lane < (int)kk && ld[j] == maxd[j] && ls[j] == maxs[j])) - 1;
                        if (lane == dst)
                        {
                            ld[j] = md;
                            ls[j] = ms;
                        }
                        if (sl == ms)
                        {
                            d = inf;
                            sl = SLOT_NONE;
                        }
                        double xd = lane < (int)kk ? ld[j] : -1.0;
                        uint32_t xs = lane < (int)kk ? ls[j] : 0u;
                        for (int o = 16; o > 0; o >>= 1)
                        {
                            const double od = __shfl_xor_sync(full, xd, o);
                            const uint32_t os = __shfl_xor_sync(full, xs, o);
                            if (xfew_less(xd, xs, od, os))
                            {
                                xd = od;
                                xs = os;
                            }
                        }
                        maxd[j] = xd;
                        maxs[j] = xs;
                    }
                }
            }
            if (lane < (int)kk)
            {
#pragma unroll
                for (int j = 0; j < NQ; ++j)
                {
                    cand_d[(gw * kk + lane) * qpad + j] = ld[j];
                    cand_slot[(gw * kk + lane) * qpad + j] = ls[j];
                }
            }
        }

        // the kk smallest (distance, slot) of n candidates per query, two levels as in knn_rerank_select_kernel
        __global__ void __launch_bounds__(SEL_THREADS) exact_few_select_kernel(const double *cand_d, const uint32_t *cand_slot,
                                                                              uint32_t n, uint32_t qpad,
                                                                              const uint64_t *slot_ids, uint32_t kk,
                                                                              uint32_t kstride, uint64_t *out_ids,
                                                                              double *out_dists)
        {
            extern __shared__ __align__(16) unsigned char sel_smem[];
            const int lane = threadIdx.x & 31;
            const uint32_t warp = threadIdx.x >> 5;
            const size_t qi = blockIdx.x;
            double *s_key = reinterpret_cast<double *>(sel_smem);
            double *s_key2 = s_key + n;
            uint32_t *s_slot = reinterpret_cast<uint32_t *>(s_key2 + (size_t)(SEL_THREADS / 32) * kk);
            uint32_t *s_slot2 = s_slot + n;
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            for (uint32_t c = threadIdx.x; c < n; c += SEL_THREADS)
            {
                s_key[c] = cand_d[(size_t)c * qpad + qi];
                s_slot[c] = cand_slot[(size_t)c * qpad + qi];
            }
            __syncthreads();
            double bk = inf;
            uint32_t bs = SLOT_NONE, bi = 0;
            auto rescan = [&](const double *key, const uint32_t *slot, uint32_t lo, uint32_t hi) {
                bk = inf;
                bs = SLOT_NONE;
                bi = 0;
                for (uint32_t c = lo + lane; c < hi; c += 32)
                {
                    const uint32_t sl = slot[c];
                    if (sl == SLOT_NONE)
                        continue;
                    const double k = key[c];
                    if (bs == SLOT_NONE || k < bk || (k == bk && sl < bs))
                    {
                        bk = k;
                        bs = sl;
                        bi = c;
                    }
                }
            };
            auto pick = [&](double &wk, uint32_t &ws, uint32_t &wi) {
                wk = bk;
                ws = bs;
                wi = bi;
                for (int o = 16; o > 0; o >>= 1)
                {
                    const double ok = __shfl_xor_sync(0xffffffffu, wk, o);
                    const uint32_t os = __shfl_xor_sync(0xffffffffu, ws, o);
                    const uint32_t oi = __shfl_xor_sync(0xffffffffu, wi, o);
                    if (os != SLOT_NONE && (ws == SLOT_NONE || ok < wk || (ok == wk && os < ws)))
                    {
                        wk = ok;
                        ws = os;
                        wi = oi;
                    }
                }
                return ws != SLOT_NONE;
            };
            {
                const uint32_t per = (n + SEL_THREADS / 32 - 1) / (SEL_THREADS / 32);
                const uint32_t lo = min(n, warp * per), hi = min(n, lo + per);
                rescan(s_key, s_slot, lo, hi);
                for (uint32_t r = 0; r < kk; ++r)
                {
                    double wk;
                    uint32_t ws, wi;
                    const bool any = pick(wk, ws, wi);
                    if (lane == 0)
                    {
                        s_key2[warp * kk + r] = any ? wk : inf;
                        s_slot2[warp * kk + r] = any ? ws : SLOT_NONE;
                    }
                    if (any && ((wi - lo) & 31u) == (uint32_t)lane)
                    {
                        s_slot[wi] = SLOT_NONE;
                        rescan(s_key, s_slot, lo, hi);
                    }
                    __syncwarp();
                }
            }
            __syncthreads();
            if (threadIdx.x >= 32)
                return;
            const uint32_t n2 = (SEL_THREADS / 32) * kk;
            rescan(s_key2, s_slot2, 0, n2);
            for (uint32_t r = 0; r < kk; ++r)
            {
                double wk;
                uint32_t ws, wi;
                if (!pick(wk, ws, wi))
                    break;  // fewer than kk live elements: the caller pads the row
                if (lane == 0)
                {
                    out_ids[qi * kstride + r] = slot_ids[ws];
                    if (out_dists)
                        out_dists[qi * kstride + r] = wk;
                }
                if ((wi & 31u) == (uint32_t)lane)
                {
                    s_slot2[wi] = SLOT_NONE;
                    rescan(s_key2, s_slot2, 0, n2);
                }
                __syncwarp();
            }
        }

        // out[row][2i] = cos(theta_i), out[row][2i+1] = sin(theta_i), theta_i = s_0 + ... + s_i
        __global__ void chain_trig_kernel(const double *rows, size_t n, int d, double *out)
        {
            size_t r = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (r >= n)
                return;
            double theta = 0.;
            for (int i = 0; i < d; ++i)
            {
                theta = dadd(theta, rows[r * d + i]);
                out[r * 2 * d + 2 * i] = omplcu_cos(theta);
                out[r * 2 * d + 2 * i + 1] = omplcu_sin(theta);
            }
        }

        __global__ void dist_rows_kernel(PairMetric pm, const uint8_t *alive, size_t nslots, const uint32_t *sel,
                                         size_t b0, size_t nb, double *keys, uint32_t *vals)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * nslots)
                return;
            size_t b = i / nslots, slot = i % nslots;
            size_t qi = sel ? sel[b0 + b] : (b0 + b);
            double d = __longlong_as_double(0x7ff0000000000000ll);
            if (alive[slot])
                d = pm(slot, qi);
            keys[i] = d;
            vals[i] = (uint32_t)slot;
        }

        __global__ void exact_emit_kernel(const double *keys, const uint32_t *vals, size_t nslots,
                                          const uint64_t *slot_ids, const uint32_t *sel, size_t b0, size_t nb,
                                          uint32_t kk, uint32_t kstride, uint64_t *out_ids, double *out_dists)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * kk)
                return;
            size_t b = i / kk, r = i % kk;
            size_t qi = sel ? sel[b0 + b] : (b0 + b);
            out_ids[qi * kstride + r] = slot_ids[vals[b * nslots + r]];
            if (out_dists)
                out_dists[qi * kstride + r] = keys[b * nslots + r];
        }

        // ---- exact path, large k: distance bound from a sample, collect, sort only what passed ----
        // sd[b][j] = exact distance from query b to every step-th slot (dead slots: +inf)
        __global__ void sample_dist_kernel(PairMetric pm, const uint8_t *alive, size_t step, uint32_t m,
                                           const uint32_t *sel, size_t b0, size_t nb, double *sd)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * m)
                return;
            const size_t b = i / m, slot = (i % m) * step;
            const size_t qi = sel ? sel[b0 + b] : (b0 + b);
            double d = __longlong_as_double(0x7ff0000000000000ll);
            if (alive[slot])
                d = pm(slot, qi);
            sd[i] = d;
        }

        // bound[b] = the R-th smallest (1-based) of the m sample distances of query b: at least R sample
        // points lie within it.  One warp per query; distances are >= 0 (or +inf), so their bit patterns
        // order like unsigned integers and the value is found by bisection on the bits.
        __global__ void __launch_bounds__(128) rank_select_kernel(const double *sd, uint32_t m, size_t nb, uint32_t R,
                                                                  double *bound)
        {
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const unsigned long long *v = reinterpret_cast<const unsigned long long *>(sd) + b * m;
            unsigned long long lo = 0, hi = 0x7ff0000000000000ull;  // answer in [lo, hi]
            while (lo < hi)
            {
                const unsigned long long mid = lo + ((hi - lo) >> 1);
                unsigned c = 0;
                for (uint32_t j = lane; j < m; j += 32)
                    c += v[j] <= mid ? 1u : 0u;
                c = __reduce_add_sync(0xffffffffu, c);
                if (c >= R)
                    hi = mid;
                else
                    lo = mid + 1;
            }
            if (lane == 0)
                bound[b] = __longlong_as_double((long long)lo);
        }

        // every live slot within bound[b] of query b is appended to the query's segment [b*cap, b*cap+cap)
        __global__ void collect_kernel(PairMetric pm, const uint8_t *alive, size_t nslots, const uint32_t *sel,
                                       size_t b0, size_t nb, const double *bound, uint32_t cap, double *keys,
                                       uint32_t *vals, uint32_t *cnt)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * nslots)
                return;
            const size_t b = i / nslots, slot = i % nslots;
            if (!alive[slot])
                return;
            const size_t qi = sel ? sel[b0 + b] : (b0 + b);
            const double d = pm(slot, qi);
            if (d <= bound[b])
            {
                const uint32_t pos = atomicAdd(&cnt[b], 1u);
                if (pos < cap)
                {
                    keys[b * cap + pos] = d;
                    vals[b * cap + pos] = (uint32_t)slot;
                }
            }
        }

        __global__ void collect_segments_kernel(const uint32_t *cnt, uint32_t cap, size_t nb, int *beg, int *end)
        {
            size_t b = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (b >= nb)
                return;
            beg[b] = (int)(b * cap);
            end[b] = (int)(b * cap + min(cnt[b], cap));
        }

        // first kk of every sorted segment; segments that overflowed or came up short are queued for
        // the full sort (redo holds query indices)
        // Kept lists of the embedded path (a few dozen to a few hundred survivors per query, at most `cap`): one warp per
        // query extracts the kk best by (distance, slot) from shared memory -- every lane caches the best of its own
        // entries, a shuffle tree picks the winner of a round, only the winning lane rescans -- instead of a segmented
        // radix sort over 65536 tiny segments (which was half of the SE(3) call).  More than `cap` survivors, or fewer
        // than kk: the query is redone unfiltered.  Dynamic shared memory: 4 warps x cap x 12 bytes.
        __global__ void __launch_bounds__(128) collect_rank_emit_kernel(const double *keys, const uint32_t *vals,
                                                                        const uint32_t *cnt, uint32_t cap,
                                                                        const uint64_t *slot_ids, size_t b0, size_t nb,
                                                                        uint32_t kk, uint32_t kstride, uint64_t *out_ids,
                                                                        double *out_dists, uint32_t *redo,
                                                                        unsigned *redo_count, const uint32_t *sel = nullptr)
        {
            extern __shared__ __align__(16) unsigned char cre_smem[];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            double *s_key = reinterpret_cast<double *>(cre_smem) + (size_t)w * cap;
            uint32_t *s_slot = reinterpret_cast<uint32_t *>(reinterpret_cast<double *>(cre_smem) + (size_t)4 * cap) + (size_t)w * cap;
            const size_t b = blockIdx.x * 4ull + w;
            if (b >= nb)
                return;
            const size_t qi = b0 + (sel ? sel[b] : b);
            const uint32_t n = cnt[b];
            if (n > cap || n < kk)
            {
                if (lane == 0)
                    redo[atomicAdd(redo_count, 1u)] = (uint32_t)qi;
                return;
            }
            for (uint32_t i = lane; i < n; i += 32)
            {
                s_key[i] = keys[b * cap + i];
                s_slot[i] = vals[b * cap + i];
            }
            __syncwarp();
            const double inf = __longlong_as_double(0x7ff0000000000000ll);
            double bk = inf;
            uint32_t bs = SLOT_NONE, bi = 0;
            auto rescan = [&]() {
                bk = inf;
                bs = SLOT_NONE;
                bi = 0;
                for (uint32_t c = lane; c < n; c += 32)
                {
                    const uint32_t sl = s_slot[c];
                    if (sl == SLOT_NONE)
                        continue;
                    const double k = s_key[c];
                    if (bs == SLOT_NONE || k < bk || (k == bk && sl < bs))
                    {
                        bk = k;
                        bs = sl;
                        bi = c;
                    }
                }
            };
            rescan();
            for (uint32_t r = 0; r < kk; ++r)
            {
                double wk = bk;
                uint32_t ws = bs, wi = bi;
                for (int o = 16; o > 0; o >>= 1)
                {
                    const double ok = __shfl_xor_sync(0xffffffffu, wk, o);
                    const uint32_t os = __shfl_xor_sync(0xffffffffu, ws, o);
                    const uint32_t oi = __shfl_xor_sync(0xffffffffu, wi, o);
                    if (os != SLOT_NONE && (ws == SLOT_NONE || ok < wk || (ok == wk && os < ws)))
                    {
                        wk = ok;
                        ws = os;
                        wi = oi;
                    }
                }
                if (ws == SLOT_NONE)
                    break;  // cannot happen (n >= kk live entries); the row would keep its previous content
                if (lane == 0)
                {
                    out_ids[qi * kstride + r] = slot_ids[ws];
                    if (out_dists)
                        out_dists[qi * kstride + r] = wk;
                }
                if ((wi & 31u) == (uint32_t)lane)
                {
                    s_slot[wi] = SLOT_NONE;
                    rescan();
                }
                __syncwarp();
            }
        }

        __global__ void collect_emit_kernel(const double *keys, const uint32_t *vals, const uint32_t *cnt,
                                            uint32_t cap, const uint64_t *slot_ids, const uint32_t *sel, size_t b0,
                                            size_t nb, uint32_t kk, uint32_t kstride, uint64_t *out_ids,
                                            double *out_dists, uint32_t *redo, unsigned *redo_count)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * kk)
                return;
            const size_t b = i / kk, r = i % kk;
            const size_t qi = sel ? sel[b0 + b] : (b0 + b);
            const uint32_t c = cnt[b];
            if (c > cap || c < kk)
            {
                if (r == 0)
                    redo[atomicAdd(redo_count, 1u)] = (uint32_t)qi;
                return;
            }
            out_ids[qi * kstride + r] = slot_ids[vals[b * cap + r]];
            if (out_dists)
                out_dists[qi * kstride + r] = keys[b * cap + r];
        }

        // pad columns [kk, kstride) with OMPLCU_ID_NONE / +inf
        __global__ void pad_rows_kernel(uint64_t *ids, double *dists, size_t nq, uint32_t kk, uint32_t kstride)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            uint32_t w = kstride - kk;
            if (i >= nq * w)
                return;
            size_t qi = i / w, r = kk + i % w;
            ids[qi * kstride + r] = OMPLCU_ID_NONE;
            if (dists)
                dists[qi * kstride + r] = __longlong_as_double(0x7ff0000000000000ll);
        }

        // radius: exact distance of every (query, slot) pair; pairs within the radius are counted per
        // query as they are written (dead slots carry +inf and never count)
        __global__ void radius_dist_kernel(PairMetric pm, const uint8_t *alive, size_t nslots, size_t b0, size_t nb,
                                           const double *radii, double *keys, uint32_t *counts)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * nslots)
                return;
            const size_t b = i / nslots, slot = i % nslots;
            double d = __longlong_as_double(0x7ff0000000000000ll);
            if (alive[slot])
                d = pm(slot, b0 + b);
            keys[i] = d;
            if (d <= radii[b0 + b] && d < __longlong_as_double(0x7ff0000000000000ll))
                atomicAdd(&counts[b], 1u);
        }

        // the counted pairs, copied into the query's segment (any order: the sort by slot follows)
        __global__ void radius_compact_kernel(const double *keys, size_t nslots, size_t b0, size_t nb,
                                              const double *radii, const unsigned long long *offs, uint32_t *fill,
                                              double *out_keys, uint32_t *out_vals)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= nb * nslots)
                return;
            const size_t b = i / nslots, slot = i % nslots;
            const double d = keys[i];
            if (d <= radii[b0 + b] && d < __longlong_as_double(0x7ff0000000000000ll))
            {
                const unsigned long long p = offs[b] + atomicAdd(&fill[b], 1u);
                out_keys[p] = d;
                out_vals[p] = (uint32_t)slot;
            }
        }

        __global__ void radius_emit_kernel(const double *keys, const uint32_t *vals, const uint64_t *slot_ids,
                                           size_t tot, uint64_t *out_ids, double *out_dists)
        {
            size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
            if (i >= tot)
                return;
            out_ids[i] = slot_ids[vals[i]];
            out_dists[i] = keys[i];
        }

        // ---- merge of per-shard partial results (one warp per query) ----
        // The nparts * k candidates of a query are staged in shared memory once; every lane then
        // ranks its candidates against all of them by (distance, id) with broadcast reads.
        __global__ void __launch_bounds__(128) merge_topk_kernel(const double *dists, const uint64_t *ids, int nparts,
                                                                 size_t part_stride, size_t nq, uint32_t k,
                                                                 double *out_d, uint64_t *out_i)
        {
            extern __shared__ __align__(16) unsigned char merge_smem[];
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            const size_t qi = blockIdx.x * 4ull + w;
            if (qi >= nq)
                return;
            const uint32_t n = (uint32_t)nparts * k;
            double *sk = reinterpret_cast<double *>(merge_smem) + (size_t)w * 2 * n;
            uint64_t *si = reinterpret_cast<uint64_t *>(sk + n);
            for (uint32_t c = lane; c < n; c += 32)
            {
                const size_t oc = (size_t)(c / k) * part_stride + qi * k + (c % k);
                const uint64_t id = ids[oc];
                si[c] = id;
                // absent entries sort after everything real
                sk[c] = id == OMPLCU_ID_NONE ? __longlong_as_double(0x7ff8000000000000ll) : dists[oc];
            }
            // pre-fill (fewer than k valid entries overall)
            for (uint32_t r = lane; r < k; r += 32)
            {
                out_i[qi * k + r] = OMPLCU_ID_NONE;
                out_d[qi * k + r] = __longlong_as_double(0x7ff0000000000000ll);
            }
            __syncwarp();
            for (uint32_t c = lane; c < n; c += 32)
            {
                const uint64_t id = si[c];
                if (id == OMPLCU_ID_NONE)
                    continue;
                const double key = sk[c];
                uint32_t rank = 0;
                for (uint32_t o = 0; o < n; ++o)
                {
                    const double ko = sk[o];  // NaN (absent) compares false both ways
                    rank += (ko < key || (ko == key && si[o] < id)) ? 1u : 0u;
                }
                if (rank < k)
                {
                    out_i[qi * k + rank] = id;
                    out_d[qi * k + rank] = key;
                }
            }
        }

        // The same merge for candidate counts that do not fit shared memory (FMT*'s k ~ 2100 on 8 shards): every
        // part is a list sorted by distance (absent entries last), so a candidate's rank among all candidates is the
        // sum over the parts of a lower bound by distance plus the members of the run of equal distances with a
        // smaller id -- nparts binary searches per candidate instead of a comparison with every other candidate, and
        // no staging: the lists of one query (nparts * k * 16 bytes) are read through L1.  A search never looks past
        // k - rank entries: a candidate that would reach rank k is dropped there, which is where most of them end.
        // One block per query.
        __global__ void __launch_bounds__(256) merge_sorted_kernel(const double *__restrict__ dists,
                                                                   const uint64_t *__restrict__ ids, int nparts,
                                                                   size_t part_stride, size_t nq, uint32_t k,
                                                                   double *__restrict__ out_d, uint64_t *__restrict__ out_i)
        {
            for (size_t qi = blockIdx.x; qi < nq; qi += gridDim.x)
            {
                for (uint32_t r = threadIdx.x; r < k; r += blockDim.x)
                {
                    out_i[qi * k + r] = OMPLCU_ID_NONE;
                    out_d[qi * k + r] = __longlong_as_double(0x7ff0000000000000ll);
                }
                __syncthreads();
                const uint32_t n = (uint32_t)nparts * k;
                for (uint32_t c = threadIdx.x; c < n; c += blockDim.x)
                {
                    const uint32_t p = c / k, j = c - p * k;
                    const size_t own = (size_t)p * part_stride + qi * k;
                    const uint64_t id = __ldg(ids + own + j);
                    if (id == OMPLCU_ID_NONE)
                        continue;
                    const double key = __ldg(dists + own + j);
                    uint32_t rank = 0;
                    bool keep = true;
                    for (int t = 0; t < nparts && keep; ++t)
                    {
                        const int pp = t == 0 ? (int)p : (t <= (int)p ? t - 1 : t);  // the own list first: rank ~ j at once
                        const size_t base = (size_t)pp * part_stride + qi * k;
                        const uint32_t lim = k - rank;
                        uint32_t lo = 0, hi = lim;
                        while (lo < hi)
                        {
                            const uint32_t mid = (lo + hi) >> 1;
                            const bool less = __ldg(ids + base + mid) != OMPLCU_ID_NONE && __ldg(dists + base + mid) < key;
                            if (less)
                                lo = mid + 1;
                            else
                                hi = mid;
                        }
                        rank += lo;
                        // equal distances: ordered by id whatever their order inside the list
                        for (uint32_t e = lo; e < k && rank < k; ++e)
                        {
                            const uint64_t ie = __ldg(ids + base + e);
                            if (ie == OMPLCU_ID_NONE || __ldg(dists + base + e) != key)
                                break;
                            rank += ie < id ? 1u : 0u;
                        }
                        keep = rank < k;
                    }
                    if (keep)
                    {
                        out_i[qi * k + rank] = id;
                        out_d[qi * k + rank] = key;
                    }
                }
                __syncthreads();
            }
        }

        inline unsigned nblk(size_t n, unsigned bs)
        {
            return (unsigned)((n + bs - 1) / bs);
        }
    }  // namespace

    // ==================================================================
    // NNIndex
    // ==================================================================
    NNIndex::NNIndex(const SpaceDesc &sp, int device) : device_(device), sp_(sp)
    {
        DeviceGuard g(device_);
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        cudaDeviceProp prop;
        CUDA_CHECK(cudaGetDeviceProperties(&prop, device_));
        sm_count_ = prop.multiProcessorCount;
        fast_space_ = (!sp_.compound && sp_.ncomp == 1 && sp_.comp[0].type == OMPLCU_RV &&
                       sp_.comp[0].dim >= 1 && sp_.comp[0].dim <= KNN_MAX_FAST_DIM);
        if (!fast_space_)
        {
            // SO(2) / SO(3) / compound spaces of them: Euclidean lower-bound embedding in front of the exact paths
            int ed = 0;
            bool ok = sp_.ncomp >= 1, rot = false;
            for (int i = 0; i < sp_.ncomp && ok; ++i)
            {
                const int e = embed_dim(sp_.comp[i]);
                ok = e > 0 && (!sp_.compound || (sp_.comp[i].weight >= 0.0 && std::isfinite(sp_.comp[i].weight)));
                ed += e;
                rot = rot || sp_.comp[i].type != OMPLCU_RV;
                if (sp_.comp[i].type == OMPLCU_SO3)
                    embed_slack_ += (sp_.compound ? sp_.comp[i].weight : 1.0) * EMBED_SO3_SLACK;
            }
            embed_ = ok && ed >= 1 && ed <= KNN_MAX_FAST_DIM && (rot || sp_.compound);
            if (embed_)
                D_ = ed;
        }
        if (fast_space_)
            D_ = sp_.comp[0].dim;
        if (fast_space_ || embed_)
        {
            rowf_ = scan_rowf(D_);
            center_.reserve(D_);
            pmax2_.reserve(1);
            embed_bad_.reserve(1);
            CUDA_CHECK(cudaMemsetAsync(pmax2_.p, 0, sizeof(unsigned), stream_));
            CUDA_CHECK(cudaMemsetAsync(embed_bad_.p, 0, sizeof(unsigned), stream_));
        }
        if (const char *e = getenv("OMPLCU_TC_SCAN"))
            tc_mode_ = atoi(e);  // default for the "tc_scan" option: -1 auto, 0 never, 1 whenever eligible
        if (const char *e = getenv("OMPLCU_TC_TS"))
            tc_ts_ = atoi(e);
        if (const char *e = getenv("OMPLCU_TC_ALIGN"))
            tc_align_ = atoi(e) != 0;
        if (const char *e = getenv("OMPLCU_TC_GRID"))
            tc_grid_ = atoi(e);
        if (const char *e = getenv("OMPLCU_TC_PAIR"))
            tc_pair_ = atoi(e);
        if (const char *e = getenv("OMPLCU_THR_SCAN"))
            thr_mode_ = atoi(e);  // default for the "thr_scan" option
        CUDA_CHECK(cudaStreamSynchronize(stream_));
    }

    // ---- per-query contexts ----
    NNIndex::Lease::Lease(NNIndex &i) : ix(i), c(nullptr)
    {
        {
            std::lock_guard<std::mutex> lk(ix.pool_mu_);
            if (!ix.free_.empty())
            {
                c = ix.free_.back();
                ix.free_.pop_back();
                return;
            }
        }
        // all contexts are busy (or this is the first query): a new one.  The caller holds a DeviceGuard.
        std::unique_ptr<QCtx> n(new QCtx);
        CUDA_CHECK(cudaStreamCreateWithFlags(&n->stream, cudaStreamNonBlocking));
        CUDA_CHECK(cudaEventCreate(&n->ev0_));
        CUDA_CHECK(cudaEventCreate(&n->ev1_));
        CUDA_CHECK(cudaEventCreate(&n->ev2_));
        CUDA_CHECK(cudaEventCreate(&n->ev3_));
        n->counters_.reserve(4);
        n->kept_.reserve(1);
        c = n.get();
        std::lock_guard<std::mutex> lk(ix.pool_mu_);
        ix.pool_.push_back(std::move(n));
    }

    NNIndex::Lease::~Lease()
    {
        std::lock_guard<std::mutex> lk(ix.pool_mu_);
        ix.free_.push_back(c);
    }

    void NNIndex::publish(const QCtx &c)
    {
        std::lock_guard<std::mutex> lk(stat_mu_);
        last_collect_redo_ = c.last_collect_redo_;
        last_filter_ = c.last_filter_;
        last_tc_ = c.last_tc_;
        last_thr_ = c.last_thr_;
        last_kept_ = c.last_kept_;
        last_rank_ = c.last_rank_;
        last_sample_ns_ = c.last_sample_ns_;
        last_main_ns_ = c.last_main_ns_;
        last_scan_ns_ = c.last_scan_ns_;
        last_fallbacks_ = c.last_fallbacks_;
        last_rescans_ = c.last_rescans_;
        last_few_ = c.last_few_;
        last_proven_ = c.last_proven_;
        last_path_ = c.last_path_;
        if (profile_ && c.last_path_ == 1)
        {
            scan_ns_total_ += c.last_scan_ns_;
            scan_count_++;
        }
    }

    NNIndex::~NNIndex()
    {
        DeviceGuard g(device_);
        if (stream_)
        {
            cudaStreamSynchronize(stream_);
            cudaStreamDestroy(stream_);
        }
        for (auto &c : pool_)
        {
            if (c->stream)
            {
                cudaStreamSynchronize(c->stream);
                cudaStreamDestroy(c->stream);
            }
            for (cudaEvent_t e : {c->ev0_, c->ev1_, c->ev2_, c->ev3_})
                if (e)
                    cudaEventDestroy(e);
        }
    }

    void NNIndex::set_id_base(uint64_t base)
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        REQUIRE(nslots_ == 0, "omplcu_nn_set_id_base: structure is not empty");
        id_base_ = base;
    }

    void NNIndex::reserve_slots(size_t n)
    {
        const size_t st = (size_t)sp_.stride;
        pts64_.reserve(n * st, true, stream_);
        ids_.reserve(n, true, stream_);
        alive_.reserve(n, true, stream_);
        if (fast_space_ || embed_)
        {
            size_t rows = ((n + SCAN_PT - 1) / SCAN_PT) * SCAN_PT;
            if (rows > filt_rows_)
            {
                // grow geometrically so repeated single adds stay amortised O(1)
                size_t want = std::max(rows, filt_rows_ * 2);
                want = ((want + SCAN_PT - 1) / SCAN_PT) * SCAN_PT;
                filt_.reserve(want * rowf_, true, stream_);
                want = (filt_.cap / rowf_ / SCAN_PT) * SCAN_PT;
                size_t fresh = want - filt_rows_;
                filt_fill_dead_kernel<<<nblk(fresh * rowf_, 256), 256, 0, stream_>>>(filt_.p, rowf_, D_, filt_rows_,
                                                                                    fresh);
                launches_++;
                CUDA_CHECK(cudaGetLastError());
                filt_rows_ = want;
            }
        }
    }

    uint64_t NNIndex::add_at(const double *rows, size_t n, uint64_t want_first)
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        if (want_first != OMPLCU_ID_NONE)
        {
            REQUIRE(want_first >= id_base_ + next_seq_, "omplcu_nn_add: ids must increase");
            next_seq_ = want_first - id_base_;
        }
        const uint64_t first_id = id_base_ + next_seq_;
        if (n == 0)
            return first_id;
        REQUIRE(rows != nullptr, "omplcu_nn_add: rows is NULL");
        REQUIRE(nslots_ + n < (size_t)0xFFFFFFF0u, "omplcu_nn_add: more than 2^32 slots");
        const size_t st = (size_t)sp_.stride;
        reserve_slots(nslots_ + n);
        CUDA_CHECK(cudaMemcpyAsync(pts64_.p + nslots_ * st, rows, n * st * sizeof(double), cudaMemcpyHostToDevice,
                                   stream_));
        fill_ids_kernel<<<nblk(n, 256), 256, 0, stream_>>>(ids_.p, alive_.p, nslots_, n, first_id);
        launches_++;
        if (fast_space_ || embed_)
        {
            if (!have_center_)
            {
                // centre = mean of the first batch (any fixed vector is valid; a
                // central one keeps the FP32 error bound tight)
                std::vector<double> c(D_, 0.0);
                const size_t nmean = embed_ ? std::min<size_t>(n, 4096) : n;
                std::vector<double> e(embed_ ? D_ : 0);
                for (size_t i = 0; i < nmean; ++i)
                {
                    if (embed_)
                        embed_state(sp_, rows + i * st, e.data());
                    for (int d = 0; d < D_; ++d)
                        c[d] += embed_ ? e[d] : rows[i * st + d];
                }
                for (int d = 0; d < D_; ++d)
                {
                    c[d] /= (double)nmean;
                    if (!std::isfinite(c[d]))
                        c[d] = 0.0;
                }
                CUDA_CHECK(
                    cudaMemcpyAsync(center_.p, c.data(), D_ * sizeof(double), cudaMemcpyHostToDevice, stream_));
                CUDA_CHECK(cudaStreamSynchronize(stream_));  // c is a stack-scoped vector
                have_center_ = true;
            }
            if (embed_)
                filt_append_embed_kernel<<<nblk(n, 128), 128, 0, stream_>>>(sp_, pts64_.p, center_.p, filt_.p, pmax2_.p, D_,
                                                                            rowf_, nslots_, n, embed_bad_.p);
            else
            filt_append_kernel<<<nblk(n, 256), 256, 0, stream_>>>(pts64_.p, center_.p, filt_.p, pmax2_.p, D_, rowf_,
                                                                  nslots_, n);
            launches_++;
        }
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(stream_));  // caller may free `rows`
        h_ids_.reserve(nslots_ + n);
        for (size_t i = 0; i < n; ++i)
            h_ids_.push_back(first_id + i);
        h_alive_.insert(h_alive_.end(), n, 1);
        nslots_ += n;
        nlive_ += n;
        next_seq_ += n;
        mut_version_++;
        return first_id;
    }

    void NNIndex::remove(uint64_t id)
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        auto it = std::lower_bound(h_ids_.begin(), h_ids_.end(), id);
        if (it == h_ids_.end() || *it != id || !h_alive_[it - h_ids_.begin()])
            throw Error(OMPLCU_ERR_NOT_FOUND, "omplcu_nn_remove: id not present");
        const size_t slot = it - h_ids_.begin();
        h_alive_[slot] = 0;
        nlive_--;
        mut_version_++;
        mark_dead_kernel<<<1, 1, 0, stream_>>>(alive_.p, (fast_space_ || embed_) ? filt_.p : nullptr, rowf_, D_, slot);
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        const size_t dead = nslots_ - nlive_;
        if (dead > 1024 && dead > nlive_)
            compact();
        CUDA_CHECK(cudaStreamSynchronize(stream_));
    }

    // Bulk removal: one kernel marks every slot, one synchronisation.  Absent ids are skipped.
    size_t NNIndex::remove_batch(const uint64_t *ids, size_t n)
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        std::vector<uint32_t> slots;
        slots.reserve(n);
        for (size_t i = 0; i < n; ++i)
        {
            auto it = std::lower_bound(h_ids_.begin(), h_ids_.end(), ids[i]);
            if (it == h_ids_.end() || *it != ids[i] || !h_alive_[it - h_ids_.begin()])
                continue;
            h_alive_[it - h_ids_.begin()] = 0;
            slots.push_back((uint32_t)(it - h_ids_.begin()));
        }
        if (slots.empty())
            return 0;
        nlive_ -= slots.size();
        mut_version_++;
        DevBuf<uint32_t> d_slots;
        d_slots.reserve(slots.size());
        CUDA_CHECK(cudaMemcpyAsync(d_slots.p, slots.data(), slots.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                   stream_));
        mark_dead_batch_kernel<<<nblk(slots.size(), 256), 256, 0, stream_>>>(alive_.p, (fast_space_ || embed_) ? filt_.p : nullptr,
                                                                             rowf_, D_, d_slots.p, slots.size());
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(stream_));
        const size_t dead = nslots_ - nlive_;
        if (dead > 1024 && dead > nlive_)
        {
            compact();
            CUDA_CHECK(cudaStreamSynchronize(stream_));
        }
        return slots.size();
    }

    // Drop dead slots, keeping insertion order (so slot order stays id order).
    void NNIndex::compact()
    {
        const size_t st = (size_t)sp_.stride;
        std::vector<uint32_t> live;
        live.reserve(nlive_);
        std::vector<uint64_t> nids;
        nids.reserve(nlive_);
        for (size_t s = 0; s < nslots_; ++s)
            if (h_alive_[s])
            {
                live.push_back((uint32_t)s);
                nids.push_back(h_ids_[s]);
            }
        if (!live.empty())
        {
            DevBuf<uint32_t> d_live;
            DevBuf<double> npts;
            DevBuf<uint64_t> nidb;
            d_live.reserve(live.size());
            npts.reserve(live.size() * st);
            nidb.reserve(live.size());
            CUDA_CHECK(cudaMemcpyAsync(d_live.p, live.data(), live.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                       stream_));
            gather_rows_kernel<<<nblk(live.size() * st, 256), 256, 0, stream_>>>(pts64_.p, npts.p, ids_.p, nidb.p,
                                                                                 d_live.p, live.size(), (int)st);
            launches_++;
            CUDA_CHECK(cudaMemcpyAsync(pts64_.p, npts.p, live.size() * st * sizeof(double), cudaMemcpyDeviceToDevice,
                                       stream_));
            CUDA_CHECK(
                cudaMemcpyAsync(ids_.p, nidb.p, live.size() * sizeof(uint64_t), cudaMemcpyDeviceToDevice, stream_));
            CUDA_CHECK(cudaMemsetAsync(alive_.p, 1, live.size(), stream_));
            CUDA_CHECK(cudaStreamSynchronize(stream_));
        }
        const size_t old_slots = nslots_;
        nslots_ = live.size();
        trig_rows_ = 0;  // slots moved: the chain trig table is rebuilt on the next query
        h_ids_.swap(nids);
        h_alive_.assign(nslots_, 1);
        if (fast_space_ || embed_)
        {
            // rebuild the filter copy for the surviving rows, kill the vacated ones
            if (nslots_)
            {
                if (embed_)
                {
                    CUDA_CHECK(cudaMemsetAsync(embed_bad_.p, 0, sizeof(unsigned), stream_));  // the offender may be gone
                    embed_ver_ = 0;
                    filt_append_embed_kernel<<<nblk(nslots_, 128), 128, 0, stream_>>>(sp_, pts64_.p, center_.p, filt_.p,
                                                                                      pmax2_.p, D_, rowf_, 0, nslots_,
                                                                                      embed_bad_.p);
                }
                else
                filt_append_kernel<<<nblk(nslots_, 256), 256, 0, stream_>>>(pts64_.p, center_.p, filt_.p, pmax2_.p,
                                                                            D_, rowf_, 0, nslots_);
                launches_++;
            }
            if (old_slots > nslots_)
            {
                filt_fill_dead_kernel<<<nblk((old_slots - nslots_) * rowf_, 256), 256, 0, stream_>>>(
                    filt_.p, rowf_, D_, nslots_, old_slots - nslots_);
                launches_++;
            }
            CUDA_CHECK(cudaGetLastError());
        }
    }

    void NNIndex::clear()
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        if ((fast_space_ || embed_) && nslots_)
        {
            filt_fill_dead_kernel<<<nblk(nslots_ * rowf_, 256), 256, 0, stream_>>>(filt_.p, rowf_, D_, 0, nslots_);
            launches_++;
            CUDA_CHECK(cudaMemsetAsync(pmax2_.p, 0, sizeof(unsigned), stream_));
            CUDA_CHECK(cudaMemsetAsync(embed_bad_.p, 0, sizeof(unsigned), stream_));
            embed_ver_ = 0;
            CUDA_CHECK(cudaStreamSynchronize(stream_));
        }
        have_center_ = false;
        nslots_ = nlive_ = 0;
        mut_version_++;
        trig_rows_ = 0;
        h_ids_.clear();
        h_alive_.clear();
        // ids keep increasing across clear() so stale ids can never alias new points
    }

    size_t NNIndex::list(uint64_t *ids, size_t cap) const
    {
        std::shared_lock<std::shared_mutex> lk(rw_);
        size_t c = 0;
        for (size_t s = 0; s < nslots_; ++s)
            if (h_alive_[s])
            {
                if (c < cap && ids)
                    ids[c] = h_ids_[s];
                c++;
            }
        return c;
    }

    void NNIndex::get(uint64_t id, double *row)
    {
        std::shared_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        auto it = std::lower_bound(h_ids_.begin(), h_ids_.end(), id);
        if (it == h_ids_.end() || *it != id || !h_alive_[it - h_ids_.begin()])
            throw Error(OMPLCU_ERR_NOT_FOUND, "omplcu_nn_get: id not present");
        const size_t slot = it - h_ids_.begin();
        CUDA_CHECK(cudaMemcpyAsync(row, pts64_.p + slot * sp_.stride, sp_.stride * sizeof(double),
                                   cudaMemcpyDeviceToHost, stream_));
        CUDA_CHECK(cudaStreamSynchronize(stream_));
    }

    void NNIndex::set_option(const char *key, long v)
    {
        std::unique_lock<std::shared_mutex> lk(rw_);
        if (!strcmp(key, "force_exact"))
            force_exact_ = v != 0;
        else if (!strcmp(key, "sample_thresholds"))
            sample_thr_ = v != 0;
        else if (!strcmp(key, "collect"))
            collect_ = v != 0;
        else if (!strcmp(key, "scan_ctas_per_sm"))
            scan_ctas_per_sm_ = (int)v;  // 0 = as many as fit
        else if (!strcmp(key, "tc_scan"))
            tc_mode_ = (int)v;  // -1 auto, 0 never, 1 whenever the space is eligible
        else if (!strcmp(key, "tc_tiles"))
            tc_na_ = (int)v;  // upper limit on resident query tiles per CTA (1, 2, 4); 0 = by shape
        else if (!strcmp(key, "tc_ts"))
            tc_ts_ = (int)v;  // one resident tile (d >= 53, or small batches): query tile in tensor memory
        else if (!strcmp(key, "sort_queries"))
            sort_queries_ = v != 0;  // batches of >= 4096 queries are scanned in Morton order (results in the caller's order)
        else if (!strcmp(key, "few_scan"))
            few_mode_ = v != 0;  // up to 8 queries: lane-per-point FP32 scan instead of a mostly empty tensor-core tile
        else if (!strcmp(key, "tc_align"))
            tc_align_ = v != 0;  // work pieces of the tensor-core scan end on query-tile boundaries when that costs < 15 % of the SMs
        else if (!strcmp(key, "tc_pair"))
            tc_pair_ = (int)v;  // with tc_ts: CTA pairs (clusters of 2) share every point chunk by TMA multicast
        else if (!strcmp(key, "l2_window"))
            l2_window_ = v != 0;  // persisting L2 window over the operand image during the tensor-core scan
        else if (!strcmp(key, "thr_scan"))
            thr_mode_ = (int)v;  // -1 auto, 0 never (list scan), 1 whenever the tensor-core scan runs
        else if (!strcmp(key, "thr_sample_step"))
        {
            thr_step_ = (int)std::max<long>(1, v);
            tc_sversion_ = 0;
        }
        else if (!strcmp(key, "tc_variant"))
        {
            tc_variant_ = (int)v;
            tc_version_ = 0;
        }
        else if (!strcmp(key, "tc_scan_values_out"))
            tc_dbg_ = reinterpret_cast<float *>(v);  // diagnostics: device buffer [qpad][units * 128] floats, 0 = off
        else if (!strcmp(key, "profile"))
        {
            // time the scan kernel with CUDA events on the launching stream
            profile_ = v != 0;
            std::lock_guard<std::mutex> sl(stat_mu_);
            scan_ns_total_ = 0;
            scan_count_ = 0;
        }
        else
            throw Error(OMPLCU_ERR_INVALID, std::string("unknown option ") + key);
    }

    long NNIndex::get_stat(const char *key) const
    {
        std::lock_guard<std::mutex> lk(stat_mu_);
        if (!strcmp(key, "last_fallbacks"))
            return last_fallbacks_;
        if (!strcmp(key, "tc_demoted"))
            return tc_demoted_.load() ? 1 : 0;  // "auto" gave up on the tensor-core filter for this set (too dense)
        if (!strcmp(key, "last_proven"))
            return last_proven_;  // embedded spaces: queries the list scan alone proved (no threshold scan needed)
        if (!strcmp(key, "last_few"))
            return last_few_;  // the last query ran the few-query scan (knn_few_kernel)
        if (!strcmp(key, "last_rescans"))
            return last_rescans_;  // queries the tensor-core filter passed on to the FP32-pipe scan
        if (!strcmp(key, "last_path"))
            return last_path_;
        if (!strcmp(key, "last_tc"))
            return last_tc_;
        if (!strcmp(key, "launches"))
            return launches_;
        if (!strcmp(key, "last_thr"))
            return last_thr_;
        if (!strcmp(key, "last_filter"))
            return last_filter_;
        if (!strcmp(key, "last_kept"))
            return last_kept_;
        if (!strcmp(key, "last_rank"))
            return last_rank_;
        if (!strcmp(key, "last_sample_ns"))
            return last_sample_ns_;
        if (!strcmp(key, "last_main_ns"))
            return last_main_ns_;
        if (!strcmp(key, "last_scan_ns"))
            return last_scan_ns_;
        if (!strcmp(key, "scan_ns_total"))
            return scan_ns_total_;
        if (!strcmp(key, "scan_count"))
            return scan_count_;
        if (!strcmp(key, "slots"))
            return (long)nslots_;
        if (!strcmp(key, "fast_space"))
            return fast_space_ ? 1 : 0;
        if (!strcmp(key, "last_collect_redo"))
            return last_collect_redo_;
        throw Error(OMPLCU_ERR_INVALID, std::string("unknown stat ") + key);
    }

    // ==================================================================
    // queries
    // ==================================================================
    bool NNIndex::fast_eligible(size_t kk) const
    {
        return fast_space_ && !force_exact_ && kk >= 1 && kk <= 40 && nslots_ >= 1;
    }

    namespace
    {
        // action 0: report resident CTAs per SM; action 1: launch
        template <int D, int RQ, int L>
        void scan_action(int action, const ScanParams *P, unsigned grid, cudaStream_t s, int *occ)
        {
            const size_t smem = scan_smem_bytes(D, RQ, L);
            static bool configured[64] = {};  // per device
            static std::mutex configured_mu;
            std::lock_guard<std::mutex> cl(configured_mu);
            int dev = 0;
            cudaGetDevice(&dev);
            if (!configured[dev & 63])
            {
                CUDA_CHECK(cudaFuncSetAttribute(knn_scan_kernel<D, RQ, L>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)smem));
                configured[dev & 63] = true;
            }
            if (action == 0)
                CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, knn_scan_kernel<D, RQ, L>, SCAN_THREADS,
                                                                         smem));
            else
                knn_scan_kernel<D, RQ, L><<<grid, SCAN_THREADS, smem, s>>>(*P);
        }

        template <int RQ, int L>
        void dispatch_scan(int D, int action, const ScanParams *P, unsigned grid, cudaStream_t s, int *occ)
        {
            switch (D)
            {
#define OMPLCU_SCAN_CASE(d)                                  \
    case d:                                                  \
        scan_action<d, RQ, L>(action, P, grid, s, occ);      \
        break;
                OMPLCU_SCAN_CASE(1)
                OMPLCU_SCAN_CASE(2)
                OMPLCU_SCAN_CASE(3)
                OMPLCU_SCAN_CASE(4)
                OMPLCU_SCAN_CASE(5)
                OMPLCU_SCAN_CASE(6)
                OMPLCU_SCAN_CASE(7)
                OMPLCU_SCAN_CASE(8)
                OMPLCU_SCAN_CASE(9)
                OMPLCU_SCAN_CASE(10)
                OMPLCU_SCAN_CASE(11)
                OMPLCU_SCAN_CASE(12)
                OMPLCU_SCAN_CASE(13)
                OMPLCU_SCAN_CASE(14)
                OMPLCU_SCAN_CASE(15)
                OMPLCU_SCAN_CASE(16)
#undef OMPLCU_SCAN_CASE
                default:
                    throw Error(OMPLCU_ERR_INTERNAL, "scan: unsupported dimension");
            }
        }

        // (RQ, L) per k range: k<=10 -> (4,12); k<=14 -> (4,16); k<=40 -> (2,42)
        void scan_any(int cfg, int D, int action, const ScanParams *P, unsigned grid, cudaStream_t s, int *occ)
        {
            if (cfg == 0)
                dispatch_scan<4, 12>(D, action, P, grid, s, occ);
            else if (cfg == 1)
                dispatch_scan<4, 16>(D, action, P, grid, s, occ);
            else
                dispatch_scan<2, 42>(D, action, P, grid, s, occ);
        }
    }  // namespace

    namespace
    {
        template <int D>
        void launch_thr_sample_d(const float4 *rows, uint32_t nsample, const float *qp, uint32_t qpad, uint32_t L,
                                 float *out, cudaStream_t s)
        {
            constexpr int SQ = D <= 8 ? 8 : 4;
            thr_sample_kernel<D, SQ><<<nblk(qpad / SQ, 4), 128, 0, s>>>(rows, nsample, qp, qpad, L, out);
        }

        void launch_thr_sample(int D, const float4 *rows, uint32_t nsample, const float *qp, uint32_t qpad, uint32_t L,
                               float *out, cudaStream_t s)
        {
            switch (D)
            {
#define OMPLCU_THR_CASE(d)                                            \
    case d:                                                           \
        launch_thr_sample_d<d>(rows, nsample, qp, qpad, L, out, s);   \
        break;
                OMPLCU_THR_CASE(1)
                OMPLCU_THR_CASE(2)
                OMPLCU_THR_CASE(3)
                OMPLCU_THR_CASE(4)
                OMPLCU_THR_CASE(5)
                OMPLCU_THR_CASE(6)
                OMPLCU_THR_CASE(7)
                OMPLCU_THR_CASE(8)
                OMPLCU_THR_CASE(9)
                OMPLCU_THR_CASE(10)
                OMPLCU_THR_CASE(11)
                OMPLCU_THR_CASE(12)
                OMPLCU_THR_CASE(13)
                OMPLCU_THR_CASE(14)
                OMPLCU_THR_CASE(15)
                OMPLCU_THR_CASE(16)
#undef OMPLCU_THR_CASE
                default:
                    throw Error(OMPLCU_ERR_INTERNAL, "threshold sample: unsupported dimension");
            }
        }
    }  // namespace

    namespace
    {
        // action 0: report resident CTAs per SM; action 1: launch
        template <int L>
        void hd_action(int action, int D, int rowf, const ScanParams *P, unsigned grid, cudaStream_t s, int *occ)
        {
            const size_t smem = hd_smem_bytes(D, L);
            CUDA_CHECK(cudaFuncSetAttribute(knn_scan_hd_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
            if (action == 0)
                CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, knn_scan_hd_kernel<L>, HD_THREADS, smem));
            else
                knn_scan_hd_kernel<L><<<grid, HD_THREADS, smem, s>>>(*P, D, rowf);
        }

        void hd_any(int cfg, int action, int D, int rowf, const ScanParams *P, unsigned grid, cudaStream_t s, int *occ)
        {
            if (cfg == 0)
                hd_action<12>(action, D, rowf, P, grid, s, occ);
            else if (cfg == 1)
                hd_action<16>(action, D, rowf, P, grid, s, occ);
            else
                hd_action<42>(action, D, rowf, P, grid, s, occ);
        }
    }  // namespace

    // scan configuration: (RQ, L) = (4, 16) for k <= 12, (2, 48) for k <= 40
    // nq <= FEW_MAXQ queries: knn_few_kernel (one warp = one segment of FEW_L candidates), then the same exact
    // re-rank, band check and fallback as every other filter.
    void NNIndex::knn_few(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids,
                          double *d_dists, cudaStream_t s)
    {
        const uint32_t qpad = 32;
        const unsigned grid = (unsigned)std::max<size_t>(1, std::min<size_t>((size_t)sm_count_, (nslots_ + 127) / 128));
        const uint32_t smax = grid * FEW_WARPS;
        c.qp_.reserve((size_t)D_ * qpad);
        c.qaux_.reserve(2 * (size_t)qpad);
        c.cand_vals_.reserve((size_t)smax * qpad);
        c.cand_ids_.reserve((size_t)smax * FEW_L * qpad);
        c.ovf_list_.reserve(nq);
        query_prep_kernel<<<1, 32, 0, s>>>(d_q, nq, center_.p, D_, c.qp_.p, c.qaux_.p, qpad, c.counters_.p);
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev0_, s));
        const size_t qsm = (size_t)D_ * FEW_MAXQ * sizeof(float);
        if (nq == 1)
            knn_few_kernel<1><<<grid, FEW_WARPS * 32, qsm, s>>>(filt_.p, rowf_, D_, nslots_, c.qp_.p, qpad, c.cand_vals_.p,
                                                                c.cand_ids_.p);
        else if (nq <= 4)
            knn_few_kernel<4><<<grid, FEW_WARPS * 32, qsm, s>>>(filt_.p, rowf_, D_, nslots_, c.qp_.p, qpad, c.cand_vals_.p,
                                                                c.cand_ids_.p);
        else
            knn_few_kernel<FEW_MAXQ><<<grid, FEW_WARPS * 32, qsm, s>>>(filt_.p, rowf_, D_, nslots_, c.qp_.p, qpad,
                                                                       c.cand_vals_.p, c.cand_ids_.p);
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev1_, s));
        knn_rerank_select_kernel<<<(unsigned)nq, SEL_THREADS,
                                   (size_t)smax * FEW_L * 12 + (size_t)D_ * 8 + 8 + (size_t)(SEL_THREADS / 32) * kk * 12, s>>>(
            c.cand_vals_.p, c.cand_ids_.p, smax, (uint32_t)FEW_L, qpad, pts64_.p, ids_.p, D_, d_q, c.qaux_.p, pmax2_.p, nq,
            (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists, c.ovf_list_.p, c.counters_.p, (double)(D_ + 3));
        launches_ += 3;
        CUDA_CHECK(cudaGetLastError());
        const bool ho = c.ho_ids_ && c.ho_entries_ == nq * kstride && kk == kstride;
        if (ho)
        {
            CUDA_CHECK(cudaMemcpyAsync(c.ho_ids_, d_ids, c.ho_entries_ * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(c.ho_dists_, d_dists, c.ho_entries_ * sizeof(double), cudaMemcpyDeviceToHost, s));
        }
        unsigned novf = 0;
        CUDA_CHECK(cudaMemcpyAsync(&novf, c.counters_.p, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        c.ho_done_ = ho && novf == 0;
        c.last_fallbacks_ = novf;
        c.last_tc_ = 0;
        c.last_thr_ = 0;
        c.last_rescans_ = 0;
        c.last_few_ = 1;
        if (profile_)
        {
            float ms = 0.f;
            CUDA_CHECK(cudaEventElapsedTime(&ms, c.ev0_, c.ev1_));
            c.last_scan_ns_ = (long)(ms * 1e6);
        }
        if (novf)
            knn_exact(c, d_q, c.ovf_list_.p, novf, kk, kstride, d_ids, d_dists, s);
    }

    void NNIndex::knn_fast(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids,
                           double *d_dists, cudaStream_t s, bool no_tc)
    {
        c.last_few_ = 0;
        if (!no_tc && few_mode_ && nq <= (size_t)FEW_MAXQ && tc_mode_ < 0 && thr_mode_ <= 0 && nslots_ < ((size_t)1 << 32) - 64)
        {
            knn_few(c, d_q, nq, kk, kstride, d_ids, d_dists, s);
            return;
        }
        const int cfg = kk <= 10 ? 0 : (kk <= 14 ? 1 : 2);
        const int RQ = cfg == 2 ? 2 : 4;
        const bool hd = D_ > KNN_MAX_REG_DIM;  // register-tiled kernel of knn_scan_hd.cuh
        // tensor cores for the dense-contraction case (knn_scan_tc.cuh): d >= 64 and a real batch.
        // Its scan values carry a wider error band than the FP32 kernels', so its lists keep more
        // spare entries (16 / 24 / 48 instead of 12 / 16 / 42): the band check then compares the
        // k-th exact distance with the (k + 6)-th scan value or later, and fall-backs stay rare.
        TcShape tc_sh = tc_shape(D_);
        if (tc_na_ > 0 && tc_na_ < tc_sh.na)
            tc_sh.na = tc_na_;
        const int tc_L = cfg == 0 ? 16 : (cfg == 1 ? 24 : 48);
        int tc_stages = tc_pick_stages(tc_sh, tc_L);
        while (tc_stages < 3 && tc_sh.na > 1)
        {
            tc_sh.na /= 2;  // long lists: fewer resident query tiles
            tc_stages = tc_pick_stages(tc_sh, tc_L);
        }
        // a small batch needs no more resident query tiles than it fills: a lone query runs 1 MMA chain per point
        // tile instead of 4
        while (tc_sh.na > 1 && nq <= (size_t)TC_QT * (tc_sh.na / 2))
        {
            tc_sh.na /= 2;
            tc_stages = tc_pick_stages(tc_sh, tc_L);
        }
        // one resident tile: the query tile goes into tensor memory and its shared-memory copy makes room for a
        // deeper TMA ring (these shapes are bound by the L2 -> SM feed)
        const bool use_ts = tc_ts_ && tc_ts_usable(tc_sh) && !tc_dbg_;
        if (use_ts)
            tc_stages = tc_pick_stages(tc_sh, tc_L, true);
        // ... and CTA pairs share every point chunk (knn_scan_tc.cu, PAIR): super-tiles of two query tiles
        const bool pair = use_ts && tc_pair_ && nq > (size_t)TC_QT && sm_count_ >= 2;
        const bool tc_auto = nq >= TC_MIN_BATCH;  // measured faster than the FP32 scans at every batched size
        if (tc_demoted_.load(std::memory_order_relaxed) && nslots_ < tc_demote_slots_.load(std::memory_order_relaxed) / 2)
            tc_demoted_.store(false, std::memory_order_relaxed);  // the set thinned out: give the tensor cores another go
        const bool tc = !no_tc && D_ <= TC_MAX_DIM && tc_stages > 0 &&
                        (tc_mode_ == 1 || (tc_mode_ < 0 && tc_auto && !tc_demoted_.load(std::memory_order_relaxed)));
        c.last_tc_ = tc ? 1 : 0;
        c.last_thr_ = 0;
        if (tc && (thr_mode_ == 1 || (thr_mode_ < 0 && nslots_ >= THR_MIN_SLOTS)) && nslots_ >= 4096)
        {
            knn_fast_thr(c, d_q, nq, kk, kstride, d_ids, d_dists, s);
            return;
        }
        const int L = tc ? tc_L : (cfg == 0 ? 12 : (cfg == 1 ? 16 : 42));
        const size_t QT = tc ? (size_t)TC_QT * (pair ? 2 : tc_sh.na) : hd ? (size_t)HD_QT : (size_t)SCAN_THREADS * RQ;
        const size_t PT = tc ? (size_t)TC_PT : hd ? (size_t)HD_PT : (size_t)SCAN_PT;
        const size_t qtiles = (nq + QT - 1) / QT;
        const uint32_t qpad = (uint32_t)(qtiles * QT);
        const unsigned long long U = (nslots_ + PT - 1) / PT;
        const unsigned long long TU = qtiles * U;
        int resident = 1;  // CTAs per SM: one persistent CTA per resident slot
        if (tc)
            resident = 1;  // all of shared memory and all 512 TMEM columns
        else if (hd)
            hd_any(cfg, 0, D_, rowf_, nullptr, 0, s, &resident);
        else
            scan_any(cfg, D_, 0, nullptr, 0, s, &resident);
        resident = std::max(resident, 1);
        if (scan_ctas_per_sm_ > 0)
            resident = std::min(resident, scan_ctas_per_sm_);
        // candidate slots per query: 512 for the all-pairs ranking of knn_rerank_kernel; a batch of one query
        // tile may be cut into up to one segment per SM (2560 slots, knn_rerank_select_kernel), so that a
        // single query -- the literal nearestK of PRM::addMilestone, prm/src/PRM.cpp:575 -- streams the image
        // on every SM instead of 31
        constexpr unsigned FEW_SLOTS = 2560;
        const bool few = tc && qtiles == 1;
        const unsigned max_seg = (few ? FEW_SLOTS : 512u) / L;
        unsigned long long G = (unsigned long long)sm_count_ * resident;  // work pieces
        if (tc && pair)
            G = (unsigned long long)sm_count_ / 2;  // one per CTA pair
        if (tc && tc_grid_ > 0)
            G = std::min<unsigned long long>(G, (unsigned long long)tc_grid_);
        G = std::min<unsigned long long>(G, TU);
        G = std::min<unsigned long long>(G, qtiles * (unsigned long long)(max_seg - 1));
        G = std::max<unsigned long long>(G, 1);
        if (tc && tc_align_ && tc_sh.na >= 2 && !pair)
        {
            // Pieces that end on query-tile boundaries: every (query, piece) pair pays a list warm-up of about
            // L (1 + ln(points / L)) insertions and adds L candidates to the re-rank, so 128 super-tiles on 128 CTAs
            // (one segment per query) beat the same work cut into 148 equal pieces (2.15 segments per query) at
            // every size measured but the largest, where it is a tie (profiles/r2_experiments.md).  Not for the
            // tensor-bound shapes (one resident tile), which want every SM.
            unsigned long long Ga = 0;
            if (qtiles <= G)
                Ga = qtiles * (G / qtiles);
            else
                for (unsigned long long x = G; x >= 1; --x)
                    if (qtiles % x == 0)
                    {
                        Ga = x;
                        break;
                    }
            if (Ga * 100 >= G * 85)
                G = Ga;
        }
        // pieces overlapping one query tile (same formulas as the kernel)
        uint32_t smax = 1;
        for (size_t t = 0; t < qtiles; ++t)
        {
            unsigned long long x0 = t * U, x1 = (t + 1) * U - 1;
            unsigned long long g0 = ((x0 + 1) * G + TU - 1) / TU - 1;
            unsigned long long g1 = ((x1 + 1) * G + TU - 1) / TU - 1;
            smax = std::max<uint32_t>(smax, (uint32_t)(g1 - g0 + 1));
        }
        if ((size_t)smax * L > (few ? FEW_SLOTS : 512u))
            throw Error(OMPLCU_ERR_INTERNAL, "scan: too many segments per query tile");

        c.qp_.reserve((size_t)D_ * qpad);
        c.qaux_.reserve(2 * (size_t)qpad);
        c.cand_vals_.reserve((size_t)smax * qpad);
        c.cand_ids_.reserve((size_t)smax * L * qpad);
        c.ovf_list_.reserve(nq);
        // Rows of segments that no piece writes must read "no candidate, not full".  When every query tile is cut
        // into exactly smax pieces (aligned pieces, or one tile) the tensor-core scan writes every row itself.
        if (!(tc && G % qtiles == 0 && G / qtiles == smax))
        {
            CUDA_CHECK(cudaMemsetAsync(c.cand_ids_.p, 0xFF, (size_t)smax * L * qpad * sizeof(uint32_t), s));
            CUDA_CHECK(cudaMemsetAsync(c.cand_vals_.p, 0x7F, (size_t)smax * qpad * sizeof(float), s));  // 3.4e38: "not full"
        }
        // large batches: scan order = Morton order of the queries (qsort_*_kernel above); results go back to the
        // caller's rows through `perm`
        const uint32_t *perm = nullptr;
        if (tc && sort_queries_ && nq >= 4096 && !tc_dbg_ && tc_sh.na >= 2)  // the one-tile shapes are tensor-bound: no gain
        {
            const int nd = std::min(D_, QSORT_DIMS);
            c.qs_keys_.reserve(2 * nq);
            c.qs_idx_.reserve(2 * nq);
            c.qs_bbox_.reserve(2 * QSORT_DIMS);
            CUDA_CHECK(cudaMemsetAsync(c.qs_bbox_.p, 0xFF, QSORT_DIMS * sizeof(uint32_t), s));
            CUDA_CHECK(cudaMemsetAsync(c.qs_bbox_.p + QSORT_DIMS, 0, QSORT_DIMS * sizeof(uint32_t), s));
            qsort_bbox_kernel<<<nblk(nq, 256), 256, 0, s>>>(d_q, nq, D_, nd, c.qs_bbox_.p);
            qsort_key_kernel<<<nblk(nq, 256), 256, 0, s>>>(d_q, nq, D_, nd, c.qs_bbox_.p, c.qs_keys_.p, c.qs_idx_.p);
            cub::DoubleBuffer<uint32_t> dk(c.qs_keys_.p, c.qs_keys_.p + nq), dv(c.qs_idx_.p, c.qs_idx_.p + nq);
            size_t tmp = 0;
            CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (int)nq, 0, 2 * nd, s));
            c.cub_tmp_.reserve(tmp);
            CUDA_CHECK(cub::DeviceRadixSort::SortPairs(c.cub_tmp_.p, tmp, dk, dv, (int)nq, 0, 2 * nd, s));
            perm = dv.Current();
            launches_ += 4;
        }
        query_prep_kernel<<<nblk(qpad, 256), 256, 0, s>>>(d_q, nq, center_.p, D_, c.qp_.p, c.qaux_.p, qpad, c.counters_.p,
                                                          perm);
        launches_++;
        ScanParams P;
        P.filt = filt_.p;
        P.qp = c.qp_.p;
        P.qpad = qpad;
        P.init_thr = nullptr;
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev0_, s));
        // Starting bounds (knn_scan.cuh): a strided sample of the points gives every query a
        // value T with at least L sample points below it; the full pass starts its lists at T
        // and skips most of the early insertions.  Self-verifying: the re-rank's band check
        // does not care where a list's threshold came from.
        const uint32_t nsample = 4096;
        if (sample_thr_ && !hd && !tc && nslots_ >= 16 * (size_t)nsample && L <= 64)
        {
            const int rw = rowf_ / 4;
            c.sample_thr_buf_.reserve(qpad + (size_t)nsample * rowf_);
            float4 *rows = reinterpret_cast<float4 *>(c.sample_thr_buf_.p + qpad);  // qpad is a multiple of 256
            const uint32_t step = (uint32_t)(nslots_ / nsample);
            sample_gather_kernel<<<nblk((size_t)nsample * rw, 256), 256, 0, s>>>(
                reinterpret_cast<const float4 *>(filt_.p), rw, step, nsample, rows);
            launch_thr_sample(D_, rows, nsample, c.qp_.p, qpad, (uint32_t)L, c.sample_thr_buf_.p, s);
            CUDA_CHECK(cudaGetLastError());
            launches_ += 2;
            P.init_thr = c.sample_thr_buf_.p;
        }
        P.cand_thr = c.cand_vals_.p;
        P.cand_ids = c.cand_ids_.p;
        P.units = U;
        P.total_units = TU;
        P.smax = smax;
        if (tc)
        {
            {
                // shared operand image: rebuilt by the first query after a mutation, finished before any
                // other query (on another stream) may read it
                std::lock_guard<std::mutex> pl(prep_mu_);
                if (tc_version_ != mut_version_)
                {
                    tc_pimg_.reserve((size_t)U * tc_sh.nkc * tc_sh.chunk_elems());
                    tc_build_point_image(filt_.p, rowf_, D_, (size_t)U, tc_pimg_.p, s);
                    CUDA_CHECK(cudaStreamSynchronize(s));
                    tc_version_ = mut_version_;
                    launches_++;
                }
            }
            c.tc_qimg_.reserve((size_t)(qpad / TC_QT) * tc_sh.nkc * tc_sh.chunk_elems());
            tc_build_query_image(c.qp_.p, qpad, D_, c.tc_qimg_.p, s);
            launches_++;
            TcScanParams T{};
            T.pimg = tc_pimg_.p;
            T.qimg = c.tc_qimg_.p;
            T.cand_thr = c.cand_vals_.p;
            T.cand_ids = c.cand_ids_.p;
            T.units = U;
            T.total_units = TU;
            T.qpad = qpad;
            T.smax = smax;
            T.sh = tc_sh;
            T.nstages = tc_stages;
            T.diag = tc_variant_;
            T.dbg = tc_dbg_;
            T.ts = use_ts ? 1 : 0;
            T.pair = pair ? 1 : 0;
            // The operand image is re-read once per query super-tile by CTAs that drift apart: ask L2 to keep it
            // (persisting access-policy window on this stream) so the re-reads do not go back to DRAM.
            bool window = false;
            if (l2_window_ && qtiles > 1)
            {
                static std::once_flag lim[64];
                std::call_once(lim[device_ & 63], [] {
                    int dev = 0, maxp = 0;
                    cudaGetDevice(&dev);
                    cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, dev);
                    if (maxp > 0)
                        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)maxp);
                    cudaGetLastError();
                });
                int maxw = 0, maxp = 0;
                cudaDeviceGetAttribute(&maxw, cudaDevAttrMaxAccessPolicyWindowSize, device_);
                cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, device_);
                const size_t bytes = (size_t)U * tc_sh.nkc * tc_sh.chunk_elems() * sizeof(uint16_t);
                if (maxw > 0 && maxp > 0)
                {
                    cudaStreamAttrValue av{};
                    av.accessPolicyWindow.base_ptr = tc_pimg_.p;
                    av.accessPolicyWindow.num_bytes = std::min(bytes, (size_t)maxw);
                    av.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)maxp / (double)std::max<size_t>(bytes, 1));
                    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                    window = cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess;
                    cudaGetLastError();
                }
            }
            tc_launch_scan(L, T, (unsigned)(pair ? 2 * G : G), s);
            if (window)
            {
                cudaStreamAttrValue av{};
                av.accessPolicyWindow.num_bytes = 0;
                cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &av);
                cudaGetLastError();
            }
        }
        else if (hd)
            hd_any(cfg, 1, D_, rowf_, &P, (unsigned)G, s, nullptr);
        else
            scan_any(cfg, D_, 1, &P, (unsigned)G, s, nullptr);
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev1_, s));

        if ((size_t)smax * L > 512)
            knn_rerank_select_kernel<<<(unsigned)nq, SEL_THREADS,
                                       (size_t)smax * L * 12 + (size_t)D_ * 8 + 8 + (size_t)(SEL_THREADS / 32) * kk * 12, s>>>(
                c.cand_vals_.p, c.cand_ids_.p, smax, (uint32_t)L, qpad, pts64_.p, ids_.p, D_, d_q, c.qaux_.p, pmax2_.p,
                nq, (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists, c.ovf_list_.p, c.counters_.p, tc_err_units(D_), perm);
        else
        knn_rerank_kernel<512><<<nblk(nq, 4), 128, 0, s>>>(c.cand_vals_.p, c.cand_ids_.p, smax, (uint32_t)L, qpad,
                                                          pts64_.p, ids_.p, D_, d_q, c.qaux_.p, pmax2_.p, nq,
                                                          (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists,
                                                          c.ovf_list_.p, c.counters_.p,
                                                          tc ? tc_err_units(D_) : (double)(D_ + 3), perm);
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        unsigned novf = 0;
        const bool ho = c.ho_ids_ && !no_tc && c.ho_entries_ == nq * kstride && kk == kstride;
        if (ho)
        {
            // small host-pointer call: the results travel with the fallback counter (pinned memory)
            CUDA_CHECK(cudaMemcpyAsync(c.ho_ids_, d_ids, c.ho_entries_ * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(c.ho_dists_, d_dists, c.ho_entries_ * sizeof(double), cudaMemcpyDeviceToHost, s));
        }
        CUDA_CHECK(cudaMemcpyAsync(&novf, c.counters_.p, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        c.ho_done_ = ho && novf == 0;
        c.last_fallbacks_ = novf;
        if (profile_)
        {
            float ms = 0.f;
            CUDA_CHECK(cudaEventElapsedTime(&ms, c.ev0_, c.ev1_));
            c.last_scan_ns_ = (long)(ms * 1e6);
        }
        if (!no_tc)
            c.last_rescans_ = 0;
        if (tc && tc_mode_ < 0)
        {
            // "auto" keeps score: once a quarter of the recent queries came back unproven the set is too dense for the
            // BF16 split's band (2-D / 3-D roadmaps at 1M states) and the FP32-pipe scan becomes the first filter
            const long seen = tc_seen_.fetch_add((long)nq) + (long)nq, unp = tc_unproven_.fetch_add((long)novf) + (long)novf;
            if (seen >= 64)
            {
                if (unp * 4 > seen)
                {
                    tc_demote_slots_.store(nslots_, std::memory_order_relaxed);
                    tc_demoted_.store(true, std::memory_order_relaxed);
                }
                tc_seen_.store(0);
                tc_unproven_.store(0);
            }
        }
        if (novf >= RESCAN_MIN && tc && tc_mode_ < 0)
        {
            // Many unproven queries: a dense low-dimensional set, where the k-th distance is inside the BF16 split's
            // error band (800 u against the FP32 scan's (d + 3) u; 17 % of the queries at 1M x 3-D).  Give them to
            // the FP32-pipe scan as one batch; only what that cannot prove either goes to the exact path.
            c.re_sel_.reserve(novf);
            c.re_q_.reserve((size_t)novf * D_);
            c.re_ids_.reserve((size_t)novf * kk);
            c.re_d_.reserve((size_t)novf * kk);
            CUDA_CHECK(cudaMemcpyAsync(c.re_sel_.p, c.ovf_list_.p, novf * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
            gather_rows_kernel<<<nblk((size_t)novf * D_, 256), 256, 0, s>>>(d_q, c.re_sel_.p, novf, D_, c.re_q_.p);
            launches_++;
            const long scan_ns = c.last_scan_ns_;
            knn_fast(c, c.re_q_.p, novf, kk, kk, c.re_ids_.p, c.re_d_.p, s, true);
            scatter_results_kernel<<<nblk((size_t)novf * kk, 256), 256, 0, s>>>(c.re_ids_.p, c.re_d_.p, c.re_sel_.p, novf,
                                                                                 (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists);
            launches_++;
            CUDA_CHECK(cudaGetLastError());
            c.last_tc_ = 1;
            c.last_rescans_ = novf;          // c.last_fallbacks_: what the second pass sent to the exact path
            c.last_scan_ns_ += scan_ns;
        }
        else if (novf)
            knn_exact(c, d_q, c.ovf_list_.p, novf, kk, kstride, d_ids, d_dists, s);
    }

    namespace
    {
        // Smallest r such that, when the bound is the r-th smallest scan value of a sample holding each
        // point with probability p, fewer than k + 3 points of the whole set fall below it with
        // probability < 1e-5 (the count is r + NegBin(r, p)).  Such a query is not wrong, only redone exactly.
        uint32_t thr_pick_rank(size_t k, double p)
        {
            const double q = 1.0 - p;
            for (uint32_t r = 1; r < 4096; ++r)
            {
                if (r >= k + 3)
                    return r;
                const uint32_t jmax = (uint32_t)(k + 2 - r);  // P(NegBin <= jmax)
                double term = std::pow(p, (double)r), cdf = term;
                for (uint32_t j = 1; j <= jmax; ++j)
                {
                    term *= q * (double)(j + r - 1) / (double)j;
                    cdf += term;
                }
                if (cdf < 1e-5)
                    return r;
            }
            return 4096;
        }
    }  // namespace

    // Threshold path of the tensor-core scan (knn_scan_tc.cu, knn_thr_scan_tc_kernel), for an index
    // large enough that a sample pass pays:
    //   1. list scan (L = 8 / 16 / 24) over the sample image: every step-th point, 1/step of the pairs
    //   2. tau_select: the r-th smallest sample value per query is the query's bound tau
    //   3. threshold scan over all points: slots with a scan value below tau are appended (~ r * step per query)
    //   4. exact epilogue: FP64 distance of every kept slot, k best by (distance, slot), band check
    //      against tau; unprovable queries (row overflow, fewer than k kept, band) go to the exact path
    // Nothing here converges, so the cost per (query, segment) is the appends, not a list warm-up:
    // the time is linear in the points a rank owns, which is what makes a sharded roadmap scale.
    void NNIndex::knn_fast_thr(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids,
                               double *d_dists, cudaStream_t s)
    {
        TcShape sh = tc_shape(D_);
        if (tc_na_ > 0 && tc_na_ < sh.na)
            sh.na = tc_na_;
        const int stages = tc_thr_pick_stages(sh);
        REQUIRE(stages > 0, "threshold scan: shape does not fit shared memory");
        const size_t QT = (size_t)TC_QT * sh.na;
        const size_t qtiles = (nq + QT - 1) / QT;
        const uint32_t qpad = (uint32_t)(qtiles * QT);
        auto segments = [&](size_t qtiles, unsigned long long U, unsigned long long G) {
            const unsigned long long TU = qtiles * U;
            uint32_t smax = 1;
            for (size_t t = 0; t < qtiles; ++t)
            {
                const unsigned long long x0 = t * U, x1 = (t + 1) * U - 1;
                const unsigned long long g0 = ((x0 + 1) * G + TU - 1) / TU - 1;
                const unsigned long long g1 = ((x1 + 1) * G + TU - 1) / TU - 1;
                smax = std::max<uint32_t>(smax, (uint32_t)(g1 - g0 + 1));
            }
            return smax;
        };
        // ---- sample ----
        const size_t step = (size_t)thr_step_;
        const size_t srows = (nslots_ + step - 1) / step;
        const unsigned long long Us = (srows + TC_PT - 1) / TC_PT;
        const uint32_t r = thr_pick_rank(kk, 1.0 / (double)step);
        REQUIRE(r <= 48, "threshold scan: k too large for the sample lists");
        const int Ls = r <= 8 ? 8 : (r <= 16 ? 16 : (r <= 24 ? 24 : 48));
        TcShape shs = sh;
        int stages_s = tc_pick_stages(shs, Ls);
        while (stages_s < 3 && shs.na > 1)
        {
            shs.na /= 2;
            stages_s = tc_pick_stages(shs, Ls);
        }
        REQUIRE(stages_s > 0, "threshold scan: sample lists do not fit shared memory");
        const size_t qtiles_s = qpad / ((size_t)TC_QT * shs.na);  // long lists: fewer resident query tiles
        const unsigned long long TUs = qtiles_s * Us;
        const unsigned long long Gs = std::max<unsigned long long>(1, std::min<unsigned long long>(sm_count_, TUs));
        const uint32_t smax_s = segments(qtiles_s, Us, Gs);
        // ---- main ----
        const unsigned long long U = (nslots_ + TC_PT - 1) / TC_PT, TU = qtiles * U;
        const unsigned long long G = std::max<unsigned long long>(1, std::min<unsigned long long>(sm_count_, TU));
        const uint32_t smax = segments(qtiles, U, G);
        // expected kept slots per query r * step, standard deviation sqrt(r (1 - p)) * step
        const double mean = (double)r * (double)step, sd = std::sqrt((double)r) * (double)step;
        constexpr int MAXC = 1024;
        uint32_t cap = (uint32_t)(mean + 6.0 * sd) + 64;
        cap = std::min<uint32_t>(MAXC, ((cap + 31) / 32) * 32);
        c.last_rank_ = r;

        c.qp_.reserve((size_t)D_ * qpad);
        c.qaux_.reserve(2 * (size_t)qpad);
        c.cand_vals_.reserve((size_t)smax_s * qpad);
        c.cand_ids_.reserve((size_t)smax_s * Ls * qpad);
        c.list_vals_.reserve((size_t)smax_s * Ls * qpad);
        c.tau_.reserve(qpad);
        c.app_cnt_.reserve((size_t)smax * qpad);
        c.app_ids_.reserve((size_t)smax * qpad * cap);
        c.ovf_list_.reserve(nq);
        CUDA_CHECK(cudaMemsetAsync(c.list_vals_.p, 0x7F, (size_t)smax_s * Ls * qpad * sizeof(float), s));  // "empty"
        CUDA_CHECK(cudaMemsetAsync(c.app_cnt_.p, 0, (size_t)smax * qpad * sizeof(uint32_t), s));
        CUDA_CHECK(cudaMemsetAsync(c.counters_.p, 0, sizeof(unsigned), s));
        CUDA_CHECK(cudaMemsetAsync(c.kept_.p, 0, sizeof(unsigned long long), s));
        query_prep_kernel<<<nblk(qpad, 256), 256, 0, s>>>(d_q, nq, center_.p, D_, c.qp_.p, c.qaux_.p, qpad);
        launches_++;
        {
            std::lock_guard<std::mutex> pl(prep_mu_);  // shared images, see knn_fast
            bool built = false;
            if (tc_version_ != mut_version_)
            {
                tc_pimg_.reserve((size_t)U * sh.nkc * sh.chunk_elems());
                tc_build_point_image(filt_.p, rowf_, D_, (size_t)U, tc_pimg_.p, s);
                built = true;
                launches_++;
            }
            if (tc_sversion_ != mut_version_ || tc_sstep_ != step)
            {
                tc_simg_.reserve((size_t)Us * sh.nkc * sh.chunk_elems());
                tc_build_sample_image(filt_.p, rowf_, D_, nslots_, step, (size_t)Us, tc_simg_.p, s);
                built = true;
                launches_++;
            }
            if (built)
                CUDA_CHECK(cudaStreamSynchronize(s));
            tc_version_ = mut_version_;
            tc_sversion_ = mut_version_;
            tc_sstep_ = step;
        }
        c.tc_qimg_.reserve((size_t)(qpad / TC_QT) * sh.nkc * sh.chunk_elems());
        tc_build_query_image(c.qp_.p, qpad, D_, c.tc_qimg_.p, s);
        launches_++;
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev0_, s));
        TcScanParams T{};
        T.qimg = c.tc_qimg_.p;
        T.qpad = qpad;
        T.sh = sh;
        T.diag = tc_variant_;
        T.dbg = nullptr;
        // 1. sample lists
        T.pimg = tc_simg_.p;
        T.cand_thr = c.cand_vals_.p;
        T.cand_ids = c.cand_ids_.p;
        T.list_vals = c.list_vals_.p;
        T.units = Us;
        T.total_units = TUs;
        T.smax = smax_s;
        T.nstages = stages_s;
        T.diag = 0;
        T.sh = shs;
        tc_launch_scan(Ls, T, (unsigned)Gs, s);
        T.sh = sh;
        // 2. bounds
        if ((size_t)smax_s * Ls > 64)
            tau_select_kernel<32><<<nblk((size_t)qpad * 32, 128), 128, 0, s>>>(c.list_vals_.p, smax_s * Ls, qpad, r, c.tau_.p);
        else
            tau_select_kernel<1><<<nblk(qpad, 128), 128, 0, s>>>(c.list_vals_.p, smax_s * Ls, qpad, r, c.tau_.p);
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev2_, s));
        // 3. threshold scan
        T.pimg = tc_pimg_.p;
        T.list_vals = nullptr;
        T.tau = c.tau_.p;
        T.app_ids = c.app_ids_.p;
        T.app_cnt = c.app_cnt_.p;
        T.cap = cap;
        T.units = U;
        T.total_units = TU;
        T.smax = smax;
        T.nstages = stages;
        T.diag = tc_variant_;
        T.dbg = tc_dbg_;
        tc_launch_thr_scan(T, (unsigned)G, s);
        launches_ += 3;
        CUDA_CHECK(cudaGetLastError());
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev1_, s));
        // 4. exact epilogue
        const size_t rr_smem = (size_t)48 * MAXC + 4096;
        static std::once_flag rr_once[64];
        std::call_once(rr_once[device_ & 63], [&] {
            CUDA_CHECK(cudaFuncSetAttribute(knn_rerank_thr_kernel<MAXC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)rr_smem));
        });
        knn_rerank_thr_kernel<MAXC><<<nblk(nq, 4), 128, rr_smem, s>>>(
            c.app_ids_.p, c.app_cnt_.p, smax, cap, qpad, c.tau_.p, pts64_.p, ids_.p, D_, d_q, c.qaux_.p, pmax2_.p, nq,
            (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists, c.ovf_list_.p, c.counters_.p, tc_err_units(D_),
            profile_ ? c.kept_.p : nullptr);
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        if (profile_)
            CUDA_CHECK(cudaEventRecord(c.ev3_, s));
        unsigned novf = 0;
        CUDA_CHECK(cudaMemcpyAsync(&novf, c.counters_.p, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        c.last_fallbacks_ = novf;
        c.last_thr_ = 1;
        if (profile_)
        {
            float ms = 0.f;
            CUDA_CHECK(cudaEventElapsedTime(&ms, c.ev0_, c.ev1_));
            c.last_scan_ns_ = (long)(ms * 1e6);
            CUDA_CHECK(cudaEventElapsedTime(&ms, c.ev0_, c.ev2_));
            c.last_sample_ns_ = (long)(ms * 1e6);
            CUDA_CHECK(cudaEventElapsedTime(&ms, c.ev2_, c.ev1_));
            c.last_main_ns_ = (long)(ms * 1e6);
            unsigned long long kept = 0;
            CUDA_CHECK(cudaMemcpy(&kept, c.kept_.p, sizeof(kept), cudaMemcpyDeviceToHost));
            c.last_kept_ = (long)kept;
        }
        if (novf)
            knn_exact(c, d_q, c.ovf_list_.p, novf, kk, kstride, d_ids, d_dists, s);
    }

    bool NNIndex::filter_usable() const
    {
        return (fast_space_ || embed_) && !force_exact_ && D_ <= TC_MAX_DIM && tc_mode_ != 0 && nslots_ >= 4096 &&
               tc_thr_pick_stages(tc_shape(D_)) > 0;
    }

    // Embedded spaces: has any stored state broken the embedding's assumptions (embed_state)?  Read once per
    // mutation; the flag only ever goes up between compactions.
    bool NNIndex::embed_ok(cudaStream_t s)
    {
        if (!embed_)
            return true;
        std::lock_guard<std::mutex> pl(prep_mu_);
        if (embed_ver_ != mut_version_)
        {
            unsigned bad = 0;
            CUDA_CHECK(cudaMemcpyAsync(&bad, embed_bad_.p, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaStreamSynchronize(s));
            embed_okay_ = bad == 0;
            embed_ver_ = mut_version_;
        }
        return embed_okay_;
    }

    // See nn.cuh.  Returns smax (rows per query); *qpad_out = the row pitch of app_cnt_ / app_ids_.
    uint32_t NNIndex::filter_scan(QCtx &c, const double *d_q, size_t nq, const double *d_bound, uint32_t cap,
                                  uint32_t *qpad_out, cudaStream_t s)
    {
        TcShape sh = tc_shape(D_);
        if (tc_na_ > 0 && tc_na_ < sh.na)
            sh.na = tc_na_;
        const int stages = tc_thr_pick_stages(sh);
        REQUIRE(stages > 0, "filter scan: shape does not fit shared memory");
        const size_t QT = (size_t)TC_QT * sh.na;
        const size_t qtiles = (nq + QT - 1) / QT;
        const uint32_t qpad = (uint32_t)(qtiles * QT);
        const unsigned long long U = (nslots_ + TC_PT - 1) / TC_PT, TU = qtiles * U;
        const unsigned long long G = std::max<unsigned long long>(1, std::min<unsigned long long>(sm_count_, TU));
        uint32_t smax = 1;
        for (size_t t = 0; t < qtiles; ++t)
        {
            const unsigned long long x0 = t * U, x1 = (t + 1) * U - 1;
            const unsigned long long g0 = ((x0 + 1) * G + TU - 1) / TU - 1;
            const unsigned long long g1 = ((x1 + 1) * G + TU - 1) / TU - 1;
            smax = std::max<uint32_t>(smax, (uint32_t)(g1 - g0 + 1));
        }
        c.qp_.reserve((size_t)D_ * qpad);
        c.qaux_.reserve(2 * (size_t)qpad);
        c.tau_.reserve(qpad);
        c.app_cnt_.reserve((size_t)smax * qpad);
        c.app_ids_.reserve((size_t)smax * qpad * cap);
        CUDA_CHECK(cudaMemsetAsync(c.app_cnt_.p, 0, (size_t)smax * qpad * sizeof(uint32_t), s));
        query_prep_kernel<<<nblk(qpad, 256), 256, 0, s>>>(d_q, nq, center_.p, D_, c.qp_.p, c.qaux_.p, qpad);
        scan_bound_kernel<<<nblk(qpad, 256), 256, 0, s>>>(d_bound, nq, qpad, c.qaux_.p, pmax2_.p, D_,
                                                          tc_err_units(D_), c.tau_.p);
        launches_ += 2;
        {
            std::lock_guard<std::mutex> pl(prep_mu_);  // shared operand image, see knn_fast
            if (tc_version_ != mut_version_)
            {
                tc_pimg_.reserve((size_t)U * sh.nkc * sh.chunk_elems());
                tc_build_point_image(filt_.p, rowf_, D_, (size_t)U, tc_pimg_.p, s);
                CUDA_CHECK(cudaStreamSynchronize(s));
                tc_version_ = mut_version_;
                launches_++;
            }
        }
        c.tc_qimg_.reserve((size_t)(qpad / TC_QT) * sh.nkc * sh.chunk_elems());
        tc_build_query_image(c.qp_.p, qpad, D_, c.tc_qimg_.p, s);
        TcScanParams T{};
        T.pimg = tc_pimg_.p;
        T.qimg = c.tc_qimg_.p;
        T.qpad = qpad;
        T.sh = sh;
        T.tau = c.tau_.p;
        T.app_ids = c.app_ids_.p;
        T.app_cnt = c.app_cnt_.p;
        T.cap = cap;
        T.units = U;
        T.total_units = TU;
        T.smax = smax;
        T.nstages = stages;
        T.diag = 0;
        T.dbg = nullptr;
        tc_launch_thr_scan(T, (unsigned)G, s);
        launches_ += 2;
        CUDA_CHECK(cudaGetLastError());
        *qpad_out = qpad;
        return smax;
    }

    // Embedded spaces: tensor-core LIST scan of the embedded queries against the embedded roadmap (the lists of the
    // fast path, no re-rank), then cand_bound_kernel: d_bound[b] = kk-th smallest exact distance among the listed slots.
    // Queries the lists alone prove (cand_bound_kernel) have their rows of d_ids / d_dists written; returns how many
    // of the batch are left for the threshold scan.
    unsigned NNIndex::embed_list_bound(QCtx &c, const double *d_qe, const void *pair_metric, size_t q0, size_t nq, size_t kk,
                                       double *d_bound, size_t kstride, uint64_t *d_ids, double *d_dists, cudaStream_t s)
    {
        const PairMetric &pm = *static_cast<const PairMetric *>(pair_metric);
        TcShape sh = tc_shape(D_);
        const int L = kk <= 10 ? 16 : (kk <= 14 ? 24 : 48);
        int stages = tc_pick_stages(sh, L);
        while (stages < 3 && sh.na > 1)
        {
            sh.na /= 2;
            stages = tc_pick_stages(sh, L);
        }
        while (sh.na > 1 && nq <= (size_t)TC_QT * (sh.na / 2))
        {
            sh.na /= 2;
            stages = tc_pick_stages(sh, L);
        }
        REQUIRE(stages > 0, "embedded list scan: shape does not fit shared memory");
        const size_t QT = (size_t)TC_QT * sh.na;
        const size_t qtiles = (nq + QT - 1) / QT;
        const uint32_t qpad = (uint32_t)(qtiles * QT);
        const unsigned max_seg = 512u / L;
        unsigned long long U = 0, TU = 0, G = 1;
        uint32_t smax = 1;
        auto plan = [&](size_t rows) {
            U = (rows + TC_PT - 1) / TC_PT;
            TU = qtiles * U;
            G = std::min<unsigned long long>(sm_count_, TU);
            G = std::min<unsigned long long>(G, qtiles * (unsigned long long)(max_seg - 1));
            G = std::max<unsigned long long>(G, 1);
            smax = 1;
            for (size_t t = 0; t < qtiles; ++t)
            {
                const unsigned long long x0 = t * U, x1 = (t + 1) * U - 1;
                const unsigned long long g0 = ((x0 + 1) * G + TU - 1) / TU - 1;
                const unsigned long long g1 = ((x1 + 1) * G + TU - 1) / TU - 1;
                smax = std::max<uint32_t>(smax, (uint32_t)(g1 - g0 + 1));
            }
        };
        plan(nslots_);
        // (A bound from the image of every 8th state was tried for the big batches, where the lists prove next to
        // nothing and this scan is only there for the bound: the threshold scan then kept more than its rows hold for
        // many queries -- 13 of 65536 at k = 10, three quarters at k = 36 on 1M SE(3) states -- and every such query costs
        // an exact pass.  The plumbing for a sampled image stays; step is 1.)
        const size_t step = 1;
        const size_t rows = (nslots_ + step - 1) / step;
        if (step > 1)
            plan(rows);
        REQUIRE((size_t)smax * L <= 512, "embedded list scan: too many segments per query tile");
        c.qp_.reserve((size_t)D_ * qpad);
        c.qaux_.reserve(2 * (size_t)qpad);
        c.cand_vals_.reserve((size_t)smax * qpad);
        c.cand_ids_.reserve((size_t)smax * L * qpad);
        CUDA_CHECK(cudaMemsetAsync(c.cand_ids_.p, 0xFF, (size_t)smax * L * qpad * sizeof(uint32_t), s));
        CUDA_CHECK(cudaMemsetAsync(c.cand_vals_.p, 0x7F, (size_t)smax * qpad * sizeof(float), s));
        query_prep_kernel<<<nblk(qpad, 256), 256, 0, s>>>(d_qe, nq, center_.p, D_, c.qp_.p, c.qaux_.p, qpad);
        {
            std::lock_guard<std::mutex> pl(prep_mu_);  // shared operand images, see knn_fast
            if (step > 1)
            {
                if (tc_sversion_ != mut_version_ || tc_sstep_ != step)
                {
                    tc_simg_.reserve((size_t)U * sh.nkc * sh.chunk_elems());
                    tc_build_sample_image(filt_.p, rowf_, D_, nslots_, step, (size_t)U, tc_simg_.p, s);
                    CUDA_CHECK(cudaStreamSynchronize(s));
                    tc_sversion_ = mut_version_;
                    tc_sstep_ = step;
                    launches_++;
                }
            }
            else if (tc_version_ != mut_version_)
            {
                tc_pimg_.reserve((size_t)U * sh.nkc * sh.chunk_elems());
                tc_build_point_image(filt_.p, rowf_, D_, (size_t)U, tc_pimg_.p, s);
                CUDA_CHECK(cudaStreamSynchronize(s));
                tc_version_ = mut_version_;
                launches_++;
            }
        }
        c.tc_qimg_.reserve((size_t)(qpad / TC_QT) * sh.nkc * sh.chunk_elems());
        tc_build_query_image(c.qp_.p, qpad, D_, c.tc_qimg_.p, s);
        TcScanParams T{};
        T.pimg = step > 1 ? tc_simg_.p : tc_pimg_.p;
        T.qimg = c.tc_qimg_.p;
        T.cand_thr = c.cand_vals_.p;
        T.cand_ids = c.cand_ids_.p;
        T.units = U;
        T.total_units = TU;
        T.qpad = qpad;
        T.smax = smax;
        T.sh = sh;
        T.nstages = stages;
        tc_launch_scan(L, T, (unsigned)G, s);
        c.unp_list_.reserve(nq);
        CUDA_CHECK(cudaMemsetAsync(c.counters_.p + 2, 0, sizeof(unsigned), s));
        cand_bound_kernel<<<nblk(nq, 4), 128, 0, s>>>(c.cand_ids_.p, c.cand_vals_.p, smax,
                                                      (uint32_t)L, qpad, pm, q0, alive_.p, nq, (uint32_t)kk, d_bound,
                                                      c.qaux_.p, pmax2_.p, D_, tc_err_units(D_), embed_slack_, ids_.p,
                                                      (uint32_t)kstride, d_ids, d_dists, c.counters_.p + 2, c.unp_list_.p,
                                                      (uint32_t)step, nslots_);
        launches_ += 4;
        CUDA_CHECK(cudaGetLastError());
        unsigned unproven = 0;
        CUDA_CHECK(cudaMemcpyAsync(&unproven, c.counters_.p + 2, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        return unproven;
    }

    // Large index (>= 32K slots), any k and any metric (FMT*'s k ~ 2100 of 100K; SE3 / chain kNN): sorting
    // every slot per query is the cost of the full path.  Here an exact distance bound per query comes from a strided sample (the R-th
    // smallest of m sample distances, R chosen 4 sigma above k*m/n), only slots within the bound are
    // kept (the same exact FP64 distances, any metric), and only those are sorted: by slot, then
    // stably by distance, so ties stay in insertion order.  A query whose bound caught fewer than k
    // or more than `cap` slots goes to the full sort.  Returns false if the sample cannot help.
    // Chain metric only: bring the per-slot trig table up to date and tabulate the queries.
    // Returns true if the exact-path kernels should use the tables.
    bool NNIndex::prepare_chain_tables(QCtx &c, const double *d_q, size_t nq, cudaStream_t s)
    {
        if (sp_.compound || sp_.ncomp != 1 || sp_.comp[0].type != OMPLCU_CHAIN)
            return false;
        const int d = sp_.comp[0].dim;
        {
            std::lock_guard<std::mutex> pl(prep_mu_);  // the per-slot table is shared between concurrent queries
            trig_.reserve(std::max<size_t>(nslots_, 1) * 2 * d, true, s);
            if (trig_rows_ < nslots_)
            {
                chain_trig_kernel<<<nblk(nslots_ - trig_rows_, 128), 128, 0, s>>>(pts64_.p + trig_rows_ * d,
                                                                                   nslots_ - trig_rows_, d,
                                                                                   trig_.p + trig_rows_ * 2 * d);
                CUDA_CHECK(cudaStreamSynchronize(s));
                trig_rows_ = nslots_;
                launches_++;
            }
        }
        c.qtrig_.reserve(std::max<size_t>(nq, 1) * 2 * d);
        chain_trig_kernel<<<nblk(nq, 128), 128, 0, s>>>(d_q, nq, d, c.qtrig_.p);
        launches_++;
        CUDA_CHECK(cudaGetLastError());
        return true;
    }

    bool NNIndex::knn_exact_collect(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                                    uint64_t *d_ids, double *d_dists, cudaStream_t s)
    {
        const uint32_t m = 1024;
        if (!collect_ || nlive_ == 0 || nslots_ >= ((size_t)1 << 31))
            return false;
        // embedded spaces, k <= 40: the bound comes from a list scan on the embedding (embed_list_bound) and is tight:
        // a few times kk slots lie within it, so the rows can be short; no sample, so it pays from 4096 slots on
        const bool tight = embed_ && kk <= 40 && d_sel == nullptr && nsel >= 16 && filter_usable() && embed_ok(s);
        if (!tight && nslots_ < 32 * (size_t)m)
            return false;
        const double mean = (double)kk / (double)nlive_ * m;
        const uint32_t R = (uint32_t)std::ceil(mean + 4.0 * std::sqrt(mean) + 6.0);
        if (!tight && R > m / 4)
            return false;
        const size_t step = std::max<size_t>(nslots_ / m, 1);
        const uint32_t cap = tight ? 1024u : (uint32_t)(1.6 * (double)R / m * (double)nslots_) + 256;
        if (cap < kk)
            return false;
        const size_t budget = (size_t)1 << 26;
        const size_t nb_max = std::max<size_t>(1, std::min(nsel, budget / cap));
        c.ex_keys_.reserve(nb_max * cap);
        c.ex_keys2_.reserve(nb_max * cap);
        c.ex_vals_.reserve(nb_max * cap);
        c.ex_vals2_.reserve(nb_max * cap);
        c.ex_offs_.reserve(2 * nb_max + 2);
        c.col_sd_.reserve(nb_max * m);
        c.col_bound_.reserve(nb_max);
        c.col_cnt_.reserve(nb_max);
        c.col_redo_.reserve(nsel);
        CUDA_CHECK(cudaMemsetAsync(c.counters_.p + 1, 0, sizeof(unsigned), s));
        // filter rows: room for every survivor of a query in each of its rows (a query's neighbours are spread
        // over the point ranges of its segments in no particular way)
        // embedded spaces: a handful of queries is cheaper unfiltered (one distance pass + one sort)
        const bool filt = d_sel == nullptr && filter_usable() && (!embed_ || nsel >= 16) && embed_ok(s);
        const uint32_t fcap = ((cap + 64 + 31) / 32) * 32;
        c.last_filter_ = filt ? 1 : 0;
        c.last_proven_ = 0;
        const bool tables = prepare_chain_tables(c, d_q, c.exact_nq_, s);
        const PairMetric pm{sp_, pts64_.p, d_q, tables ? trig_.p : nullptr, tables ? c.qtrig_.p : nullptr};
        int *beg = c.ex_offs_.p, *end = c.ex_offs_.p + nb_max + 1;
        for (size_t b0 = 0; b0 < nsel; b0 += nb_max)
        {
            const size_t nb = std::min(nb_max, nsel - b0);
            const size_t items = nb * cap;
            REQUIRE(items < ((size_t)1 << 31), "exact path: batch too large");
            CUDA_CHECK(cudaMemsetAsync(c.col_cnt_.p, 0, nb * sizeof(uint32_t), s));
            if (tight)
            {
                // embedded spaces, k <= 40: the bound comes from a list scan on the embedding instead of the sample
                c.q_emb_.reserve(nb * (size_t)D_);
                c.emb_bound_.reserve(nb);
                embed_queries_kernel<<<nblk(nb, 128), 128, 0, s>>>(sp_, d_q + b0 * sp_.stride, nb, D_, embed_slack_,
                                                                   c.col_bound_.p, c.q_emb_.p, c.emb_bound_.p);
                const unsigned left = embed_list_bound(c, c.q_emb_.p, &pm, b0, nb, kk, c.col_bound_.p, kstride, d_ids, d_dists, s);
                c.last_proven_ += (long)nb - (long)left;
                if (left == 0)
                    continue;  // many segments per query tile (batches of a few thousand): the list scan proves them all
                if (few_mode_ && left <= (unsigned)XFEW_MAXQ && kk <= (size_t)XFEW_MAXK && (size_t)left * nslots_ <= (size_t)5000000)
                {
                    // a handful left: the exact one-pass kernel on them is cheaper than a scan of the whole image
                    c.re_q_.reserve((size_t)left * sp_.stride);
                    c.re_ids_.reserve((size_t)left * kk);
                    c.re_d_.reserve((size_t)left * kk);
                    gather_rows_kernel<<<nblk((size_t)left * sp_.stride, 256), 256, 0, s>>>(d_q + b0 * sp_.stride, c.unp_list_.p,
                                                                                          left, sp_.stride, c.re_q_.p);
                    knn_exact_few(c, c.re_q_.p, left, kk, kk, c.re_ids_.p, c.re_d_.p, s);
                    scatter_results_kernel<<<nblk((size_t)left * kk, 256), 256, 0, s>>>(
                        c.re_ids_.p, c.re_d_.p, c.unp_list_.p, left, (uint32_t)kk, (uint32_t)kstride, d_ids + b0 * kstride,
                        d_dists ? d_dists + b0 * kstride : nullptr);
                    c.last_few_ = 0;
                    launches_ += 2;
                    CUDA_CHECK(cudaGetLastError());
                    continue;
                }
                // the threshold scan for the rest only (compacted; row i of the scan is query b0 + unp_list[i])
                c.q_emb2_.reserve((size_t)left * D_);
                c.emb_bound_.reserve(nb);
                c.col_bound2_.reserve(left);
                gather_unproven_kernel<<<nblk(left, 128), 128, 0, s>>>(c.unp_list_.p, left, D_, c.q_emb_.p, c.col_bound_.p,
                                                                      embed_slack_, c.q_emb2_.p, c.emb_bound_.p,
                                                                      c.col_bound2_.p);
                uint32_t fq = 0;
                const uint32_t rows = filter_scan(c, c.q_emb2_.p, left, c.emb_bound_.p, fcap, &fq, s);
                kept_collect_metric_kernel<<<nblk(left, 4), 128, 0, s>>>(c.app_ids_.p, c.app_cnt_.p, rows, fcap, fq, pm, b0,
                                                                         alive_.p, c.col_bound2_.p, left, cap, c.ex_keys2_.p,
                                                                         c.ex_vals2_.p, c.col_cnt_.p, c.unp_list_.p);
                collect_rank_emit_kernel<<<nblk(left, 4), 128, (size_t)4 * cap * 12, s>>>(c.ex_keys2_.p, c.ex_vals2_.p, c.col_cnt_.p, cap,
                                                                            ids_.p, b0, left, (uint32_t)kk, (uint32_t)kstride,
                                                                            d_ids, d_dists, c.col_redo_.p, c.counters_.p + 1,
                                                                            c.unp_list_.p);
                CUDA_CHECK(cudaGetLastError());
                launches_ += 10;
                continue;
            }
            else
            {
            sample_dist_kernel<<<nblk(nb * m, 256), 256, 0, s>>>(pm, alive_.p, step, m, d_sel, b0, nb, c.col_sd_.p);
            rank_select_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.col_sd_.p, m, nb, R, c.col_bound_.p);
            }
            if (filt)
            {
                // Plain RealVector space: the tensor-core threshold scan keeps every slot whose scan value is
                // below the scan-value image of the query's exact distance bound (a superset of the slots within
                // the bound); only those get an FP64 distance.  Everything downstream is unchanged: the same
                // exact distances, the same bound, the same redo rule.  The survivors arrive in slot order, so
                // one stable sort by distance gives the (distance, slot) order.
                uint32_t fq = 0;
                if (embed_)
                {
                    // SO(2) / SO(3) / compound: the filter runs on the Euclidean lower-bound embedding (embed_state),
                    // the kept slots get the space's own exact distance
                    c.q_emb_.reserve(nb * (size_t)D_);
                    c.emb_bound_.reserve(nb);
                    embed_queries_kernel<<<nblk(nb, 128), 128, 0, s>>>(sp_, d_q + b0 * sp_.stride, nb, D_, embed_slack_,
                                                                       c.col_bound_.p, c.q_emb_.p, c.emb_bound_.p);
                    const uint32_t rows = filter_scan(c, c.q_emb_.p, nb, c.emb_bound_.p, fcap, &fq, s);
                    kept_collect_metric_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.app_ids_.p, c.app_cnt_.p, rows, fcap, fq, pm,
                                                                           b0, alive_.p, c.col_bound_.p, nb, cap,
                                                                           c.ex_keys2_.p, c.ex_vals2_.p, c.col_cnt_.p);
                    if (tight)
                    {
                        collect_rank_emit_kernel<<<nblk(nb, 4), 128, (size_t)4 * cap * 12, s>>>(
                            c.ex_keys2_.p, c.ex_vals2_.p, c.col_cnt_.p, cap, ids_.p, b0, nb, (uint32_t)kk, (uint32_t)kstride,
                            d_ids, d_dists, c.col_redo_.p, c.counters_.p + 1);
                        CUDA_CHECK(cudaGetLastError());
                        launches_ += 8;
                        continue;
                    }
                }
                else
                {
                const uint32_t rows = filter_scan(c, d_q + b0 * sp_.stride, nb, c.col_bound_.p, fcap, &fq, s);
                kept_collect_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.app_ids_.p, c.app_cnt_.p, rows, fcap, fq, pts64_.p,
                                                                alive_.p, D_, d_q + b0 * sp_.stride, c.col_bound_.p,
                                                                nb, cap, c.ex_keys2_.p, c.ex_vals2_.p, c.col_cnt_.p);
                }
                collect_segments_kernel<<<nblk(nb, 256), 256, 0, s>>>(c.col_cnt_.p, cap, nb, beg, end);
                CUDA_CHECK(cudaGetLastError());
                size_t tmp2 = 0;
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp2, c.ex_keys2_.p, c.ex_keys_.p,
                                                                    c.ex_vals2_.p, c.ex_vals_.p, (int)items, (int)nb,
                                                                    beg, end, 0, 64, s));
                c.cub_tmp_.reserve(tmp2);
                tmp2 = c.cub_tmp_.cap;
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp2, c.ex_keys2_.p, c.ex_keys_.p,
                                                                    c.ex_vals2_.p, c.ex_vals_.p, (int)items, (int)nb,
                                                                    beg, end, 0, 64, s));
            }
            else
            {
            collect_kernel<<<nblk(nb * nslots_, 256), 256, 0, s>>>(pm, alive_.p, nslots_, d_sel, b0, nb, c.col_bound_.p,
                                                                   cap, c.ex_keys_.p, c.ex_vals_.p, c.col_cnt_.p);
            collect_segments_kernel<<<nblk(nb, 256), 256, 0, s>>>(c.col_cnt_.p, cap, nb, beg, end);
            CUDA_CHECK(cudaGetLastError());
            size_t tmp = 0, tmp2 = 0;
            // pass 1: by slot (payload: distance); pass 2: stable by distance (payload: slot)
            CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp, c.ex_vals_.p, c.ex_vals2_.p, c.ex_keys_.p,
                                                                c.ex_keys2_.p, (int)items, (int)nb, beg, end, 0, 32, s));
            CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp2, c.ex_keys2_.p, c.ex_keys_.p, c.ex_vals2_.p,
                                                                c.ex_vals_.p, (int)items, (int)nb, beg, end, 0, 64, s));
            c.cub_tmp_.reserve(std::max(tmp, tmp2));
            tmp = tmp2 = c.cub_tmp_.cap;
            CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp, c.ex_vals_.p, c.ex_vals2_.p, c.ex_keys_.p,
                                                                c.ex_keys2_.p, (int)items, (int)nb, beg, end, 0, 32, s));
            CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp2, c.ex_keys2_.p, c.ex_keys_.p,
                                                                c.ex_vals2_.p, c.ex_vals_.p, (int)items, (int)nb, beg, end,
                                                                0, 64, s));
            }
            collect_emit_kernel<<<nblk(nb * kk, 256), 256, 0, s>>>(c.ex_keys_.p, c.ex_vals_.p, c.col_cnt_.p, cap, ids_.p,
                                                                   d_sel, b0, nb, (uint32_t)kk, (uint32_t)kstride,
                                                                   d_ids, d_dists, c.col_redo_.p, c.counters_.p + 1);
            CUDA_CHECK(cudaGetLastError());
            launches_ += 16;
        }
        unsigned nredo = 0;
        CUDA_CHECK(cudaMemcpyAsync(&nredo, c.counters_.p + 1, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        c.last_collect_redo_ = nredo;
        if (nredo)
            knn_exact_sort(c, d_q, c.col_redo_.p, nredo, kk, kstride, d_ids, d_dists, s);
        return true;
    }

    // Up to 8 queries, k <= 32, any space: exact_few_kernel + exact_few_select_kernel (one pass, no sort of all slots).
    void NNIndex::knn_exact_few(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids,
                                double *d_dists, cudaStream_t s)
    {
        const uint32_t qpad = XFEW_MAXQ;
        const unsigned grid = (unsigned)std::max<size_t>(
            1, std::min<size_t>({(size_t)sm_count_, (size_t)8192 / (XFEW_WARPS * kk), (nslots_ + 127) / 128}));
        const uint32_t n = grid * XFEW_WARPS * (uint32_t)kk;
        c.xf_d_.reserve((size_t)n * qpad);
        c.xf_s_.reserve((size_t)n * qpad);
        const bool tables = prepare_chain_tables(c, d_q, nq, s);
        const PairMetric pm{sp_, pts64_.p, d_q, tables ? trig_.p : nullptr, tables ? c.qtrig_.p : nullptr};
        if (nq == 1)
            exact_few_kernel<1><<<grid, XFEW_WARPS * 32, 0, s>>>(pm, alive_.p, nslots_, (uint32_t)nq, (uint32_t)kk, qpad,
                                                                  c.xf_d_.p, c.xf_s_.p);
        else if (nq <= 4)
            exact_few_kernel<4><<<grid, XFEW_WARPS * 32, 0, s>>>(pm, alive_.p, nslots_, (uint32_t)nq, (uint32_t)kk, qpad,
                                                                  c.xf_d_.p, c.xf_s_.p);
        else
            exact_few_kernel<XFEW_MAXQ><<<grid, XFEW_WARPS * 32, 0, s>>>(pm, alive_.p, nslots_, (uint32_t)nq, (uint32_t)kk,
                                                                          qpad, c.xf_d_.p, c.xf_s_.p);
        const size_t smem = (size_t)n * 12 + (size_t)(SEL_THREADS / 32) * kk * 12 + 16;
        CUDA_CHECK(cudaFuncSetAttribute(exact_few_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        exact_few_select_kernel<<<(unsigned)nq, SEL_THREADS, smem, s>>>(c.xf_d_.p, c.xf_s_.p, n, qpad, ids_.p, (uint32_t)kk,
                                                                        (uint32_t)kstride, d_ids, d_dists);
        launches_ += 2;
        CUDA_CHECK(cudaGetLastError());
        c.last_few_ = 2;
    }

    void NNIndex::knn_exact(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                            uint64_t *d_ids, double *d_dists, cudaStream_t s)
    {
        if (nsel == 0 || nslots_ == 0)
            return;
        if (knn_exact_collect(c, d_q, d_sel, nsel, kk, kstride, d_ids, d_dists, s))
            return;
        knn_exact_sort(c, d_q, d_sel, nsel, kk, kstride, d_ids, d_dists, s);
    }

    void NNIndex::knn_exact_sort(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                                 uint64_t *d_ids, double *d_dists, cudaStream_t s)
    {
        if (nsel == 0 || nslots_ == 0)
            return;
        REQUIRE(nslots_ < ((size_t)1 << 31), "exact path: too many slots");
        size_t nb_max = std::max<size_t>(1, std::min(nsel, EXACT_PAIR_BUDGET / nslots_));
        c.ex_keys_.reserve(nb_max * nslots_);
        c.ex_keys2_.reserve(nb_max * nslots_);
        c.ex_vals_.reserve(nb_max * nslots_);
        c.ex_vals2_.reserve(nb_max * nslots_);
        c.ex_offs_.reserve(nb_max + 1);
        std::vector<int> offs(nb_max + 1);
        for (size_t b = 0; b <= nb_max; ++b)
            offs[b] = (int)(b * nslots_);
        CUDA_CHECK(cudaMemcpyAsync(c.ex_offs_.p, offs.data(), (nb_max + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        const bool tables = prepare_chain_tables(c, d_q, c.exact_nq_, s);
        const PairMetric pm{sp_, pts64_.p, d_q, tables ? trig_.p : nullptr, tables ? c.qtrig_.p : nullptr};
        for (size_t b0 = 0; b0 < nsel; b0 += nb_max)
        {
            const size_t nb = std::min(nb_max, nsel - b0);
            const size_t items = nb * nslots_;
            REQUIRE(items < ((size_t)1 << 31), "exact path: batch too large");
            dist_rows_kernel<<<nblk(items, 256), 256, 0, s>>>(pm, alive_.p, nslots_, d_sel, b0, nb, c.ex_keys_.p,
                                                              c.ex_vals_.p);
            launches_++;
            CUDA_CHECK(cudaGetLastError());
            size_t tmp = 0;
            if (nb == 1)
            {
                CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp, c.ex_keys_.p, c.ex_keys2_.p, c.ex_vals_.p,
                                                           c.ex_vals2_.p, (int)items, 0, 64, s));
                c.cub_tmp_.reserve(tmp);
                CUDA_CHECK(cub::DeviceRadixSort::SortPairs(c.cub_tmp_.p, tmp, c.ex_keys_.p, c.ex_keys2_.p, c.ex_vals_.p,
                                                           c.ex_vals2_.p, (int)items, 0, 64, s));
            }
            else
            {
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp, c.ex_keys_.p, c.ex_keys2_.p, c.ex_vals_.p,
                                                                    c.ex_vals2_.p, (int)items, (int)nb, c.ex_offs_.p,
                                                                    c.ex_offs_.p + 1, 0, 64, s));
                c.cub_tmp_.reserve(tmp);
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp, c.ex_keys_.p, c.ex_keys2_.p,
                                                                    c.ex_vals_.p, c.ex_vals2_.p, (int)items, (int)nb,
                                                                    c.ex_offs_.p, c.ex_offs_.p + 1, 0, 64, s));
            }
            launches_ += 8;
            exact_emit_kernel<<<nblk(nb * kk, 256), 256, 0, s>>>(c.ex_keys2_.p, c.ex_vals2_.p, nslots_, ids_.p, d_sel, b0,
                                                                 nb, (uint32_t)kk, (uint32_t)kstride, d_ids, d_dists);
            launches_++;
            CUDA_CHECK(cudaGetLastError());
        }
        CUDA_CHECK(cudaStreamSynchronize(s));
    }

    void NNIndex::knn_device(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                             cudaStream_t s)
    {
        std::shared_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        Lease L(*this);
        knn_locked(*L.c, d_q, nq, k, d_ids, d_dists, s);
        publish(*L.c);
    }

    // rw_ is held (shared) by the caller; `c` is the caller's leased context
    void NNIndex::knn_locked(QCtx &c, const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                             cudaStream_t s, bool sync)
    {
        if (nq == 0 || k == 0)
            return;
        REQUIRE(d_q && d_ids, "omplcu_nn_knn: NULL buffer");
        REQUIRE(k < ((size_t)1 << 31), "omplcu_nn_knn: k too large");
        const size_t kk = std::min(k, nlive_);
        c.last_fallbacks_ = 0;
        c.last_path_ = 0;
        c.exact_nq_ = nq;
        if (kk > 0)
        {
            if (fast_eligible(kk))
            {
                c.last_path_ = 1;
                knn_fast(c, d_q, nq, kk, k, d_ids, d_dists, s);
            }
            else
            {
                c.last_path_ = 2;
                c.last_few_ = 0;
                // the one-pass kernel evaluates every (query, slot) distance: it wins up to ~5M pairs (8 queries on 100K
                // states, 4 on 1M: 0.24 / 0.43 ms), beyond that the sample-bounded collect path does (scripts/time_se3_small.py)
                if (few_mode_ && !force_exact_ && nq <= (size_t)XFEW_MAXQ && kk <= (size_t)XFEW_MAXK && nslots_ >= 1024 &&
                    nq * nslots_ <= (size_t)5000000 && nslots_ < ((size_t)1 << 32) - 64)
                    knn_exact_few(c, d_q, nq, kk, k, d_ids, d_dists, s);
                else
                    knn_exact(c, d_q, nullptr, nq, kk, k, d_ids, d_dists, s);
            }
        }
        if (kk < k)
        {
            pad_rows_kernel<<<nblk(nq * (k - kk), 256), 256, 0, s>>>(d_ids, d_dists, nq, (uint32_t)kk, (uint32_t)k);
            launches_++;
            CUDA_CHECK(cudaGetLastError());
        }
        if (sync)
            CUDA_CHECK(cudaStreamSynchronize(s));
    }

    // Host pointers in, host pointers out.  One shared lock spans staging, query and copy-out, so the
    // count reported (min(k, size)) is the one the query used; staging lives in the leased context, so
    // host threads do not serialise on it.
    void NNIndex::knn_host(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists, uint32_t *counts)
    {
        if (nq == 0)
            return;
        REQUIRE(q && ids, "omplcu_nn_knn: NULL buffer");
        const size_t st = (size_t)sp_.stride;
        size_t kk;
        {
            std::shared_lock<std::shared_mutex> lk(rw_);
            DeviceGuard g(device_);
            Lease L(*this);
            QCtx &c = *L.c;
            c.q_stage_.reserve(nq * st);
            c.out_i_.reserve(nq * std::max<size_t>(k, 1));
            c.out_d_.reserve(nq * std::max<size_t>(k, 1));
            kk = std::min(k, nlive_);
            // Small calls (the literal nearestK of PRM::addMilestone, prm/src/PRM.cpp:575) go through pinned staging:
            // pageable copies are synchronous round trips of 10 - 15 us each, more than the scan of a 100K roadmap.
            const size_t qbytes = nq * st * sizeof(double), entries = nq * k;
            const bool small = k > 0 && qbytes + entries * 16 <= ((size_t)1 << 20);
            if (small)
            {
                c.pin_.reserve(qbytes + entries * 16 + 16);
                memcpy(c.pin_.p, q, qbytes);
                const size_t off = (qbytes + 15) & ~(size_t)15;
                c.ho_ids_ = reinterpret_cast<uint64_t *>(c.pin_.p + off);
                c.ho_dists_ = reinterpret_cast<double *>(c.pin_.p + off + entries * 8);
                c.ho_entries_ = entries;
                c.ho_done_ = false;
            }
            CUDA_CHECK(cudaMemcpyAsync(c.q_stage_.p, small ? (const void *)c.pin_.p : (const void *)q, qbytes,
                                       cudaMemcpyHostToDevice, c.stream));
            if (k > 0)
            {
                struct Clear
                {
                    QCtx &c;
                    ~Clear()
                    {
                        c.ho_ids_ = nullptr;
                        c.ho_dists_ = nullptr;
                        c.ho_entries_ = 0;
                    }
                } clear{c};
                knn_locked(c, c.q_stage_.p, nq, k, c.out_i_.p, c.out_d_.p, c.stream, !small);
                if (small)
                {
                    if (!c.ho_done_)  // another path than the fast one, or some queries were redone exactly
                    {
                        CUDA_CHECK(cudaMemcpyAsync(c.ho_ids_, c.out_i_.p, entries * 8, cudaMemcpyDeviceToHost, c.stream));
                        CUDA_CHECK(cudaMemcpyAsync(c.ho_dists_, c.out_d_.p, entries * 8, cudaMemcpyDeviceToHost, c.stream));
                        CUDA_CHECK(cudaStreamSynchronize(c.stream));
                    }
                    memcpy(ids, c.ho_ids_, entries * 8);
                    if (dists)
                        memcpy(dists, c.ho_dists_, entries * 8);
                }
                else
                {
                    CUDA_CHECK(cudaMemcpyAsync(ids, c.out_i_.p, nq * k * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
                    if (dists)
                        CUDA_CHECK(cudaMemcpyAsync(dists, c.out_d_.p, nq * k * sizeof(double), cudaMemcpyDeviceToHost,
                                                   c.stream));
                }
            }
            CUDA_CHECK(cudaStreamSynchronize(c.stream));
            publish(c);
        }
        if (counts)
            for (size_t i = 0; i < nq; ++i)
                counts[i] = (uint32_t)kk;
    }

    // NearestNeighborsLinear::nearestR (NearestNeighborsLinear.h:135-142): all elements with
    // distance <= radius (inclusive), sorted by (distance, insertion order).  Exact FP64 distances
    // for every (query, slot) pair; the pairs within the radius are counted while the distances are
    // written, copied out per query, and only they are sorted (by slot, then stably by distance).
    void NNIndex::radius_host(const double *q, size_t nq, const double *radii, uint64_t **ids, double **dists,
                              uint64_t *offsets)
    {
        REQUIRE(ids && offsets, "omplcu_nn_radius: NULL output");
        *ids = nullptr;
        if (dists)
            *dists = nullptr;
        offsets[0] = 0;
        if (nq == 0)
            return;
        REQUIRE(q && radii, "omplcu_nn_radius: NULL buffer");
        std::shared_lock<std::shared_mutex> lk(rw_);
        DeviceGuard g(device_);
        Lease L(*this);
        QCtx &c = *L.c;
        cudaStream_t s = c.stream;
        const size_t st = (size_t)sp_.stride;
        size_t total = 0;  // neighbours found so far (all batches), kept on the device in c.rad_ids_/c.rad_d_
        if (nslots_ > 0 && nlive_ > 0)
        {
            REQUIRE(nslots_ < ((size_t)1 << 31), "radius: too many slots");
            c.q_stage_.reserve(nq * st);
            c.rad_r_.reserve(nq);
            CUDA_CHECK(cudaMemcpyAsync(c.q_stage_.p, q, nq * st * sizeof(double), cudaMemcpyHostToDevice, s));
            CUDA_CHECK(cudaMemcpyAsync(c.rad_r_.p, radii, nq * sizeof(double), cudaMemcpyHostToDevice, s));
            // Plain RealVector spaces: the tensor-core threshold scan keeps, per query, every slot whose scan
            // value is below the scan-value image of the radius (a superset of the slots within it); only those get
            // an FP64 distance and the inclusive test d <= r of NearestNeighborsLinear.h:139.  A batch in which
            // some query has more than RCAP neighbours (or an unbounded radius) takes the all-pairs path below.
            const bool filt = filter_usable() && embed_ok(s);
            constexpr uint32_t RCAP = 2048;
            c.last_filter_ = 0;
            const size_t nb_max = filt ? std::min<size_t>(nq, 8192)
                                       : std::max<size_t>(1, std::min(nq, RADIUS_PAIR_BUDGET / nslots_));
            const size_t nb_pairs = std::max<size_t>(1, std::min(nb_max, RADIUS_PAIR_BUDGET / nslots_));  // all-pairs sub-batch
            if (!filt)
                c.ex_keys_.reserve(nb_max * nslots_);
            c.col_cnt_.reserve(2 * nb_max);
            c.rad_off_.reserve(nb_max);
            c.ex_offs_.reserve(nb_max + 1);
            std::vector<uint32_t> cnt(nb_max);
            std::vector<unsigned long long> boff(nb_max);
            std::vector<int> seg(nb_max + 1);
            const bool tables = prepare_chain_tables(c, c.q_stage_.p, nq, s);
            const PairMetric pm{sp_, pts64_.p, c.q_stage_.p, tables ? trig_.p : nullptr, tables ? c.qtrig_.p : nullptr};
            size_t nb_done = 0;
            for (size_t b0 = 0; b0 < nq; b0 += nb_done)
            {
                size_t nb = std::min(nb_max, nq - b0);
                nb_done = nb;
                uint32_t *d_cnt = c.col_cnt_.p, *d_fill = c.col_cnt_.p + nb_max;
                if (filt)
                {
                    uint32_t fq = 0;
                    c.col_sd_.reserve(nb_max * RCAP);
                    c.col_redo_.reserve(nb_max * RCAP);
                    if (embed_)
                    {
                        // the filter runs on the lower-bound embedding, the kept slots get the space's own distance
                        c.q_emb_.reserve(nb * (size_t)D_);
                        c.emb_bound_.reserve(nb);
                        embed_queries_kernel<<<nblk(nb, 128), 128, 0, s>>>(sp_, c.q_stage_.p + b0 * st, nb, D_, embed_slack_,
                                                                           c.rad_r_.p + b0, c.q_emb_.p, c.emb_bound_.p);
                        const uint32_t rows = filter_scan(c, c.q_emb_.p, nb, c.emb_bound_.p, RCAP, &fq, s);
                        kept_collect_metric_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.app_ids_.p, c.app_cnt_.p, rows, RCAP, fq, pm,
                                                                               b0, alive_.p, c.rad_r_.p + b0, nb, RCAP,
                                                                               c.col_sd_.p, c.col_redo_.p, d_cnt);
                        launches_++;
                    }
                    else
                    {
                    const uint32_t rows = filter_scan(c, c.q_stage_.p + b0 * st, nb, c.rad_r_.p + b0, RCAP, &fq, s);
                    kept_collect_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.app_ids_.p, c.app_cnt_.p, rows, RCAP, fq, pts64_.p,
                                                                    alive_.p, D_, c.q_stage_.p + b0 * st,
                                                                    c.rad_r_.p + b0, nb, RCAP, c.col_sd_.p,
                                                                    c.col_redo_.p, d_cnt);
                    }
                    launches_++;
                    CUDA_CHECK(cudaGetLastError());
                    CUDA_CHECK(cudaMemcpyAsync(cnt.data(), d_cnt, nb * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
                    CUDA_CHECK(cudaStreamSynchronize(s));
                    bool fits = true;
                    for (size_t b = 0; b < nb; ++b)
                        fits = fits && cnt[b] <= RCAP;
                    if (fits)
                    {
                        unsigned long long tot = 0;
                        for (size_t b = 0; b < nb; ++b)
                        {
                            boff[b] = tot;
                            seg[b] = (int)tot;
                            tot += cnt[b];
                            offsets[b0 + b + 1] = offsets[b0 + b] + cnt[b];
                        }
                        seg[nb] = (int)tot;
                        c.last_filter_ = 1;
                        if (tot == 0)
                            continue;
                        c.ex_keys2_.reserve(tot);
                        c.ex_keys3_.reserve(tot);
                        c.ex_vals_.reserve(tot);
                        c.ex_vals2_.reserve(tot);
                        c.rad_ids_.reserve(total + tot, true, s);
                        c.rad_d_.reserve(total + tot, true, s);
                        CUDA_CHECK(cudaMemcpyAsync(c.rad_off_.p, boff.data(), nb * sizeof(unsigned long long),
                                                   cudaMemcpyHostToDevice, s));
                        CUDA_CHECK(cudaMemcpyAsync(c.ex_offs_.p, seg.data(), (nb + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
                        uint32_t cmax = 0;
                        for (size_t b = 0; b < nb; ++b)
                            cmax = std::max(cmax, cnt[b]);
                        if (cmax <= 256)
                        {
                            radius_sort_emit_kernel<256><<<nblk(nb, 4), 128, 0, s>>>(
                                c.col_sd_.p, c.col_redo_.p, d_cnt, RCAP, c.rad_off_.p, nb, ids_.p, c.rad_ids_.p + total,
                                c.rad_d_.p + total);
                            launches_++;
                            CUDA_CHECK(cudaGetLastError());
                            total += tot;
                            continue;
                        }
                        // survivors are in slot order already: one stable sort by distance
                        segment_pack_kernel<<<nblk(nb, 4), 128, 0, s>>>(c.col_sd_.p, c.col_redo_.p, d_cnt, RCAP, c.rad_off_.p,
                                                                        nb, c.ex_keys3_.p, c.ex_vals2_.p);
                        size_t tmp2 = 0;
                        CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp2, c.ex_keys3_.p, c.ex_keys2_.p,
                                                                            c.ex_vals2_.p, c.ex_vals_.p, (int)tot, (int)nb,
                                                                            c.ex_offs_.p, c.ex_offs_.p + 1, 0, 64, s));
                        c.cub_tmp_.reserve(tmp2);
                        tmp2 = c.cub_tmp_.cap;
                        CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp2, c.ex_keys3_.p, c.ex_keys2_.p,
                                                                            c.ex_vals2_.p, c.ex_vals_.p, (int)tot, (int)nb,
                                                                            c.ex_offs_.p, c.ex_offs_.p + 1, 0, 64, s));
                        radius_emit_kernel<<<nblk(tot, 256), 256, 0, s>>>(c.ex_keys2_.p, c.ex_vals_.p, ids_.p, tot,
                                                                          c.rad_ids_.p + total, c.rad_d_.p + total);
                        launches_ += 10;
                        CUDA_CHECK(cudaGetLastError());
                        total += tot;
                        continue;
                    }
                    // too many neighbours for the filtered path: all pairs, in sub-batches the distance matrix allows
                    nb = std::min(nb, nb_pairs);
                    nb_done = nb;
                    c.ex_keys_.reserve(nb_pairs * nslots_);
                }
                const size_t items = nb * nslots_;
                CUDA_CHECK(cudaMemsetAsync(c.col_cnt_.p, 0, 2 * nb_max * sizeof(uint32_t), s));
                radius_dist_kernel<<<nblk(items, 256), 256, 0, s>>>(pm, alive_.p, nslots_, b0, nb, c.rad_r_.p, c.ex_keys_.p,
                                                                    d_cnt);
                launches_++;
                CUDA_CHECK(cudaGetLastError());
                CUDA_CHECK(cudaMemcpyAsync(cnt.data(), d_cnt, nb * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
                CUDA_CHECK(cudaStreamSynchronize(s));
                unsigned long long tot = 0;
                for (size_t b = 0; b < nb; ++b)
                {
                    boff[b] = tot;
                    seg[b] = (int)tot;
                    tot += cnt[b];
                    offsets[b0 + b + 1] = offsets[b0 + b] + cnt[b];
                }
                seg[nb] = (int)tot;
                if (tot == 0)
                    continue;
                REQUIRE(tot < ((unsigned long long)1 << 31), "radius: too many neighbours in one batch");
                c.ex_keys2_.reserve(tot);
                c.ex_keys3_.reserve(tot);
                c.ex_vals_.reserve(tot);
                c.ex_vals2_.reserve(tot);
                c.rad_ids_.reserve(total + tot, true, s);
                c.rad_d_.reserve(total + tot, true, s);
                CUDA_CHECK(cudaMemcpyAsync(c.rad_off_.p, boff.data(), nb * sizeof(unsigned long long),
                                           cudaMemcpyHostToDevice, s));
                CUDA_CHECK(cudaMemcpyAsync(c.ex_offs_.p, seg.data(), (nb + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
                radius_compact_kernel<<<nblk(items, 256), 256, 0, s>>>(c.ex_keys_.p, nslots_, b0, nb, c.rad_r_.p, c.rad_off_.p,
                                                                       d_fill, c.ex_keys2_.p, c.ex_vals_.p);
                // by slot (payload distance), then stably by distance (payload slot)
                size_t tmp = 0, tmp2 = 0;
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp, c.ex_vals_.p, c.ex_vals2_.p, c.ex_keys2_.p,
                                                                    c.ex_keys3_.p, (int)tot, (int)nb, c.ex_offs_.p,
                                                                    c.ex_offs_.p + 1, 0, 32, s));
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp2, c.ex_keys3_.p, c.ex_keys2_.p,
                                                                    c.ex_vals2_.p, c.ex_vals_.p, (int)tot, (int)nb,
                                                                    c.ex_offs_.p, c.ex_offs_.p + 1, 0, 64, s));
                c.cub_tmp_.reserve(std::max(tmp, tmp2));
                tmp = tmp2 = c.cub_tmp_.cap;
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp, c.ex_vals_.p, c.ex_vals2_.p,
                                                                    c.ex_keys2_.p, c.ex_keys3_.p, (int)tot, (int)nb,
                                                                    c.ex_offs_.p, c.ex_offs_.p + 1, 0, 32, s));
                CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortPairs(c.cub_tmp_.p, tmp2, c.ex_keys3_.p, c.ex_keys2_.p,
                                                                    c.ex_vals2_.p, c.ex_vals_.p, (int)tot, (int)nb,
                                                                    c.ex_offs_.p, c.ex_offs_.p + 1, 0, 64, s));
                radius_emit_kernel<<<nblk(tot, 256), 256, 0, s>>>(c.ex_keys2_.p, c.ex_vals_.p, ids_.p, tot,
                                                                  c.rad_ids_.p + total, c.rad_d_.p + total);
                launches_ += 18;
                CUDA_CHECK(cudaGetLastError());
                total += tot;
            }
        }
        else
            for (size_t i = 0; i < nq; ++i)
                offsets[i + 1] = 0;
        *ids = (uint64_t *)malloc(std::max<size_t>(total, 1) * sizeof(uint64_t));
        REQUIRE(*ids != nullptr, "omplcu_nn_radius: out of host memory");
        if (dists)
        {
            *dists = (double *)malloc(std::max<size_t>(total, 1) * sizeof(double));
            REQUIRE(*dists != nullptr, "omplcu_nn_radius: out of host memory");
        }
        if (total)
        {
            CUDA_CHECK(cudaMemcpyAsync(*ids, c.rad_ids_.p, total * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
            if (dists)
                CUDA_CHECK(cudaMemcpyAsync(*dists, c.rad_d_.p, total * sizeof(double), cudaMemcpyDeviceToHost, s));
        }
        CUDA_CHECK(cudaStreamSynchronize(s));
        c.last_path_ = 2;
        publish(c);
    }

    void merge_topk_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts, size_t part_stride,
                           size_t nq, size_t k, double *d_out_dists, uint64_t *d_out_ids, cudaStream_t s)
    {
        DeviceGuard g(device);
        REQUIRE(nparts >= 1 && d_dists && d_ids && d_out_dists && d_out_ids, "omplcu_merge_topk: bad arguments");
        REQUIRE(part_stride >= nq * k, "omplcu_merge_topk: part stride smaller than one part");
        if (nq == 0 || k == 0)
            return;
        REQUIRE(k < ((size_t)1 << 31) / (size_t)nparts, "omplcu_merge_topk: nparts * k too large");
        const size_t smem = 4 * (size_t)nparts * k * 16;
        if (smem > 200 * 1024)
        {
            // more candidates per query than shared memory stages: ranks by binary search in the sorted lists
            merge_sorted_kernel<<<(unsigned)std::min<size_t>(nq, 148 * 64), 256, 0, s>>>(
                d_dists, d_ids, nparts, part_stride, nq, (uint32_t)k, d_out_dists, d_out_ids);
            CUDA_CHECK(cudaGetLastError());
            CUDA_CHECK(cudaStreamSynchronize(s));
            return;
        }
        if (smem > 48 * 1024)
            CUDA_CHECK(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        merge_topk_kernel<<<nblk(nq, 4), 128, smem, s>>>(d_dists, d_ids, nparts, part_stride, nq, (uint32_t)k,
                                                         d_out_dists, d_out_ids);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(s));
    }
}  // namespace omplcu
