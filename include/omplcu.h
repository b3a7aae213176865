/*
 * omplcu.h -- C ABI of the B200-native OMPL hot path (libomplcu.so).
 *
 * Drop-in boundary.  Plain pointers and sizes only; every entry
 * point returns an int status (0 = OMPLCU_OK) and never throws; the message of
 * the last failure on the calling thread is omplcu_last_error().  Host-pointer
 * entry points copy in/out themselves; *_device entry points take device
 * pointers on the handle's GPU and a cudaStream_t passed as void*.
 *
 * What each entry point replaces in the reference (/root/reference/src/ompl):
 *   omplcu_space_*         base/StateSpace.h distance / interpolate / validSegmentCount
 *                          (base/src/StateSpace.cpp:851-854,1068-1116;
 *                           base/spaces/src/RealVectorStateSpace.cpp:230-265,
 *                           SO2StateSpace.cpp:116-155, SO3StateSpace.cpp:254-318)
 *   omplcu_nn_add          NearestNeighbors<_T>::add            datastructures/NearestNeighbors.h:78-85
 *   omplcu_nn_remove       NearestNeighbors<_T>::remove         NearestNeighbors.h:88
 *   omplcu_nn_clear/size/list                                   NearestNeighbors.h:75,106,109
 *   omplcu_nn_knn          NearestNeighborsLinear::nearestK     NearestNeighborsLinear.h:119-132
 *   omplcu_nn_radius       NearestNeighborsLinear::nearestR     NearestNeighborsLinear.h:135-142
 *   omplcu_nn_knn(k=1)     NearestNeighborsLinear::nearest      NearestNeighborsLinear.h:98-116
 *   omplcu_check_states    StateValidityChecker::isValid        base/StateValidityChecker.h:111
 *   omplcu_check_motions   DiscreteMotionValidator::checkMotion base/src/DiscreteMotionValidator.cpp:48-145
 *   omplcu_checker_create_hypercube   demos/HypercubeBenchmark.cpp:57-72
 *   omplcu_checker_create_chain       demos/KinematicChain.h:193-276
 *   omplcu_checker_create_circles     tests/resources/circles2D.h:139-149
 *   omplcu_checker_create_se3         (not in the reference; specified in DESIGN.md)
 *
 * Neighbour order is the total order (distance, id): NearestNeighborsLinear's
 * result with ties broken by insertion index.  Ids are insertion sequence
 * numbers (first id of a handle is its id base, default 0) and never change.
 */
#ifndef OMPLCU_H
#define OMPLCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OMPLCU_OK 0
#define OMPLCU_ERR_INVALID 1   /* bad argument */
#define OMPLCU_ERR_CUDA 2      /* CUDA runtime failure */
#define OMPLCU_ERR_NOT_FOUND 3 /* remove(): id absent (reference returns false) */
#define OMPLCU_ERR_EMPTY 4     /* nearest on empty structure ("No elements found ...") */
#define OMPLCU_ERR_INTERNAL 5

#define OMPLCU_ID_NONE UINT64_MAX

enum { OMPLCU_RV = 0, OMPLCU_SO2 = 1, OMPLCU_SO3 = 2, OMPLCU_CHAIN = 3 };
#define OMPLCU_MAX_COMPONENTS 8

/* One (sub)space.  dim = stored doubles (RV: dimension; SO2: 1; SO3: 4 in
 * x,y,z,w order; CHAIN: links).  lvs = StateSpace::getLongestValidSegmentLength()
 * of that (sub)space, factor = getValidSegmentCountFactor(). */
typedef struct
{
    int32_t type;
    int32_t dim;
    double weight;
    double lvs;
    uint32_t factor;
    uint32_t _pad;
    double param; /* CHAIN: link length */
} omplcu_component;

typedef struct omplcu_space_s *omplcu_space_t;
typedef struct omplcu_nn_s *omplcu_nn_t;
typedef struct omplcu_checker_s *omplcu_checker_t;

const char *omplcu_last_error(void);
int omplcu_version(void);
int omplcu_device_count(int *count);
void omplcu_free(void *p);
/* Benchmark hygiene (no reference counterpart): evict the L2 of `device` by writing a library-owned 256 MB buffer. */
int omplcu_device_flush_l2(int device);
/* FP32 FMA pipe peak of `device` in TFLOP/s (FMA = 2 flop), measured with a
 * register-only micro-benchmark: scalar FFMA and packed FFMA2.  This is the
 * roofline denominator for the FP32-bound kNN scan. */
int omplcu_measure_fp32_peak(int device, double *tflops_ffma, double *tflops_ffma2);

/* ---- state space ---- */
/* compound != 0: CompoundStateSpace semantics (weighted sum of component
 * distances; validSegmentCount = max over components).  compound == 0
 * requires ncomp == 1. */
int omplcu_space_create(const omplcu_component *comps, int ncomp, int compound, omplcu_space_t *out);
int omplcu_space_destroy(omplcu_space_t sp);
int omplcu_space_stride(omplcu_space_t sp, int *stride);
/* batched space math on `device` (host buffers): rows of `stride` doubles */
int omplcu_space_distance(omplcu_space_t sp, int device, const double *a, const double *b, size_t n, double *out);
int omplcu_space_interpolate(omplcu_space_t sp, int device, const double *from, const double *to, const double *t,
                             size_t n, double *out);
int omplcu_space_segment_count(omplcu_space_t sp, int device, const double *a, const double *b, size_t n,
                               uint32_t *out);

/* out[i] = StateSpace::satisfiesBounds(states[i]) (RealVectorStateSpace.cpp:200-208 with its +-DBL_EPSILON
 * slack, SO2StateSpace.cpp:90-94, SO3StateSpace.cpp:213-216, compound: all components, StateSpace.cpp:1021-1028).
 * low / high: `stride` doubles in state layout; entries of SO2 / SO3 components are ignored.  Valid-state
 * samplers call it before isValid (base/samplers/src/UniformValidStateSampler.cpp:53). */
int omplcu_space_satisfies_bounds(omplcu_space_t sp, int device, const double *low, const double *high,
                                  const double *states, size_t n, uint8_t *out);
/* StateSpace::getMaximumExtent (RealVectorStateSpace.cpp:167-176, SO2StateSpace.cpp:70-73, SO3StateSpace.cpp:166-169,
 * StateSpace.cpp:996-1003); the longest valid segment of a space is this times the fraction (StateSpace.cpp:237-248). */
int omplcu_space_maximum_extent(omplcu_space_t sp, const double *low, const double *high, double *extent);

/* ---- nearest neighbours ---- */
int omplcu_nn_create(omplcu_space_t sp, int device, omplcu_nn_t *out);
/* The same structure with its states sharded over `ndev` GPUs of this process ("omplcu_nn_create(space,
 * devices[], ndev, &h)"; 8(e)).  Every other omplcu_nn_* call works on the handle unchanged and returns exactly
 * what a single-device handle holding all the states returns: states are split by insertion index (a bulk add
 * into ndev contiguous blocks, a small add to the least-loaded shard), ids stay global insertion sequence
 * numbers, every shard answers all queries on its own GPU, the per-shard (distance, id) lists are gathered on
 * devices[0] over NVLink peer copies and merged there by (distance, id).  Device-pointer entry points take and
 * return buffers on devices[0].  A device may be named more than once (several shards on one GPU). */
int omplcu_nn_create_sharded(omplcu_space_t sp, const int *devices, int ndev, omplcu_nn_t *out);
int omplcu_nn_destroy(omplcu_nn_t nn);
/* ids handed out are id_base + insertion sequence (sharded roadmaps give each
 * rank a disjoint base).  Must be called while the structure is empty. */
int omplcu_nn_set_id_base(omplcu_nn_t nn, uint64_t id_base);
int omplcu_nn_add(omplcu_nn_t nn, const double *rows, size_t n, uint64_t *first_id);
int omplcu_nn_remove(omplcu_nn_t nn, uint64_t id);
/* n removals in one call (BIT* prunes samples in bulk: ImplicitGraph.cpp, pruneSamples): ids that are
 * not present are skipped, *removed (may be NULL) receives how many were. */
int omplcu_nn_remove_batch(omplcu_nn_t nn, const uint64_t *ids, size_t n, size_t *removed);
int omplcu_nn_clear(omplcu_nn_t nn);
int omplcu_nn_size(omplcu_nn_t nn, size_t *n);
int omplcu_nn_list(omplcu_nn_t nn, uint64_t *ids, size_t cap, size_t *n);
/* Read back the stored coordinates of `id` (stride doubles). */
int omplcu_nn_get(omplcu_nn_t nn, uint64_t id, double *row);
/* k nearest per query.  ids/dists are nq x k row-major, rows padded with
 * OMPLCU_ID_NONE / +inf beyond counts[i] = min(k, size).  dists and counts may
 * be NULL. */
int omplcu_nn_knn(omplcu_nn_t nn, const double *queries, size_t nq, size_t k, uint64_t *ids, double *dists,
                  uint32_t *counts);
/* All neighbours with distance <= radii[i], sorted.  offsets has nq+1
 * entries; *ids / *dists are allocated by the library (release with
 * omplcu_free); dists may be NULL. */
int omplcu_nn_radius(omplcu_nn_t nn, const double *queries, size_t nq, const double *radii, uint64_t **ids,
                     double **dists, uint64_t *offsets);
/* device-resident variant: d_queries (nq x stride), d_ids / d_dists (nq x k)
 * live on the handle's GPU.  Synchronises `stream` before returning. */
int omplcu_nn_knn_device(omplcu_nn_t nn, const double *d_queries, size_t nq, size_t k, uint64_t *d_ids,
                         double *d_dists, void *stream);
/* Merge nparts partial results laid out [part][nq][k] (as produced by an
 * all-gather of omplcu_nn_knn_device outputs) into the k best per query by
 * (distance, id).  Device pointers.  Up to 3200 candidates per query
 * (nparts * k) are ranked in shared memory in any order; beyond that every
 * part's row must be sorted by distance with absent entries (OMPLCU_ID_NONE)
 * last, as omplcu_nn_knn_device writes them, and ranks come from binary
 * searches in the rows (equal distances are ordered by id wherever they
 * stand).  nparts * k < 2^31. */
int omplcu_merge_topk_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts, size_t nq,
                             size_t k, double *d_out_dists, uint64_t *d_out_ids, void *stream);
/* The same with `part_stride` elements between consecutive parts (>= nq * k): lets a rank keep its ids and
 * distances in ONE buffer ([ids | dists], both 8 bytes wide) so that a single all-gather moves both. */
int omplcu_merge_topk_strided_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts,
                                     size_t part_stride, size_t nq, size_t k, double *d_out_dists,
                                     uint64_t *d_out_ids, void *stream);
/* Tuning / introspection.  Options: "force_exact" (0/1); "tc_scan" (-1 auto and 1: tensor-core filter scan
 * whenever the space is eligible (plain RealVector, d <= 128, k <= 40), 0 never: FP32-pipe scans; the environment
 * variable OMPLCU_TC_SCAN sets the default for new handles); "sample_thresholds", "collect",
 * "scan_ctas_per_sm", "tc_tiles", "profile" (tuning); "few_scan" (1: calls of up to 8 queries in auto mode run the
 * lane-per-point scan, default), "tc_ts" (query tile in tensor memory), "tc_pair" (CTA pairs share the point chunks),
 * "tc_align" (work pieces end on query-tile boundaries), "thr_scan", "l2_window" (tuning);
 * "tc_scan_values_out" (diagnostics: device pointer that receives the raw scan values); sharded handles only:
 * "gather_entries" ((distance, id) entries the gather buffer on the first device may hold, default 2^27 = 2 GB; larger
 * batches are answered in pieces).  Stats: "last_fallbacks"
 * (queries redone by the exact path), "last_rescans" (queries the tensor-core filter passed on to the FP32-pipe scan),
 * "tc_demoted" (auto mode made the FP32-pipe scan the first filter: the set is too dense for the BF16 band),
 * "last_path" (1 filter scan + exact re-rank, 2 exact sort), "last_tc" (1 if the last scan ran on the tensor cores),
 * "last_few" (1 if it was the few-query scan), "launches" (kernels launched by this handle so far), "last_scan_ns",
 * "slots". */
int omplcu_nn_set_option(omplcu_nn_t nn, const char *key, long value);
int omplcu_nn_get_stat(omplcu_nn_t nn, const char *key, long *value);

/* ---- validity ---- */
int omplcu_checker_create_allvalid(int device, omplcu_checker_t *out);
int omplcu_checker_create_hypercube(int device, int dim, double edge_width, omplcu_checker_t *out);
/* env: nenv x (x0,y0,x1,y1) */
int omplcu_checker_create_chain(int device, int nlinks, double link_length, const double *env, int nenv,
                                omplcu_checker_t *out);
/* robot: nrobot x (ox,oy,oz,r); sph: nsph x (cx,cy,cz,r); box: nbox x (cx,cy,cz,hx,hy,hz) */
int omplcu_checker_create_se3(int device, const double *robot, int nrobot, const double *sph, int nsph,
                              const double *box, int nbox, omplcu_checker_t *out);
/* circ: n x (x,y,r) */
int omplcu_checker_create_circles(int device, const double *circ, int n, omplcu_checker_t *out);
int omplcu_checker_destroy(omplcu_checker_t c);
int omplcu_checker_get_stat(omplcu_checker_t c, const char *key, long *value);

/* valid[i] = isValid(states[i]); rows of `stride` doubles. */
int omplcu_check_states(omplcu_checker_t c, int stride, const double *states, size_t n, uint8_t *valid);
/* valid[i] = checkMotion(s1[i], s2[i]).  If last_t != NULL the lastValid
 * overload is evaluated: last_t[i] = lastValid.second on failure, 1.0 on
 * success; last_state (n x stride, may be NULL) = lastValid.first on failure. */
int omplcu_check_motions(omplcu_checker_t c, omplcu_space_t sp, const double *s1, const double *s2, size_t n,
                         uint8_t *valid, double *last_t, double *last_state);
/* One batch split by index over `nchk` checkers, normally one per GPU holding the same obstacle set (SURVEY 8(e):
 * validity shards by segment, no exchange).  Results are those of omplcu_check_states / omplcu_check_motions. */
int omplcu_check_states_multi(const omplcu_checker_t *checkers, int nchk, int stride, const double *states, size_t n,
                              uint8_t *valid);
int omplcu_check_motions_multi(const omplcu_checker_t *checkers, int nchk, omplcu_space_t sp, const double *s1,
                               const double *s2, size_t n, uint8_t *valid, double *last_t, double *last_state);
int omplcu_check_states_device(omplcu_checker_t c, int stride, const double *d_states, size_t n, uint8_t *d_valid,
                               void *stream);
int omplcu_check_motions_device(omplcu_checker_t c, omplcu_space_t sp, const double *d_s1, const double *d_s2,
                                size_t n, uint8_t *d_valid, double *d_last_t, double *d_last_state, void *stream);

/* Allocation-free variant of omplcu_check_motions_device.  It has NO counterpart in the reference; it was
 * added for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * omplcu_check_motions_device, but the per-segment scratch (segment counts, first-failure codes, the work
 * counter) lives in a file-scope static __device__ array of the library instead of the checker's growable
 * buffers: the call performs no cudaMalloc / cudaFree and uses no allocating library routine.
 * Limit: n <= OMPLCU_NOALLOC_MAX_SEGMENTS, larger batches return OMPLCU_ERR_INVALID (split them).  The
 * array (2 MiB + 256 B) is held by every GPU that loads the library until the process exits.  It is
 * shared by ALL checkers and calls on a device and keeps the last call's contents; the call clears what
 * must start at zero itself.  Calls that overlap on one GPU (different streams or host threads) RACE on
 * it: issue them one at a time per device.  Synchronises `stream` before returning. */
#define OMPLCU_NOALLOC_MAX_SEGMENTS 262144
int omplcu_check_motions_device_noalloc(omplcu_checker_t c, omplcu_space_t sp, const double *d_s1,
                                        const double *d_s2, size_t n, uint8_t *d_valid, double *d_last_t,
                                        double *d_last_state, void *stream);

#ifdef __cplusplus
}
#endif
#endif
