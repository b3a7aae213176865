// capi.cu -- extern "C" boundary of libomplcu.so (declared in include/omplcu.h).
// No exception crosses the boundary: every entry point returns a status code and
// stores the message in a thread-local string (omplcu_last_error).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "nn.cuh"
#include "validity.cuh"

using namespace omplcu;
namespace omplcu
{
    void measure_fp32_peak(int device, double *tflops_ffma, double *tflops_ffma2);
}

struct omplcu_space_s
{
    SpaceDesc desc;
};
struct omplcu_nn_s
{
    NNApi *impl;
};
struct omplcu_checker_s
{
    Checker *impl;
};

namespace
{
    thread_local std::string g_last_error;

    template <typename F>
    int guarded(F &&f)
    {
        try
        {
            f();
            return OMPLCU_OK;
        }
        catch (const Error &e)
        {
            g_last_error = e.what();
            return e.code;
        }
        catch (const std::bad_alloc &)
        {
            g_last_error = "out of host memory";
            return OMPLCU_ERR_INTERNAL;
        }
        catch (const std::exception &e)
        {
            g_last_error = e.what();
            return OMPLCU_ERR_INTERNAL;
        }
    }

    void need_gpu(int device)
    {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess)
            throw Error(OMPLCU_ERR_CUDA, std::string("no usable CUDA device: ") + cudaGetErrorString(e));
        REQUIRE(device >= 0 && device < n, "device index out of range");
    }
}  // namespace

// A batch split by index over per-device checkers ("segments split evenly by index; the
// obstacle set and space description are replicated"; no exchange).  One host thread per checker.
namespace
{
    template <typename F>
    void split_over(int nc, size_t n, F &&f)
    {
        std::vector<std::thread> th;
        std::vector<std::exception_ptr> err(nc);
        for (int i = 0; i < nc; ++i)
        {
            const size_t b0 = n * (size_t)i / nc, b1 = n * (size_t)(i + 1) / nc;
            if (b1 == b0)
                continue;
            th.emplace_back([&, i, b0, b1] {
                try
                {
                    f(i, b0, b1 - b0);
                }
                catch (...)
                {
                    err[i] = std::current_exception();
                }
            });
        }
        for (auto &t : th)
            t.join();
        for (auto &e : err)
            if (e)
                std::rethrow_exception(e);
    }
}  // namespace

extern "C" {

const char *omplcu_last_error(void)
{
    return g_last_error.c_str();
}

int omplcu_version(void)
{
    return 100;
}

int omplcu_device_count(int *count)
{
    return guarded([&] {
        REQUIRE(count, "count is NULL");
        *count = 0;
        cudaError_t e = cudaGetDeviceCount(count);
        if (e != cudaSuccess)
        {
            *count = 0;
            throw Error(OMPLCU_ERR_CUDA, cudaGetErrorString(e));
        }
    });
}

void omplcu_free(void *p)
{
    free(p);
}

// Benchmark hygiene: evict `device`'s L2 by writing a library-owned 256 MB buffer (> 126 MB L2) and waiting
// for it.  Lets one process flush the L2 of every GPU a sharded handle uses without creating framework
// contexts on them.
int omplcu_device_flush_l2(int device)
{
    return guarded([&] {
        need_gpu(device);
        REQUIRE(device < 64, "device index out of range");
        static std::mutex mu;
        static void *buf[64] = {};
        const size_t bytes = (size_t)256 << 20;
        DeviceGuard g(device);
        {
            std::lock_guard<std::mutex> lk(mu);
            if (!buf[device])
                CUDA_CHECK(cudaMalloc(&buf[device], bytes));
        }
        CUDA_CHECK(cudaMemsetAsync(buf[device], 1, bytes, 0));
        CUDA_CHECK(cudaStreamSynchronize(0));
    });
}

int omplcu_measure_fp32_peak(int device, double *tflops_ffma, double *tflops_ffma2)
{
    return guarded([&] {
        need_gpu(device);
        measure_fp32_peak(device, tflops_ffma, tflops_ffma2);
    });
}

/* ---- space ---- */
int omplcu_space_create(const omplcu_component *comps, int ncomp, int compound, omplcu_space_t *out)
{
    return guarded([&] {
        REQUIRE(comps && out, "omplcu_space_create: NULL argument");
        REQUIRE(ncomp >= 1 && ncomp <= OMPLCU_MAX_COMPONENTS, "omplcu_space_create: 1..8 components");
        REQUIRE(compound || ncomp == 1, "omplcu_space_create: a plain space has exactly one component");
        auto *s = new omplcu_space_s;
        memset(&s->desc, 0, sizeof(s->desc));
        s->desc.ncomp = ncomp;
        s->desc.compound = compound ? 1 : 0;
        int stride = 0;
        for (int i = 0; i < ncomp; ++i)
        {
            const omplcu_component &c = comps[i];
            bool ok = c.type >= OMPLCU_RV && c.type <= OMPLCU_CHAIN && c.dim >= 1 && c.lvs > 0.0 && c.factor >= 1;
            if (c.type == OMPLCU_SO2)
                ok = ok && c.dim == 1;
            if (c.type == OMPLCU_SO3)
                ok = ok && c.dim == 4;
            if (!ok)
            {
                delete s;
                throw Error(OMPLCU_ERR_INVALID, "omplcu_space_create: bad component (type/dim/lvs/factor)");
            }
            s->desc.comp[i] = c;
            stride += c.dim;
        }
        if (stride > MAX_STRIDE)
        {
            delete s;
            throw Error(OMPLCU_ERR_INVALID, "omplcu_space_create: more than 128 values per state");
        }
        s->desc.stride = stride;
        *out = s;
    });
}

int omplcu_space_destroy(omplcu_space_t sp)
{
    delete sp;
    return OMPLCU_OK;
}

int omplcu_space_stride(omplcu_space_t sp, int *stride)
{
    return guarded([&] {
        REQUIRE(sp && stride, "omplcu_space_stride: NULL argument");
        *stride = sp->desc.stride;
    });
}

int omplcu_space_distance(omplcu_space_t sp, int device, const double *a, const double *b, size_t n, double *out)
{
    return guarded([&] {
        REQUIRE(sp, "space is NULL");
        need_gpu(device);
        space_distance_host(sp->desc, device, a, b, n, out);
    });
}

int omplcu_space_interpolate(omplcu_space_t sp, int device, const double *from, const double *to, const double *t,
                             size_t n, double *out)
{
    return guarded([&] {
        REQUIRE(sp, "space is NULL");
        need_gpu(device);
        space_interpolate_host(sp->desc, device, from, to, t, n, out);
    });
}

int omplcu_space_segment_count(omplcu_space_t sp, int device, const double *a, const double *b, size_t n,
                               uint32_t *out)
{
    return guarded([&] {
        REQUIRE(sp, "space is NULL");
        need_gpu(device);
        space_segment_count_host(sp->desc, device, a, b, n, out);
    });
}

int omplcu_space_satisfies_bounds(omplcu_space_t sp, int device, const double *low, const double *high,
                                  const double *states, size_t n, uint8_t *out)
{
    return guarded([&] {
        REQUIRE(sp, "space is NULL");
        need_gpu(device);
        space_satisfies_bounds_host(sp->desc, device, low, high, states, n, out);
    });
}

// StateSpace::getMaximumExtent: RealVectorStateSpace.cpp:167-176 (also the chain space, which derives from
// it), SO2StateSpace.cpp:70-73 (pi), SO3StateSpace.cpp:166-169 (pi / 2), CompoundStateSpace
// (StateSpace.cpp:996-1003: sum of weight * extent over components with weight >= DBL_EPSILON).  One scalar
// per space: host arithmetic, the same operations in the same order as the reference.
int omplcu_space_maximum_extent(omplcu_space_t sp, const double *low, const double *high, double *extent)
{
    return guarded([&] {
        REQUIRE(sp && extent, "omplcu_space_maximum_extent: NULL argument");
        const SpaceDesc &d = sp->desc;
        double tot = 0.0;
        int off = 0;
        for (int c = 0; c < d.ncomp; ++c)
        {
            const omplcu_component &k = d.comp[c];
            double e = 0.0;
            if (k.type == OMPLCU_RV || k.type == OMPLCU_CHAIN)
            {
                REQUIRE(low && high, "omplcu_space_maximum_extent: bounds are NULL");
                for (int j = 0; j < k.dim; ++j)
                {
                    const double w = high[off + j] - low[off + j];
                    e += w * w;
                }
                e = sqrt(e);
            }
            else if (k.type == OMPLCU_SO2)
                e = 3.14159265358979323846;
            else
                e = .5 * 3.14159265358979323846;
            if (!d.compound)
                tot = e;
            else if (k.weight >= 2.220446049250313e-16)
                tot += k.weight * e;
            off += k.dim;
        }
        *extent = tot;
    });
}

/* ---- nearest neighbours ---- */
int omplcu_nn_create(omplcu_space_t sp, int device, omplcu_nn_t *out)
{
    return guarded([&] {
        REQUIRE(sp && out, "omplcu_nn_create: NULL argument");
        need_gpu(device);
        auto *h = new omplcu_nn_s;
        try
        {
            h->impl = new NNIndex(sp->desc, device);
        }
        catch (...)
        {
            delete h;
            throw;
        }
        *out = h;
    });
}

int omplcu_nn_create_sharded(omplcu_space_t sp, const int *devices, int ndev, omplcu_nn_t *out)
{
    return guarded([&] {
        REQUIRE(sp && out && devices, "omplcu_nn_create_sharded: NULL argument");
        REQUIRE(ndev >= 1 && ndev <= 64, "omplcu_nn_create_sharded: 1..64 devices");
        for (int i = 0; i < ndev; ++i)
            need_gpu(devices[i]);
        auto *h = new omplcu_nn_s;
        try
        {
            h->impl = make_sharded_nn(sp->desc, devices, ndev);
        }
        catch (...)
        {
            delete h;
            throw;
        }
        *out = h;
    });
}

int omplcu_nn_destroy(omplcu_nn_t nn)
{
    if (nn)
    {
        delete nn->impl;
        delete nn;
    }
    return OMPLCU_OK;
}

int omplcu_nn_set_id_base(omplcu_nn_t nn, uint64_t id_base)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->set_id_base(id_base);
    });
}

int omplcu_nn_add(omplcu_nn_t nn, const double *rows, size_t n, uint64_t *first_id)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        uint64_t id = nn->impl->add(rows, n);
        if (first_id)
            *first_id = id;
    });
}

int omplcu_nn_remove(omplcu_nn_t nn, uint64_t id)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->remove(id);
    });
}

int omplcu_nn_remove_batch(omplcu_nn_t nn, const uint64_t *ids, size_t n, size_t *removed)
{
    return guarded([&] {
        REQUIRE(nn && (ids || n == 0), "omplcu_nn_remove_batch: NULL argument");
        const size_t r = nn->impl->remove_batch(ids, n);
        if (removed)
            *removed = r;
    });
}

int omplcu_nn_clear(omplcu_nn_t nn)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->clear();
    });
}

int omplcu_nn_size(omplcu_nn_t nn, size_t *n)
{
    return guarded([&] {
        REQUIRE(nn && n, "omplcu_nn_size: NULL argument");
        *n = nn->impl->size();
    });
}

int omplcu_nn_list(omplcu_nn_t nn, uint64_t *ids, size_t cap, size_t *n)
{
    return guarded([&] {
        REQUIRE(nn && n, "omplcu_nn_list: NULL argument");
        *n = nn->impl->list(ids, cap);
    });
}

int omplcu_nn_get(omplcu_nn_t nn, uint64_t id, double *row)
{
    return guarded([&] {
        REQUIRE(nn && row, "omplcu_nn_get: NULL argument");
        nn->impl->get(id, row);
    });
}

int omplcu_nn_knn(omplcu_nn_t nn, const double *queries, size_t nq, size_t k, uint64_t *ids, double *dists,
                  uint32_t *counts)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->knn_host(queries, nq, k, ids, dists, counts);
    });
}

int omplcu_nn_radius(omplcu_nn_t nn, const double *queries, size_t nq, const double *radii, uint64_t **ids,
                     double **dists, uint64_t *offsets)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->radius_host(queries, nq, radii, ids, dists, offsets);
    });
}

int omplcu_nn_knn_device(omplcu_nn_t nn, const double *d_queries, size_t nq, size_t k, uint64_t *d_ids,
                         double *d_dists, void *stream)
{
    return guarded([&] {
        REQUIRE(nn, "nn is NULL");
        nn->impl->knn_device(d_queries, nq, k, d_ids, d_dists, (cudaStream_t)stream);
    });
}

int omplcu_merge_topk_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts, size_t nq, size_t k,
                             double *d_out_dists, uint64_t *d_out_ids, void *stream)
{
    return guarded([&] {
        need_gpu(device);
        merge_topk_device(device, d_dists, d_ids, nparts, nq * k, nq, k, d_out_dists, d_out_ids, (cudaStream_t)stream);
    });
}

int omplcu_merge_topk_strided_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts,
                                     size_t part_stride, size_t nq, size_t k, double *d_out_dists,
                                     uint64_t *d_out_ids, void *stream)
{
    return guarded([&] {
        need_gpu(device);
        merge_topk_device(device, d_dists, d_ids, nparts, part_stride, nq, k, d_out_dists, d_out_ids,
                          (cudaStream_t)stream);
    });
}

int omplcu_nn_set_option(omplcu_nn_t nn, const char *key, long value)
{
    return guarded([&] {
        REQUIRE(nn && key, "omplcu_nn_set_option: NULL argument");
        nn->impl->set_option(key, value);
    });
}

int omplcu_nn_get_stat(omplcu_nn_t nn, const char *key, long *value)
{
    return guarded([&] {
        REQUIRE(nn && key && value, "omplcu_nn_get_stat: NULL argument");
        *value = nn->impl->get_stat(key);
    });
}

/* ---- validity ---- */
static int make_checker(int device, const CheckerDesc &d, const double *env, const double *robot, const double *sph,
                        const double *box, const double *circ, omplcu_checker_t *out)
{
    return guarded([&] {
        REQUIRE(out, "checker output is NULL");
        need_gpu(device);
        auto *h = new omplcu_checker_s;
        try
        {
            h->impl = new Checker(device, d, env, robot, sph, box, circ);
        }
        catch (...)
        {
            delete h;
            throw;
        }
        *out = h;
    });
}

int omplcu_checker_create_allvalid(int device, omplcu_checker_t *out)
{
    CheckerDesc d;
    memset(&d, 0, sizeof(d));
    d.kind = OMPLCU_CHK_ALLVALID;
    return make_checker(device, d, nullptr, nullptr, nullptr, nullptr, nullptr, out);
}

int omplcu_checker_create_hypercube(int device, int dim, double edge_width, omplcu_checker_t *out)
{
    if (dim < 1 || dim > MAX_STRIDE)
    {
        g_last_error = "omplcu_checker_create_hypercube: dim out of range (1..128)";
        return OMPLCU_ERR_INVALID;
    }
    CheckerDesc d;
    memset(&d, 0, sizeof(d));
    d.kind = OMPLCU_CHK_HYPERCUBE;
    d.dim = dim;
    d.edge_width = edge_width;
    return make_checker(device, d, nullptr, nullptr, nullptr, nullptr, nullptr, out);
}

int omplcu_checker_create_chain(int device, int nlinks, double link_length, const double *env, int nenv,
                                omplcu_checker_t *out)
{
    if (nlinks < 1 || nlinks > MAX_CHAIN_LINKS || nenv < 0 || (nenv > 0 && !env))
    {
        g_last_error = "omplcu_checker_create_chain: links must be 1..128 and env non-NULL when nenv > 0";
        return OMPLCU_ERR_INVALID;
    }
    CheckerDesc d;
    memset(&d, 0, sizeof(d));
    d.kind = OMPLCU_CHK_CHAIN;
    d.nlinks = nlinks;
    d.link_length = link_length;
    d.nenv = nenv;
    return make_checker(device, d, env, nullptr, nullptr, nullptr, nullptr, out);
}

int omplcu_checker_create_se3(int device, const double *robot, int nrobot, const double *sph, int nsph,
                              const double *box, int nbox, omplcu_checker_t *out)
{
    if (nrobot < 1 || !robot || nsph < 0 || nbox < 0 || (nsph > 0 && !sph) || (nbox > 0 && !box))
    {
        g_last_error = "omplcu_checker_create_se3: need >= 1 robot sphere and non-NULL obstacle arrays";
        return OMPLCU_ERR_INVALID;
    }
    for (int i = 0; i < nrobot; ++i)
        if (!(robot[4 * i + 3] > 0.0))
        {
            g_last_error = "omplcu_checker_create_se3: robot sphere radius must be > 0";
            return OMPLCU_ERR_INVALID;
        }
    CheckerDesc d;
    memset(&d, 0, sizeof(d));
    d.kind = OMPLCU_CHK_SE3;
    d.nrobot = nrobot;
    d.nsph = nsph;
    d.nbox = nbox;
    return make_checker(device, d, nullptr, robot, sph, box, nullptr, out);
}

int omplcu_checker_create_circles(int device, const double *circ, int n, omplcu_checker_t *out)
{
    if (n < 0 || (n > 0 && !circ))
    {
        g_last_error = "omplcu_checker_create_circles: circ is NULL";
        return OMPLCU_ERR_INVALID;
    }
    CheckerDesc d;
    memset(&d, 0, sizeof(d));
    d.kind = OMPLCU_CHK_CIRCLES;
    d.ncirc = n;
    return make_checker(device, d, nullptr, nullptr, nullptr, nullptr, circ, out);
}

int omplcu_checker_destroy(omplcu_checker_t c)
{
    if (c)
    {
        delete c->impl;
        delete c;
    }
    return OMPLCU_OK;
}

int omplcu_checker_get_stat(omplcu_checker_t c, const char *key, long *value)
{
    return guarded([&] {
        REQUIRE(c && key && value, "omplcu_checker_get_stat: NULL argument");
        *value = c->impl->get_stat(key);
    });
}

int omplcu_check_states(omplcu_checker_t c, int stride, const double *states, size_t n, uint8_t *valid)
{
    return guarded([&] {
        REQUIRE(c, "checker is NULL");
        c->impl->check_states_host(stride, states, n, valid);
    });
}

int omplcu_check_motions(omplcu_checker_t c, omplcu_space_t sp, const double *s1, const double *s2, size_t n,
                         uint8_t *valid, double *last_t, double *last_state)
{
    return guarded([&] {
        REQUIRE(c && sp, "checker or space is NULL");
        c->impl->check_motions_host(sp->desc, s1, s2, n, valid, last_t, last_state);
    });
}

int omplcu_check_states_device(omplcu_checker_t c, int stride, const double *d_states, size_t n, uint8_t *d_valid,
                               void *stream)
{
    return guarded([&] {
        REQUIRE(c, "checker is NULL");
        c->impl->check_states_device(stride, d_states, n, d_valid, (cudaStream_t)stream);
    });
}

int omplcu_check_motions_device(omplcu_checker_t c, omplcu_space_t sp, const double *d_s1, const double *d_s2,
                                size_t n, uint8_t *d_valid, double *d_last_t, double *d_last_state, void *stream)
{
    return guarded([&] {
        REQUIRE(c && sp, "checker or space is NULL");
        c->impl->check_motions_device(sp->desc, d_s1, d_s2, n, d_valid, d_last_t, d_last_state, (cudaStream_t)stream);
    });
}

int omplcu_check_states_multi(const omplcu_checker_t *checkers, int nchk, int stride, const double *states, size_t n,
                              uint8_t *valid)
{
    return guarded([&] {
        REQUIRE(checkers && nchk >= 1, "omplcu_check_states_multi: no checkers");
        for (int i = 0; i < nchk; ++i)
            REQUIRE(checkers[i], "checker is NULL");
        REQUIRE(n == 0 || (states && valid), "omplcu_check_states: NULL buffer");
        split_over(nchk, n, [&](int i, size_t o, size_t cn) {
            checkers[i]->impl->check_states_host(stride, states + o * (size_t)stride, cn, valid + o);
        });
    });
}

int omplcu_check_motions_multi(const omplcu_checker_t *checkers, int nchk, omplcu_space_t sp, const double *s1,
                               const double *s2, size_t n, uint8_t *valid, double *last_t, double *last_state)
{
    return guarded([&] {
        REQUIRE(checkers && nchk >= 1 && sp, "omplcu_check_motions_multi: no checkers or no space");
        for (int i = 0; i < nchk; ++i)
            REQUIRE(checkers[i], "checker is NULL");
        REQUIRE(n == 0 || (s1 && s2 && valid), "omplcu_check_motions: NULL buffer");
        const size_t st = (size_t)sp->desc.stride;
        split_over(nchk, n, [&](int i, size_t o, size_t cn) {
            checkers[i]->impl->check_motions_host(sp->desc, s1 + o * st, s2 + o * st, cn, valid + o,
                                                  last_t ? last_t + o : nullptr,
                                                  last_state ? last_state + o * st : nullptr);
        });
    });
}

int omplcu_check_motions_device_noalloc(omplcu_checker_t c, omplcu_space_t sp, const double *d_s1,
                                        const double *d_s2, size_t n, uint8_t *d_valid, double *d_last_t,
                                        double *d_last_state, void *stream)
{
    return guarded([&] {
        REQUIRE(c && sp, "checker or space is NULL");
        c->impl->check_motions_device_noalloc(sp->desc, d_s1, d_s2, n, d_valid, d_last_t, d_last_state,
                                              (cudaStream_t)stream);
    });
}
}
