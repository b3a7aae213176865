// ref_nn.cpp -- C wrapper around the reference's OWN nearest-neighbour headers.
// TEST INFRASTRUCTURE ONLY.  Compiles NearestNeighborsLinear.h and
// NearestNeighborsGNAT.h UNMODIFIED from /root/reference/src (include path
// only; no reference source is copied into this repo).  Elements are point
// indices (like PRM's Vertex = unsigned long, prm/PRM.h:125); the distance
// function is a std::function ending in the oracle's space distance, which is
// how planners wire it (prm/PRM.h:345-348).
//
// Used (a) to validate the oracle's kNN restatement and (b) as the timed
// "reference" CPU baseline (GNAT, threaded const queries).
#include "ompl/datastructures/NearestNeighborsLinear.h"
#include "ompl/datastructures/NearestNeighborsGNAT.h"
#include "ompl_oracle.h"

#include <chrono>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

// RNG's out-of-line constructor lives in util/src/RandomNumbers.cpp, which
// needs Boost; GreedyKCenters only needs a seeded generator.
ompl::RNG::RNG() : localSeed_(1), generator_(1)
{
}

namespace
{
    struct RefNN
    {
        orc_space sp;
        std::vector<double> pts;    // row-major, stride sp.stride, indexed by element id
        std::vector<double> qbuf;   // scratch rows for queries appended after pts (per-thread slots)
        std::unique_ptr<ompl::NearestNeighbors<unsigned long>> nn;
        int kind;
    };
}

extern "C" {

// kind: 0 Linear, 1 GNAT (default parameters), 2 GNAT(4,2,6,5,5) as in tests/datastructures/nearestneighbors.cpp:88-96
void *ref_nn_create(int kind, const orc_space *sp)
{
    auto *h = new RefNN;
    h->sp = *sp;
    h->kind = kind;
    if (kind == 0)
        h->nn.reset(new ompl::NearestNeighborsLinear<unsigned long>());
    else if (kind == 1)
        h->nn.reset(new ompl::NearestNeighborsGNAT<unsigned long>());
    else
        h->nn.reset(new ompl::NearestNeighborsGNAT<unsigned long>(4, 2, 6, 5, 5));
    return h;
}

void ref_nn_destroy(void *hv)
{
    delete static_cast<RefNN *>(hv);
}

static void bind(RefNN *h, const double *base)
{
    const orc_space *sp = &h->sp;
    size_t st = (size_t)sp->stride;
    h->nn->setDistanceFunction([sp, base, st](const unsigned long &a, const unsigned long &b) {
        return orc_distance(sp, base + a * st, base + b * st);
    });
}

// Bulk add(vector): ids are first..first+n-1.  All queries must be passed via
// ref_nn_knn/ref_nn_radius, which place the query row in a reserved slot.
void ref_nn_add(void *hv, const double *rows, size_t n, size_t reserve_query_slots)
{
    auto *h = static_cast<RefNN *>(hv);
    size_t st = (size_t)h->sp.stride;
    size_t old = h->pts.size() / st;
    // keep reserve_query_slots rows free at the end of the buffer for queries
    h->pts.reserve((old + n + reserve_query_slots) * st);
    h->pts.insert(h->pts.end(), rows, rows + n * st);
    bind(h, h->pts.data());
    std::vector<unsigned long> ids(n);
    for (size_t i = 0; i < n; ++i)
        ids[i] = old + i;
    h->nn->add(ids);
}

int ref_nn_remove(void *hv, unsigned long id)
{
    auto *h = static_cast<RefNN *>(hv);
    return h->nn->remove(id) ? 1 : 0;
}

size_t ref_nn_size(void *hv)
{
    return static_cast<RefNN *>(hv)->nn->size();
}

size_t ref_nn_list(void *hv, unsigned long *out, size_t cap)
{
    std::vector<unsigned long> v;
    static_cast<RefNN *>(hv)->nn->list(v);
    for (size_t i = 0; i < v.size() && i < cap; ++i)
        out[i] = v[i];
    return v.size();
}

// Batched nearestK over nq queries with nthreads threads (const queries are
// thread safe on Linear and GNAT).  out_idx rows hold kk = min(k,size) ids.
// Returns seconds spent inside the query loop.
double ref_nn_knn(void *hv, const double *q, size_t nq, size_t k, unsigned long *out_idx, int nthreads)
{
    auto *h = static_cast<RefNN *>(hv);
    size_t st = (size_t)h->sp.stride;
    size_t npts = h->pts.size() / st;
    if (nthreads < 1)
        nthreads = 1;
    // query rows live after the points: slot t belongs to thread t
    h->pts.resize((npts + (size_t)nthreads) * st);
    bind(h, h->pts.data());
    size_t kk = std::min(k, h->nn->size());
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](int t) {
        std::vector<unsigned long> nbh;
        double *slot = h->pts.data() + (npts + (size_t)t) * st;
        for (size_t i = nq * (size_t)t / (size_t)nthreads; i < nq * (size_t)(t + 1) / (size_t)nthreads; ++i)
        {
            std::memcpy(slot, q + i * st, st * sizeof(double));
            h->nn->nearestK(npts + (size_t)t, k, nbh);
            for (size_t j = 0; j < kk && j < nbh.size(); ++j)
                out_idx[i * kk + j] = nbh[j];
        }
    };
    if (nthreads == 1)
        work(0);
    else
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t)
            th.emplace_back(work, t);
        for (auto &x : th)
            x.join();
    }
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    h->pts.resize(npts * st);
    return dt;
}

// Single nearestR; returns count, writes up to cap ids.
size_t ref_nn_radius(void *hv, const double *q, double radius, unsigned long *out_idx, size_t cap)
{
    auto *h = static_cast<RefNN *>(hv);
    size_t st = (size_t)h->sp.stride;
    size_t npts = h->pts.size() / st;
    h->pts.resize((npts + 1) * st);
    bind(h, h->pts.data());
    std::memcpy(h->pts.data() + npts * st, q, st * sizeof(double));
    std::vector<unsigned long> nbh;
    h->nn->nearestR(npts, radius, nbh);
    for (size_t j = 0; j < nbh.size() && j < cap; ++j)
        out_idx[j] = nbh[j];
    h->pts.resize(npts * st);
    return nbh.size();
}

// nearest(); returns -1 and copies the exception text when empty.
long ref_nn_nearest(void *hv, const double *q, char *err, size_t errcap)
{
    auto *h = static_cast<RefNN *>(hv);
    size_t st = (size_t)h->sp.stride;
    size_t npts = h->pts.size() / st;
    h->pts.resize((npts + 1) * st);
    bind(h, h->pts.data());
    std::memcpy(h->pts.data() + npts * st, q, st * sizeof(double));
    long r = -1;
    try
    {
        r = (long)h->nn->nearest(npts);
    }
    catch (const std::exception &e)
    {
        if (err && errcap)
        {
            std::strncpy(err, e.what(), errcap - 1);
            err[errcap - 1] = 0;
        }
    }
    h->pts.resize(npts * st);
    return r;
}
}
