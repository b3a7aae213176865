/*
 * NearestNeighborsCUDA.h -- ompl::NearestNeighbors<_T> with the element set
 * mirrored on a B200 (libomplcu.so, C ABI in include/omplcu.h).
 *
 * Drop-in for NearestNeighborsLinear / NearestNeighborsGNAT behind the
 * reference interface datastructures/NearestNeighbors.h:46-115: same virtuals,
 * same results as NearestNeighborsLinear (NearestNeighborsLinear.h:98-142) with
 * ties broken by insertion order, same exception text on an empty structure
 * (NearestNeighborsLinear.h:115).  Header-only; needs only the reference's
 * NearestNeighbors.h / Exception.h (no Boost), so it also builds in this
 * repository's test tree.
 *
 * The reference hands a structure nothing but an opaque distance function.
 * A GPU structure needs coordinates, so two extra setters exist:
 *   setSpace(desc)        which metric: RealVector / SO2 / SO3 / compound ...
 *   setRowFunction(f)     f(elem, double *row) writes the element's state in
 *                         StateSpace::copyToReals order (StateSpace.cpp:329-335)
 * ompl_b200/OmplBridge.h derives both from a base::StateSpacePtr and a
 * State accessor.  The distance function is kept for getDistanceFunction()
 * and is NOT called by queries.
 *
 * Batch entry points (what PRM/FMT-star/BIT-star batch call sites use):
 *   nearestKBatch / nearestRBatch.
 */
#ifndef OMPL_B200_NEAREST_NEIGHBORS_CUDA_H
#define OMPL_B200_NEAREST_NEIGHBORS_CUDA_H

#include <algorithm>
#include <cstdint>
#include <functional>
#include <limits>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "ompl/datastructures/NearestNeighbors.h"
#include "ompl/util/Exception.h"
#include "omplcu.h"

namespace ompl
{
    namespace b200
    {
        /** What the device must know about the metric space. */
        struct SpaceDescription
        {
            std::vector<omplcu_component> components;
            bool compound{false};

            unsigned int stride() const
            {
                unsigned int s = 0;
                for (const auto &c : components)
                    s += c.dim;
                return s;
            }
            static omplcu_component component(int type, int dim, double weight, double lvs, unsigned factor = 1,
                                              double param = 0.)
            {
                omplcu_component c{};
                c.type = type;
                c.dim = dim;
                c.weight = weight;
                c.lvs = lvs;
                c.factor = factor;
                c.param = param;
                return c;
            }
            /** KinematicChainSpace of demos/KinematicChain.h:88-191: a RealVectorStateSpace subclass whose
                distance (:105-121) and interpolate (:150-176) are overridden.  Its getType() is still
                STATE_SPACE_REAL_VECTOR, so it can only be described explicitly (OmplBridge.h: describe()
                refuses subclasses it cannot vouch for). */
            static SpaceDescription kinematicChain(int links, double linkLength, double lvs = 1.0, unsigned factor = 1)
            {
                SpaceDescription d;
                d.components.push_back(component(OMPLCU_CHAIN, links, 1.0, lvs, factor, linkLength));
                return d;
            }
            /** RealVectorStateSpace(dim) with longest valid segment lvs (only used by motion checks). */
            static SpaceDescription realVector(int dim, double lvs = 1.0)
            {
                SpaceDescription d;
                d.components.push_back(component(OMPLCU_RV, dim, 1.0, lvs));
                return d;
            }
        };

        inline void check(int rc)
        {
            if (rc != OMPLCU_OK)
                throw ompl::Exception(std::string(omplcu_last_error()));
        }

        /** RAII owner of an omplcu_space_t. */
        class SpaceHandle
        {
        public:
            SpaceHandle() = default;
            explicit SpaceHandle(const SpaceDescription &d)
            {
                reset(d);
            }
            SpaceHandle(const SpaceHandle &) = delete;
            SpaceHandle &operator=(const SpaceHandle &) = delete;
            ~SpaceHandle()
            {
                omplcu_space_destroy(h_);
            }
            void reset(const SpaceDescription &d)
            {
                omplcu_space_destroy(h_);
                h_ = nullptr;
                check(omplcu_space_create(d.components.data(), (int)d.components.size(), d.compound ? 1 : 0, &h_));
                stride_ = d.stride();
            }
            omplcu_space_t get() const
            {
                return h_;
            }
            unsigned int stride() const
            {
                return stride_;
            }

        private:
            omplcu_space_t h_{nullptr};
            unsigned int stride_{0};
        };
    }  // namespace b200

    template <typename _T>
    class NearestNeighborsCUDA : public NearestNeighbors<_T>
    {
    public:
        using RowFunction = std::function<void(const _T &, double *)>;

        NearestNeighborsCUDA() = default;
        explicit NearestNeighborsCUDA(const b200::SpaceDescription &space, RowFunction rows = RowFunction(),
                                      int device = 0)
          : rowFun_(std::move(rows)), device_(device)
        {
            setSpace(space);
        }
        NearestNeighborsCUDA(const NearestNeighborsCUDA &) = delete;
        NearestNeighborsCUDA &operator=(const NearestNeighborsCUDA &) = delete;

        ~NearestNeighborsCUDA() override
        {
            omplcu_nn_destroy(nn_);
        }

        /** Must be called while the structure is empty. */
        void setSpace(const b200::SpaceDescription &space)
        {
            if (nn_ != nullptr && size() != 0)
                throw Exception("NearestNeighborsCUDA: setSpace() on a non-empty structure");
            omplcu_nn_destroy(nn_);
            nn_ = nullptr;
            byId_.clear();
            idOffset_ = kUnset;
            idsOf_.clear();
            space_.reset(space);
            desc_ = space;
            haveDesc_ = true;
            if (devices_.empty())
                b200::check(omplcu_nn_create(space_.get(), device_, &nn_));
            else
                b200::check(omplcu_nn_create_sharded(space_.get(), devices_.data(), (int)devices_.size(), &nn_));
        }
        void setRowFunction(RowFunction f)
        {
            rowFun_ = std::move(f);
        }
        /** GPU of the structure; takes effect at the next setSpace() or, if the space is known and the
            structure is empty, at once. */
        void setDevice(int device)
        {
            device_ = device;
            devices_.clear();
            if (haveDesc_)
                setSpace(b200::SpaceDescription(desc_));
        }
        /** Shard the states over several GPUs of this process (omplcu_nn_create_sharded):
            every query is answered by all of them and merged by (distance, insertion order), so results are
            those of the single-device structure.  Same timing rule as setDevice(). */
        void setDevices(const std::vector<int> &devices)
        {
            devices_ = devices;
            if (haveDesc_)
                setSpace(b200::SpaceDescription(desc_));
        }

        bool reportsSortedResults() const override
        {
            return true;  // FMT relies on it (fmt/src/FMT.cpp:105-110)
        }

        void clear() override
        {
            if (nn_ != nullptr)
                b200::check(omplcu_nn_clear(nn_));
            byId_.clear();
            idOffset_ = kUnset;  // ids keep increasing across clear(): re-based at the next add
            idsOf_.clear();
        }

        void add(const _T &data) override
        {
            ready();
            row_.resize(space_.stride());
            rowFun_(data, row_.data());
            std::uint64_t id = 0;
            b200::check(omplcu_nn_add(nn_, row_.data(), 1, &id));
            remember(id, data);
        }

        /** One bulk upload (NearestNeighbors.h:81-85 loops over add(x)). */
        void add(const std::vector<_T> &data) override
        {
            if (data.empty())
                return;
            ready();
            const std::size_t st = space_.stride();
            std::vector<double> rows(data.size() * st);
            for (std::size_t i = 0; i < data.size(); ++i)
                rowFun_(data[i], rows.data() + i * st);
            std::uint64_t first = 0;
            b200::check(omplcu_nn_add(nn_, rows.data(), data.size(), &first));
            for (std::size_t i = 0; i < data.size(); ++i)
                remember(first + i, data[i]);
        }

        /** Removes the most recently added element equal to data
            (NearestNeighborsLinear::remove searches from the back, :86-96). */
        bool remove(const _T &data) override
        {
            auto range = idsOf_.equal_range(data);
            if (range.first == range.second)
                return false;
            auto best = range.first;
            for (auto it = range.first; it != range.second; ++it)
                if (it->second > best->second)
                    best = it;
            const std::uint64_t id = best->second;
            const int rc = omplcu_nn_remove(nn_, id);
            if (rc == OMPLCU_ERR_NOT_FOUND)
                return false;
            b200::check(rc);
            idsOf_.erase(best);
            byId_[id - idOffset_] = _T();
            return true;
        }

        /** Bulk remove (BIT* prunes samples in batches: ImplicitGraph.cpp, pruneSamples): for every element
            the most recently added equal one, as remove(); one device call.  Returns how many were removed. */
        std::size_t removeBatch(const std::vector<_T> &data)
        {
            std::vector<std::uint64_t> ids;
            ids.reserve(data.size());
            for (const _T &d : data)
            {
                auto range = idsOf_.equal_range(d);
                if (range.first == range.second)
                    continue;
                auto best = range.first;
                for (auto it = range.first; it != range.second; ++it)
                    if (it->second > best->second)
                        best = it;
                ids.push_back(best->second);
                byId_[best->second - idOffset_] = _T();
                idsOf_.erase(best);
            }
            std::size_t removed = 0;
            b200::check(omplcu_nn_remove_batch(nn_, ids.data(), ids.size(), &removed));
            return removed;
        }

        _T nearest(const _T &data) const override
        {
            if (size() == 0)
                throw Exception("No elements found in nearest neighbors data structure");
            std::vector<_T> nbh;
            nearestK(data, 1, nbh);
            return nbh[0];
        }

        void nearestK(const _T &data, std::size_t k, std::vector<_T> &nbh) const override
        {
            std::vector<std::vector<_T>> out;
            nearestKBatch(std::vector<_T>(1, data), k, out);
            nbh.swap(out[0]);
        }

        void nearestR(const _T &data, double radius, std::vector<_T> &nbh) const override
        {
            std::vector<std::vector<_T>> out;
            nearestRBatch(std::vector<_T>(1, data), std::vector<double>(1, radius), out);
            nbh.swap(out[0]);
        }

        std::size_t size() const override
        {
            std::size_t n = 0;
            if (nn_ != nullptr)
                b200::check(omplcu_nn_size(nn_, &n));
            return n;
        }

        /** Elements in insertion order (NearestNeighborsLinear::list, :149-152). */
        void list(std::vector<_T> &data) const override
        {
            data.clear();
            const std::size_t n = size();
            std::vector<std::uint64_t> ids(n);
            std::size_t got = 0;
            if (n != 0)
                b200::check(omplcu_nn_list(nn_, ids.data(), n, &got));
            data.reserve(got);
            for (std::size_t i = 0; i < got; ++i)
                data.push_back(byId_[ids[i] - idOffset_]);
        }

        /** f(lo, hi) over [0, n) on up to 8 host threads (one below 8192 items: a thread costs more than it saves). */
        template <typename F>
        static void parallelFor(std::size_t n, F f)
        {
            const unsigned hw = std::thread::hardware_concurrency();
            const std::size_t nt = std::min<std::size_t>({(std::size_t)(hw ? hw : 1u), (std::size_t)8, n / 8192});
            if (nt <= 1)
            {
                f((std::size_t)0, n);
                return;
            }
            std::vector<std::thread> th;
            th.reserve(nt - 1);
            for (std::size_t t = 1; t < nt; ++t)
                th.emplace_back(f, n * t / nt, n * (t + 1) / nt);
            f((std::size_t)0, n / nt);
            for (auto &x : th)
                x.join();
        }

        // ---- batch-submit entry points ----
        /** out[i] = the min(k, size()) nearest elements of queries[i], sorted by (distance, insertion order).
            dists (optional) receives the exact distances, row-major with out[i].size() entries per row. */
        void nearestKBatch(const std::vector<_T> &queries, std::size_t k, std::vector<std::vector<_T>> &out,
                           std::vector<double> *dists = nullptr) const
        {
            ready();
            const std::size_t nq = queries.size(), st = space_.stride();
            out.assign(nq, std::vector<_T>());
            if (dists != nullptr)
                dists->clear();
            if (nq == 0 || k == 0)
                return;
            // call scratch lives per host thread and only grows: fresh 9 MB vectors cost a millisecond of page faults and
            // zero-filling per 65536-query call
            static thread_local std::vector<double> rows, d;
            static thread_local std::vector<std::uint64_t> ids;
            static thread_local std::vector<std::uint32_t> counts;
            if (rows.size() < nq * st)
                rows.resize(nq * st);
            // (thread_local objects are not captured by a lambda: a worker would see its OWN empty vector, so the
            // workers get plain pointers)
            double *const rowp = rows.data();
            // large batches: the host side (copyToReals of every query, one result vector per query) is spread over
            // a few threads; the row function only reads the states
            parallelFor(nq, [&, rowp](std::size_t lo, std::size_t hi) {
                for (std::size_t i = lo; i < hi; ++i)
                    rowFun_(queries[i], rowp + i * st);
            });
            if (ids.size() < nq * k)
                ids.resize(nq * k);
            if (dists != nullptr && d.size() < nq * k)
                d.resize(nq * k);
            if (counts.size() < nq)
                counts.resize(nq);
            b200::check(omplcu_nn_knn(nn_, rows.data(), nq, k, ids.data(), dists != nullptr ? d.data() : nullptr,
                                      counts.data()));
            const std::size_t kk = counts[0];  // min(k, size()): the same for every query of a call
            if (dists != nullptr)
                dists->resize(nq * kk);
            const std::uint64_t *const idp = ids.data();
            const double *const dp = d.data();
            parallelFor(nq, [&, idp, dp](std::size_t lo, std::size_t hi) {
                for (std::size_t i = lo; i < hi; ++i)
                {
                    out[i].resize(kk);
                    for (std::size_t j = 0; j < kk; ++j)
                    {
                        out[i][j] = byId_[idp[i * k + j] - idOffset_];
                        if (dists != nullptr)
                            (*dists)[i * kk + j] = dp[i * k + j];
                    }
                }
            });
        }

        /** The same without a vector per query: flat[i * kk + j] is the j-th neighbour of queries[i], kk = min(k, size())
            is returned.  What a batch call site that only walks the neighbours wants (no 65536 small allocations). */
        std::size_t nearestKBatchFlat(const std::vector<_T> &queries, std::size_t k, std::vector<_T> &flat,
                                      std::vector<double> *dists = nullptr) const
        {
            ready();
            const std::size_t nq = queries.size(), st = space_.stride();
            flat.clear();
            if (dists != nullptr)
                dists->clear();
            if (nq == 0 || k == 0)
                return 0;
            static thread_local std::vector<double> rows, d;
            static thread_local std::vector<std::uint64_t> ids;
            static thread_local std::vector<std::uint32_t> counts;
            if (rows.size() < nq * st)
                rows.resize(nq * st);
            double *const rowp = rows.data();
            parallelFor(nq, [&, rowp](std::size_t lo, std::size_t hi) {
                for (std::size_t i = lo; i < hi; ++i)
                    rowFun_(queries[i], rowp + i * st);
            });
            if (ids.size() < nq * k)
                ids.resize(nq * k);
            if (dists != nullptr && d.size() < nq * k)
                d.resize(nq * k);
            if (counts.size() < nq)
                counts.resize(nq);
            b200::check(omplcu_nn_knn(nn_, rows.data(), nq, k, ids.data(), dists != nullptr ? d.data() : nullptr,
                                      counts.data()));
            const std::size_t kk = counts[0];
            flat.resize(nq * kk);
            if (dists != nullptr)
                dists->resize(nq * kk);
            const std::uint64_t *const idp = ids.data();
            const double *const dp = d.data();
            parallelFor(nq, [&, idp, dp](std::size_t lo, std::size_t hi) {
                for (std::size_t i = lo; i < hi; ++i)
                    for (std::size_t j = 0; j < kk; ++j)
                    {
                        flat[i * kk + j] = byId_[idp[i * k + j] - idOffset_];
                        if (dists != nullptr)
                            (*dists)[i * kk + j] = dp[i * k + j];
                    }
            });
            return kk;
        }

        /** out[i] = all elements within radii[i] of queries[i] (inclusive), sorted. */
        void nearestRBatch(const std::vector<_T> &queries, const std::vector<double> &radii,
                           std::vector<std::vector<_T>> &out) const
        {
            ready();
            const std::size_t nq = queries.size(), st = space_.stride();
            if (radii.size() != nq)
                throw Exception("NearestNeighborsCUDA::nearestRBatch: one radius per query");
            out.assign(nq, std::vector<_T>());
            if (nq == 0)
                return;
            std::vector<double> rows(nq * st);
            for (std::size_t i = 0; i < nq; ++i)
                rowFun_(queries[i], rows.data() + i * st);
            std::vector<std::uint64_t> offs(nq + 1);
            std::uint64_t *ids = nullptr;
            b200::check(omplcu_nn_radius(nn_, rows.data(), nq, radii.data(), &ids, nullptr, offs.data()));
            for (std::size_t i = 0; i < nq; ++i)
            {
                out[i].reserve(offs[i + 1] - offs[i]);
                for (std::uint64_t j = offs[i]; j < offs[i + 1]; ++j)
                    out[i].push_back(byId_[ids[j] - idOffset_]);
            }
            omplcu_free(ids);
        }

        omplcu_nn_t handle() const
        {
            return nn_;
        }

    private:
        void ready() const
        {
            if (nn_ == nullptr)
                throw Exception("NearestNeighborsCUDA: setSpace() was not called");
            if (!rowFun_)
                throw Exception("NearestNeighborsCUDA: setRowFunction() was not called");
        }
        // Ids are insertion sequence numbers, handed out consecutively by the device structure: the
        // element of an id is a vector lookup (655 360 of them per headline batch), not a hash probe.
        void remember(std::uint64_t id, const _T &e)
        {
            if (idOffset_ == kUnset)
                idOffset_ = id;
            const std::uint64_t at = id - idOffset_;
            if (byId_.size() <= at)
                byId_.resize(at + 1);
            byId_[at] = e;
            idsOf_.emplace(e, id);
        }

        RowFunction rowFun_;
        int device_{0};
        b200::SpaceHandle space_;
        omplcu_nn_t nn_{nullptr};
        static constexpr std::uint64_t kUnset = ~std::uint64_t(0);
        std::vector<int> devices_;                          // non-empty: sharded over these GPUs
        b200::SpaceDescription desc_;
        bool haveDesc_{false};
        std::vector<_T> byId_;                              // id - idOffset_ -> element (_T() once removed)
        std::uint64_t idOffset_{kUnset};
        // element -> ids, for remove(); needs std::hash<_T> and operator== (pointers, integers such as
        // PRM's Vertex, std::shared_ptr<...> as in BIT* all have them)
        std::unordered_multimap<_T, std::uint64_t> idsOf_;
        mutable std::vector<double> row_;
    };
}  // namespace ompl

#endif
