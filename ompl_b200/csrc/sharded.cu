// sharded.cu -- one roadmap sharded over several GPUs of ONE process, behind the same omplcu_nn_t
// (omplcu_nn_create_sharded).  SURVEY 8(b) / 8(e): a C++ OMPL process that holds a
// NearestNeighborsCUDA<_T> (what replaces datastructures/NearestNeighborsGNAT.h:72 at the call sites
// geometric/planners/prm/ConnectionStrategy.h:77-81) shards its states across the GPUs of the box.
//
//   states    split by insertion index: a bulk add is cut into ndev contiguous blocks, a small add goes
//             to the least-loaded shard.  Ids stay GLOBAL insertion sequence numbers; inside a shard they
//             increase with the slot, so "ties by insertion index" holds per shard and, because the order
//             (distance, id) is total, for the merged result too.
//   kNN       queries land on the first device (one H2D), go to the other devices over NVLink
//             (cudaMemcpyPeerAsync), every shard answers them on its own GPU from its own host thread
//             (local exact top-k), the per-shard (distance, id) lists are peer-copied into one
//             [part][query][k] buffer on the first device, merge_topk_kernel selects the k best, ONE D2H
//             (FMT*-sized k: merge_sorted_kernel, ranks by binary search in the sorted lists, and the
//             batch goes in pieces that bound the gather buffer).
//             It is the all-gather + merge of 8(e) with a single consumer: in one process only the
//             caller's device needs the gathered lists, so the exchange is ndev - 1 peer copies into
//             disjoint slices instead of a collective.
//   radius    every shard answers (sorted CSR lists), the lists are merged per query by (distance, id).
//   validity  omplcu_check_states_multi / omplcu_check_motions_multi split a batch by segment over
//             per-device checkers, no exchange (validity.cu).
#include <algorithm>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <chrono>
#include <map>
#include <memory>
#include <thread>

#include "nn.cuh"

namespace omplcu
{
    namespace
    {
        // One host thread per shard: NNIndex calls synchronise their stream, so shards run side by side
        // only when they are driven from different threads.
        class Worker
        {
        public:
            Worker() : th_([this] { loop(); })
            {
            }
            ~Worker()
            {
                {
                    std::lock_guard<std::mutex> lk(mu_);
                    stop_ = true;
                }
                cv_.notify_all();
                th_.join();
            }
            void submit(std::function<void()> f)
            {
                std::lock_guard<std::mutex> lk(mu_);
                task_ = std::move(f);
                busy_ = true;
                err_ = nullptr;
                cv_.notify_all();
            }
            // rethrows what the task threw
            void wait()
            {
                std::unique_lock<std::mutex> lk(mu_);
                done_.wait(lk, [this] { return !busy_; });
                if (err_)
                {
                    std::exception_ptr e = err_;
                    err_ = nullptr;
                    std::rethrow_exception(e);
                }
            }

        private:
            void loop()
            {
                std::unique_lock<std::mutex> lk(mu_);
                for (;;)
                {
                    cv_.wait(lk, [this] { return stop_ || (busy_ && task_); });
                    if (stop_)
                        return;
                    std::function<void()> f = std::move(task_);
                    task_ = nullptr;
                    lk.unlock();
                    std::exception_ptr e = nullptr;
                    try
                    {
                        f();
                    }
                    catch (...)
                    {
                        e = std::current_exception();
                    }
                    lk.lock();
                    err_ = e;
                    busy_ = false;
                    done_.notify_all();
                }
            }
            std::mutex mu_;
            std::condition_variable cv_, done_;
            std::function<void()> task_;
            bool busy_ = false, stop_ = false;
            std::exception_ptr err_ = nullptr;
            std::thread th_;
        };

        struct Shard
        {
            int device = 0;
            std::unique_ptr<NNIndex> index;
            std::unique_ptr<Worker> worker;
            DevBuf<double> q, out_d;     // queries and local results on this shard's device
            DevBuf<uint64_t> out_i;
            DevBuf<double> blk_d, mrg_d;  // knn_host_distributed: every shard's lists for this shard's query block, merged block
            DevBuf<uint64_t> blk_i, mrg_i;
            cudaEvent_t done = nullptr;  // local results are in the gather buffer
        };
    }  // namespace

    class ShardedNN : public NNApi
    {
    public:
        ShardedNN(const SpaceDesc &sp, const int *devices, int ndev);
        ~ShardedNN() override;
        void set_id_base(uint64_t base) override;
        uint64_t add(const double *rows, size_t n) override;
        void remove(uint64_t id) override;
        size_t remove_batch(const uint64_t *ids, size_t n) override;
        void clear() override;
        size_t size() const override;
        size_t list(uint64_t *ids, size_t cap) const override;
        void get(uint64_t id, double *row) override;
        void knn_device(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                        cudaStream_t s) override;
        void knn_host(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists, uint32_t *counts) override;
        void radius_host(const double *q, size_t nq, const double *radii, uint64_t **ids, double **dists,
                         uint64_t *offsets) override;
        void set_option(const char *key, long v) override;
        long get_stat(const char *key) const override;

    private:
        int owner(uint64_t id) const;  // shard holding the block of `id`, -1 if none
        template <typename F>
        void on_all(F &&f);            // f(shard index) on every shard's worker, in parallel
        void knn_gather(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists, cudaStream_t s);
        void knn_host_distributed(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists);

        SpaceDesc sp_;
        std::vector<Shard> sh_;
        mutable std::mutex mu_, host_mu_;
        uint64_t next_id_ = 0;
        bool any_added_ = false;
        // blocks of consecutive ids: first id -> (count, shard)
        std::map<uint64_t, std::pair<uint64_t, int>> blocks_;
        // first device: staging of host queries, gathered partial results, merged results
        DevBuf<double> q0_, gat_d_, mrg_d_;
        DevBuf<uint64_t> gat_i_, mrg_i_;
        long last_merge_ns_ = 0, last_scan_phase_ns_ = 0;
        size_t gather_entries_ = (size_t)1 << 27;  // 2 GB of gathered lists
    };

    ShardedNN::ShardedNN(const SpaceDesc &sp, const int *devices, int ndev) : sp_(sp)
    {
        sh_.resize(ndev);
        for (int i = 0; i < ndev; ++i)
        {
            sh_[i].device = devices[i];
            sh_[i].index.reset(new NNIndex(sp, devices[i]));
            sh_[i].worker.reset(new Worker());
            DeviceGuard g(devices[i]);
            CUDA_CHECK(cudaEventCreateWithFlags(&sh_[i].done, cudaEventDisableTiming));
            // direct NVLink copies between the shards' devices and the first one (a missing or already
            // enabled link is not an error: cudaMemcpyPeerAsync then goes the slower way)
            for (int j = 0; j < ndev; ++j)
                if (devices[j] != devices[i])
                {
                    int can = 0;
                    if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can)
                        cudaDeviceEnablePeerAccess(devices[j], 0);
                    cudaGetLastError();
                }
        }
    }

    ShardedNN::~ShardedNN()
    {
        for (auto &s : sh_)
        {
            s.worker.reset();  // joins the thread before the index goes
            DeviceGuard g(s.device);
            if (s.done)
                cudaEventDestroy(s.done);
            s.q.release();
            s.out_d.release();
            s.out_i.release();
        }
        DeviceGuard g(sh_[0].device);
        q0_.release();
        gat_d_.release();
        mrg_d_.release();
        gat_i_.release();
        mrg_i_.release();
    }

    template <typename F>
    void ShardedNN::on_all(F &&f)
    {
        for (size_t i = 0; i < sh_.size(); ++i)
            sh_[i].worker->submit([&f, i] { f((int)i); });
        std::exception_ptr first = nullptr;
        for (auto &s : sh_)
        {
            try
            {
                s.worker->wait();
            }
            catch (...)
            {
                if (!first)
                    first = std::current_exception();
            }
        }
        if (first)
            std::rethrow_exception(first);
    }

    void ShardedNN::set_id_base(uint64_t base)
    {
        std::lock_guard<std::mutex> lk(mu_);
        REQUIRE(!any_added_ || size() == 0, "omplcu_nn_set_id_base: structure is not empty");
        REQUIRE(base >= next_id_ || !any_added_, "omplcu_nn_set_id_base: ids must increase");
        next_id_ = base;
    }

    size_t ShardedNN::size() const
    {
        size_t n = 0;
        for (const auto &s : sh_)
            n += s.index->size();
        return n;
    }

    int ShardedNN::owner(uint64_t id) const
    {
        auto it = blocks_.upper_bound(id);
        if (it == blocks_.begin())
            return -1;
        --it;
        return id < it->first + it->second.first ? it->second.second : -1;
    }

    uint64_t ShardedNN::add(const double *rows, size_t n)
    {
        std::lock_guard<std::mutex> lk(mu_);
        const uint64_t first = next_id_;
        if (n == 0)
            return first;
        REQUIRE(rows != nullptr, "omplcu_nn_add: rows is NULL");
        const size_t nd = sh_.size(), st = (size_t)sp_.stride;
        if (n < nd * 256)
        {
            // a small add: the least-loaded shard takes all of it
            int best = 0;
            for (size_t i = 1; i < nd; ++i)
                if (sh_[i].index->size() < sh_[best].index->size())
                    best = (int)i;
            sh_[best].index->add_at(rows, n, first);
            blocks_[first] = {n, best};
        }
        else
        {
            // a bulk add: ndev contiguous blocks, uploaded side by side (every GPU has its own PCIe link)
            on_all([&](int i) {
                const size_t b0 = n * (size_t)i / nd, b1 = n * (size_t)(i + 1) / nd;
                if (b1 > b0)
                    sh_[i].index->add_at(rows + b0 * st, b1 - b0, first + b0);
            });
            for (size_t i = 0; i < nd; ++i)
            {
                const size_t b0 = n * i / nd, b1 = n * (i + 1) / nd;
                if (b1 > b0)
                    blocks_[first + b0] = {b1 - b0, (int)i};
            }
        }
        next_id_ += n;
        any_added_ = true;
        return first;
    }

    void ShardedNN::remove(uint64_t id)
    {
        std::lock_guard<std::mutex> lk(mu_);
        const int o = owner(id);
        if (o < 0)
            throw Error(OMPLCU_ERR_NOT_FOUND, "omplcu_nn_remove: id not present");
        sh_[o].index->remove(id);
    }

    size_t ShardedNN::remove_batch(const uint64_t *ids, size_t n)
    {
        std::lock_guard<std::mutex> lk(mu_);
        std::vector<std::vector<uint64_t>> per(sh_.size());
        for (size_t i = 0; i < n; ++i)
        {
            const int o = owner(ids[i]);
            if (o >= 0)
                per[o].push_back(ids[i]);
        }
        std::vector<size_t> removed(sh_.size(), 0);
        on_all([&](int i) {
            if (!per[i].empty())
                removed[i] = sh_[i].index->remove_batch(per[i].data(), per[i].size());
        });
        size_t tot = 0;
        for (size_t r : removed)
            tot += r;
        return tot;
    }

    void ShardedNN::clear()
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &s : sh_)
            s.index->clear();
        blocks_.clear();  // ids keep increasing across clear(), as in NNIndex
    }

    size_t ShardedNN::list(uint64_t *ids, size_t cap) const
    {
        std::lock_guard<std::mutex> lk(mu_);
        std::vector<uint64_t> all;
        for (const auto &s : sh_)
        {
            const size_t n = s.index->size(), at = all.size();
            all.resize(at + n);
            const size_t got = s.index->list(all.data() + at, n);
            all.resize(at + got);
        }
        std::sort(all.begin(), all.end());  // insertion order == id order
        if (ids)
            for (size_t i = 0; i < all.size() && i < cap; ++i)
                ids[i] = all[i];
        return all.size();
    }

    void ShardedNN::get(uint64_t id, double *row)
    {
        std::lock_guard<std::mutex> lk(mu_);
        const int o = owner(id);
        if (o < 0)
            throw Error(OMPLCU_ERR_NOT_FOUND, "omplcu_nn_get: id not present");
        sh_[o].index->get(id, row);
    }

    void ShardedNN::set_option(const char *key, long v)
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (!strcmp(key, "gather_entries"))
        {
            // (distance, id) entries the gather buffer on the first device may hold: larger batches go in pieces
            gather_entries_ = (size_t)std::max<long>(v, 1);
            return;
        }
        for (auto &s : sh_)
            s.index->set_option(key, v);
    }

    long ShardedNN::get_stat(const char *key) const
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (!strcmp(key, "shards"))
            return (long)sh_.size();
        if (!strcmp(key, "last_merge_ns"))
            return last_merge_ns_;
        if (!strcmp(key, "last_scan_phase_ns"))
            return last_scan_phase_ns_;  // large host batches: H2D + per-shard kNN, wall clock over all shards
        if (!strcmp(key, "last_scan_ns"))
        {
            long m = 0;  // slowest shard's scan kernel (with option "profile")
            for (const auto &h : sh_)
                m = std::max(m, h.index->get_stat("last_scan_ns"));
            return m;
        }
        if (!strcmp(key, "launches") || !strcmp(key, "last_fallbacks") || !strcmp(key, "slots") ||
            !strcmp(key, "last_collect_redo") || !strcmp(key, "scan_ns_total") || !strcmp(key, "scan_count"))
        {
            long t = 0;
            for (const auto &s : sh_)
                t += s.index->get_stat(key);
            return t;
        }
        if (!strcmp(key, "last_scan_ns"))
        {
            long t = 0;
            for (const auto &s : sh_)
                t = std::max(t, s.index->get_stat(key));
            return t;
        }
        return sh_[0].index->get_stat(key);
    }

    // ---- queries ----
    void ShardedNN::knn_device(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                               cudaStream_t s)
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (nq == 0 || k == 0)
            return;
        REQUIRE(d_q && d_ids, "omplcu_nn_knn: NULL buffer");
        const int dev0 = sh_[0].device;
        const size_t nd = sh_.size(), st = (size_t)sp_.stride;
        {
            DeviceGuard g(dev0);
            CUDA_CHECK(cudaStreamSynchronize(s));  // the queries may still be on their way on the caller's stream
        }
        if (nd == 1)
        {
            sh_[0].index->knn_device(d_q, nq, k, d_ids, d_dists, s);
            return;
        }
        // FMT*-sized k: the gathered lists (nd * nq * k * 16 bytes) are bounded by answering the batch in pieces
        const size_t piece = std::max<size_t>(1, gather_entries_ / (nd * k));
        for (size_t q0 = 0; q0 < nq; q0 += piece)
            knn_gather(d_q + q0 * st, std::min(piece, nq - q0), k, d_ids + q0 * k, d_dists ? d_dists + q0 * k : nullptr, s);
    }

    void ShardedNN::knn_gather(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                               cudaStream_t s)
    {
        const int dev0 = sh_[0].device;
        const size_t nd = sh_.size(), st = (size_t)sp_.stride;
        const size_t part = nq * k;
        {
            DeviceGuard g(dev0);
            gat_d_.reserve(nd * part);
            gat_i_.reserve(nd * part);
            if (!d_dists)
                mrg_d_.reserve(part);
        }
        on_all([&](int i) {
            Shard &h = sh_[i];
            DeviceGuard g(h.device);
            cudaStream_t hs = h.index->stream();
            if (h.device == dev0)
            {
                // same GPU as the gather buffer: answer straight into this shard's slice
                h.index->knn_device(d_q, nq, k, gat_i_.p + i * part, gat_d_.p + i * part, hs);
                return;
            }
            h.q.reserve(nq * st);
            h.out_i.reserve(part);
            h.out_d.reserve(part);
            CUDA_CHECK(cudaMemcpyPeerAsync(h.q.p, h.device, d_q, dev0, nq * st * sizeof(double), hs));
            h.index->knn_device(h.q.p, nq, k, h.out_i.p, h.out_d.p, hs);
            CUDA_CHECK(cudaMemcpyPeerAsync(gat_i_.p + i * part, dev0, h.out_i.p, h.device, part * sizeof(uint64_t), hs));
            CUDA_CHECK(cudaMemcpyPeerAsync(gat_d_.p + i * part, dev0, h.out_d.p, h.device, part * sizeof(double), hs));
            CUDA_CHECK(cudaStreamSynchronize(hs));
        });
        merge_topk_device(dev0, gat_d_.p, gat_i_.p, (int)nd, part, nq, k, d_dists ? d_dists : mrg_d_.p, d_ids, s);
    }

    // Large host-pointer batches: nothing funnels through the first GPU.  Every shard fetches the queries over its own
    // PCIe link and answers them on its states; then shard i pulls every shard's lists for query block i (peer copies:
    // an all-to-all of nq * k * 16 / nd bytes per pair over NVLink), merges them and copies its block to the caller,
    // again over its own link.  One H2D of 3.7 MB to one GPU, a gather of 84 MB into it and one D2H of 10.5 MB from it
    // were 0.7 of the 2.1 ms of a 65536-query call on 8 GPUs.
    void ShardedNN::knn_host_distributed(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists)
    {
        std::lock_guard<std::mutex> lk(mu_);
        const size_t nd = sh_.size(), st = (size_t)sp_.stride, part = nq * k;
        const auto t0 = std::chrono::steady_clock::now();
        on_all([&](int i) {
            Shard &h = sh_[i];
            DeviceGuard g(h.device);
            cudaStream_t hs = h.index->stream();
            h.q.reserve(nq * st);
            h.out_i.reserve(part);
            h.out_d.reserve(part);
            CUDA_CHECK(cudaMemcpyAsync(h.q.p, q, nq * st * sizeof(double), cudaMemcpyHostToDevice, hs));
            h.index->knn_device(h.q.p, nq, k, h.out_i.p, h.out_d.p, hs);
            CUDA_CHECK(cudaStreamSynchronize(hs));
        });
        const auto t1 = std::chrono::steady_clock::now();
        on_all([&](int i) {
            Shard &h = sh_[i];
            const size_t lo = nq * (size_t)i / nd, hi = nq * ((size_t)i + 1) / nd, nb = hi - lo;
            if (nb == 0)
                return;
            DeviceGuard g(h.device);
            cudaStream_t hs = h.index->stream();
            h.blk_i.reserve(nd * nb * k);
            h.blk_d.reserve(nd * nb * k);
            h.mrg_i.reserve(nb * k);
            h.mrg_d.reserve(nb * k);
            for (size_t j = 0; j < nd; ++j)
            {
                CUDA_CHECK(cudaMemcpyPeerAsync(h.blk_i.p + j * nb * k, h.device, sh_[j].out_i.p + lo * k, sh_[j].device,
                                               nb * k * sizeof(uint64_t), hs));
                CUDA_CHECK(cudaMemcpyPeerAsync(h.blk_d.p + j * nb * k, h.device, sh_[j].out_d.p + lo * k, sh_[j].device,
                                               nb * k * sizeof(double), hs));
            }
            merge_topk_device(h.device, h.blk_d.p, h.blk_i.p, (int)nd, nb * k, nb, k, h.mrg_d.p, h.mrg_i.p, hs);
            CUDA_CHECK(cudaMemcpyAsync(ids + lo * k, h.mrg_i.p, nb * k * sizeof(uint64_t), cudaMemcpyDeviceToHost, hs));
            if (dists)
                CUDA_CHECK(cudaMemcpyAsync(dists + lo * k, h.mrg_d.p, nb * k * sizeof(double), cudaMemcpyDeviceToHost, hs));
            CUDA_CHECK(cudaStreamSynchronize(hs));
        });
        last_scan_phase_ns_ = (long)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
        last_merge_ns_ = (long)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t1).count();
    }

    void ShardedNN::knn_host(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists, uint32_t *counts)
    {
        if (nq == 0)
            return;
        REQUIRE(q && ids, "omplcu_nn_knn: NULL buffer");
        if (sh_.size() >= 2 && k > 0 && nq >= 4096 && 4 * sh_.size() * k * 16 <= 200 * 1024)
        {
            std::lock_guard<std::mutex> hl(host_mu_);
            const size_t kk = std::min(k, size());
            knn_host_distributed(q, nq, k, ids, dists);
            if (counts)
                for (size_t i = 0; i < nq; ++i)
                    counts[i] = (uint32_t)kk;
            return;
        }
        const size_t st = (size_t)sp_.stride;
        const int dev0 = sh_[0].device;
        cudaStream_t s0 = sh_[0].index->stream();
        std::lock_guard<std::mutex> hl(host_mu_);  // q0_ / mrg_* are per-handle staging
        size_t kk;
        {
            DeviceGuard g(dev0);
            q0_.reserve(nq * st);
            mrg_i_.reserve(nq * std::max<size_t>(k, 1));
            mrg_d_.reserve(nq * std::max<size_t>(k, 1));
            kk = std::min(k, size());
            CUDA_CHECK(cudaMemcpyAsync(q0_.p, q, nq * st * sizeof(double), cudaMemcpyHostToDevice, s0));
        }
        if (k > 0)
        {
            // mrg_d_ doubles as the caller-side distance buffer: pass it so knn_device does not reuse it as scratch
            knn_device(q0_.p, nq, k, mrg_i_.p, mrg_d_.p, s0);
            DeviceGuard g(dev0);
            CUDA_CHECK(cudaMemcpyAsync(ids, mrg_i_.p, nq * k * sizeof(uint64_t), cudaMemcpyDeviceToHost, s0));
            if (dists)
                CUDA_CHECK(cudaMemcpyAsync(dists, mrg_d_.p, nq * k * sizeof(double), cudaMemcpyDeviceToHost, s0));
            CUDA_CHECK(cudaStreamSynchronize(s0));
        }
        if (counts)
            for (size_t i = 0; i < nq; ++i)
                counts[i] = (uint32_t)kk;
    }

    void ShardedNN::radius_host(const double *q, size_t nq, const double *radii, uint64_t **ids, double **dists,
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
        std::lock_guard<std::mutex> lk(mu_);
        const size_t nd = sh_.size();
        std::vector<uint64_t *> pi(nd, nullptr);
        std::vector<double *> pd(nd, nullptr);
        std::vector<std::vector<uint64_t>> po(nd, std::vector<uint64_t>(nq + 1, 0));
        struct Free
        {
            std::vector<uint64_t *> &a;
            std::vector<double *> &b;
            ~Free()
            {
                for (auto p : a)
                    free(p);
                for (auto p : b)
                    free(p);
            }
        } guard{pi, pd};
        on_all([&](int i) { sh_[i].index->radius_host(q, nq, radii, &pi[i], &pd[i], po[i].data()); });
        uint64_t total = 0;
        for (size_t i = 0; i < nq; ++i)
        {
            for (size_t p = 0; p < nd; ++p)
                total += po[p][i + 1] - po[p][i];
            offsets[i + 1] = total;
        }
        uint64_t *oi = (uint64_t *)malloc(std::max<uint64_t>(total, 1) * sizeof(uint64_t));
        double *od = (double *)malloc(std::max<uint64_t>(total, 1) * sizeof(double));
        if (!oi || !od)
        {
            free(oi);
            free(od);
            throw Error(OMPLCU_ERR_INTERNAL, "omplcu_nn_radius: out of host memory");
        }
        std::vector<uint64_t> pos(nd);
        for (size_t i = 0; i < nq; ++i)
        {
            for (size_t p = 0; p < nd; ++p)
                pos[p] = po[p][i];
            for (uint64_t w = offsets[i]; w < offsets[i + 1]; ++w)
            {
                int best = -1;
                for (size_t p = 0; p < nd; ++p)
                {
                    if (pos[p] >= po[p][i + 1])
                        continue;
                    if (best < 0 || pd[p][pos[p]] < pd[best][pos[best]] ||
                        (pd[p][pos[p]] == pd[best][pos[best]] && pi[p][pos[p]] < pi[best][pos[best]]))
                        best = (int)p;
                }
                oi[w] = pi[best][pos[best]];
                od[w] = pd[best][pos[best]];
                pos[best]++;
            }
        }
        *ids = oi;
        if (dists)
            *dists = od;
        else
            free(od);
    }

    NNApi *make_sharded_nn(const SpaceDesc &sp, const int *devices, int ndev)
    {
        return new ShardedNN(sp, devices, ndev);
    }
}  // namespace omplcu
