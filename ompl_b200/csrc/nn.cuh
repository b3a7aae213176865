// nn.cuh -- device-resident nearest-neighbour index (host-side class).
#pragma once
#include <atomic>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <vector>

#include "common.cuh"

namespace omplcu
{
    struct Space
    {
        SpaceDesc desc;
    };

    // Device mirror of NearestNeighbors<_T>'s element set
    // (datastructures/NearestNeighbors.h:46-115): coordinates only; the C++
    // adapter keeps id -> _T.  Slots are kept in insertion order, so slot order
    // == id order and "ties by insertion index" is "ties by slot".
    // What an omplcu_nn_t can do: one index on one GPU (NNIndex) or a roadmap sharded over several
    // (ShardedNN, sharded.cu).
    class NNApi
    {
    public:
        virtual ~NNApi() = default;
        virtual void set_id_base(uint64_t base) = 0;
        virtual uint64_t add(const double *rows, size_t n) = 0;
        virtual void remove(uint64_t id) = 0;
        virtual size_t remove_batch(const uint64_t *ids, size_t n) = 0;
        virtual void clear() = 0;
        virtual size_t size() const = 0;
        virtual size_t list(uint64_t *ids, size_t cap) const = 0;
        virtual void get(uint64_t id, double *row) = 0;
        virtual void knn_device(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                                cudaStream_t s) = 0;
        virtual void knn_host(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists,
                              uint32_t *counts) = 0;
        virtual void radius_host(const double *q, size_t nq, const double *radii, uint64_t **ids, double **dists,
                                 uint64_t *offsets) = 0;
        virtual void set_option(const char *key, long v) = 0;
        virtual long get_stat(const char *key) const = 0;
    };

    class NNIndex : public NNApi
    {
    public:
        NNIndex(const SpaceDesc &sp, int device);
        ~NNIndex() override;

        void set_id_base(uint64_t base) override;
        uint64_t add(const double *rows, size_t n) override
        {
            return add_at(rows, n, OMPLCU_ID_NONE);
        }
        // first_id == OMPLCU_ID_NONE: the next sequence number; otherwise the ids of this batch are
        // first_id .. first_id + n - 1, which must lie beyond every id handed out so far (a sharded
        // roadmap gives each shard increasing blocks of the global insertion sequence)
        uint64_t add_at(const double *rows, size_t n, uint64_t first_id);
        void remove(uint64_t id) override;
        size_t remove_batch(const uint64_t *ids, size_t n) override;
        void clear() override;
        size_t size() const override
        {
            return nlive_;
        }
        size_t list(uint64_t *ids, size_t cap) const override;
        void get(uint64_t id, double *row) override;

        // device pointers; synchronises the stream
        void knn_device(const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                        cudaStream_t s) override;
        void knn_host(const double *q, size_t nq, size_t k, uint64_t *ids, double *dists, uint32_t *counts) override;
        void radius_host(const double *q, size_t nq, const double *radii, uint64_t **ids, double **dists,
                         uint64_t *offsets) override;

        void set_option(const char *key, long v) override;
        long get_stat(const char *key) const override;

        int device() const
        {
            return device_;
        }
        cudaStream_t stream() const
        {
            return stream_;
        }
        const SpaceDesc &space() const
        {
            return sp_;
        }

    private:
        struct QCtx;
        void reserve_slots(size_t n);
        void compact();
        bool fast_eligible(size_t kk) const;
        void knn_locked(QCtx &c, const double *d_q, size_t nq, size_t k, uint64_t *d_ids, double *d_dists,
                        cudaStream_t s, bool sync = true);
        void knn_fast(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids, double *d_dists,
                      cudaStream_t s, bool no_tc = false);
        void knn_few(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids, double *d_dists,
                     cudaStream_t s);
        void knn_exact_few(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids,
                           double *d_dists, cudaStream_t s);
        // tensor-core threshold path: sample scan -> per-query bound -> append scan -> exact epilogue
        // Tensor-core filter in front of an exact path that has a per-query exact distance bound (the sample
        // bound of the large-k path, the radius of nearestR): every slot within d_bound[q] of query q ends up
        // in the query's rows of c.app_ids_ / c.app_cnt_ (smax rows of `cap` slots, in slot order).
        bool filter_usable() const;
        bool embed_ok(cudaStream_t s);
        unsigned embed_list_bound(QCtx &c, const double *d_qe, const void *pair_metric, size_t q0, size_t nq, size_t kk,
                                  double *d_bound, size_t kstride, uint64_t *d_ids, double *d_dists, cudaStream_t s);
        uint32_t filter_scan(QCtx &c, const double *d_q, size_t nq, const double *d_bound, uint32_t cap,
                             uint32_t *qpad_out, cudaStream_t s);
        void knn_fast_thr(QCtx &c, const double *d_q, size_t nq, size_t kk, size_t kstride, uint64_t *d_ids, double *d_dists,
                          cudaStream_t s);
        // exact sort path over the selected queries (d_sel == nullptr: all nq)
        bool prepare_chain_tables(QCtx &c, const double *d_q, size_t nq, cudaStream_t s);
        bool knn_exact_collect(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                               uint64_t *d_ids, double *d_dists, cudaStream_t s);
        void knn_exact_sort(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                            uint64_t *d_ids, double *d_dists, cudaStream_t s);
        void knn_exact(QCtx &c, const double *d_q, const uint32_t *d_sel, size_t nsel, size_t kk, size_t kstride,
                       uint64_t *d_ids, double *d_dists, cudaStream_t s);

        int device_;
        SpaceDesc sp_;
        cudaStream_t stream_ = nullptr;  // mutations (add / remove / compact / clear)
        // Concurrency (SURVEY 8(b) "Threading"; GNAT is what OMPL hands multithreaded planners because const
        // queries run concurrently, tools/config/SelfConfig.h:112-113): queries hold rw_ shared and run on
        // their own leased context (stream + scratch), mutations hold it exclusively.  prep_mu_ guards the
        // operand images that queries rebuild lazily after a mutation; stat_mu_ the published statistics.
        mutable std::shared_mutex rw_;
        std::mutex prep_mu_;
        mutable std::mutex stat_mu_, pool_mu_;

        size_t nslots_ = 0, nlive_ = 0;
        uint64_t id_base_ = 0, next_seq_ = 0;
        std::vector<uint64_t> h_ids_;
        std::vector<uint8_t> h_alive_;

        DevBuf<double> pts64_;
        DevBuf<uint64_t> ids_;
        DevBuf<uint8_t> alive_;

        // FP32 filter copy (plain RealVector spaces with dim <= 128)
        bool fast_space_ = false;
        int D_ = 0, rowf_ = 0;
        size_t filt_rows_ = 0;  // rows initialised (multiple of SCAN_PT)
        DevBuf<float> filt_;
        DevBuf<double> center_;
        DevBuf<unsigned> pmax2_;  // bits of max fl32(|pt|^2)
        bool have_center_ = false;
        // SO(2) / SO(3) / compound spaces: filt_ holds a Euclidean lower-bound embedding of the states (nn.cu, embed_state);
        // only the filter in front of the exact paths uses it
        bool embed_ = false, embed_okay_ = true;
        double embed_slack_ = 0.0;
        unsigned long embed_ver_ = 0;
        DevBuf<unsigned> embed_bad_;

        // shared, rebuilt lazily by the first query after a mutation (under prep_mu_, then synchronised):
        // BF16 split-operand images of the filter copy (knn_scan_tc.cuh), the strided sample image of the
        // threshold path, the chain metric's per-slot trig table
        DevBuf<uint16_t> tc_pimg_, tc_simg_;
        unsigned long mut_version_ = 1, tc_version_ = 0, tc_sversion_ = 0;
        size_t tc_sstep_ = 0;
        DevBuf<double> trig_;
        size_t trig_rows_ = 0;

        // Everything one query in flight needs: a stream, its scratch, its statistics.
        struct QCtx
        {
            cudaStream_t stream = nullptr;
            cudaEvent_t ev0_ = nullptr, ev1_ = nullptr, ev2_ = nullptr, ev3_ = nullptr;
            DevBuf<double> q_stage_, out_d_;
            DevBuf<uint64_t> out_i_;
            DevBuf<float> qp_, cand_vals_, sample_thr_buf_;
            DevBuf<double> qaux_;
            DevBuf<uint32_t> cand_ids_, ovf_list_;
            DevBuf<unsigned> counters_;
            DevBuf<double> ex_keys_, ex_keys2_;
            DevBuf<uint32_t> ex_vals_, ex_vals2_;
            DevBuf<int> ex_offs_;
            DevBuf<double> col_sd_, col_bound_, qtrig_, rad_r_, rad_d_, ex_keys3_;
            DevBuf<uint64_t> rad_ids_;
            DevBuf<unsigned long long> rad_off_;
            size_t exact_nq_ = 0;
            DevBuf<uint32_t> col_cnt_, col_redo_;
            DevBuf<unsigned char> cub_tmp_;
            DevBuf<uint16_t> tc_qimg_;
            DevBuf<float> tau_, list_vals_;
            DevBuf<uint32_t> app_ids_, app_cnt_;
            DevBuf<unsigned long long> kept_;
            // small host-pointer calls (knn_host): pinned staging, and the place knn_fast copies the results to
            // together with its fallback counter, so that the common call synchronises once
            PinBuf pin_;
            uint64_t *ho_ids_ = nullptr;
            double *ho_dists_ = nullptr;
            size_t ho_entries_ = 0;
            bool ho_done_ = false;
            DevBuf<double> xf_d_;       // exact few-query path: per-warp lists of (distance, slot)
            DevBuf<uint32_t> xf_s_;
            DevBuf<double> q_emb_, emb_bound_, q_emb2_, col_bound2_;
            DevBuf<uint32_t> unp_list_;  // queries the list scan on the embedding did not prove  // embedded queries and their filter bounds (embedded spaces)
            DevBuf<uint32_t> qs_keys_, qs_idx_, qs_bbox_;  // Morton order of a large query batch (knn_fast)
            DevBuf<uint32_t> re_sel_;   // second pass of knn_fast: the unproven queries, gathered
            DevBuf<double> re_q_, re_d_;
            DevBuf<uint64_t> re_ids_;
            long last_rescans_ = 0, last_few_ = 0, last_proven_ = 0;
            long last_filter_ = 0;
            long last_collect_redo_ = 0, last_tc_ = 0, last_thr_ = 0, last_kept_ = 0, last_rank_ = 0;
            long last_sample_ns_ = 0, last_main_ns_ = 0, last_scan_ns_ = 0, last_fallbacks_ = 0, last_path_ = 0;
        };
        // leases a context for the duration of a query (creates one if all are busy)
        struct Lease
        {
            NNIndex &ix;
            QCtx *c;
            explicit Lease(NNIndex &i);
            ~Lease();
        };
        std::vector<std::unique_ptr<QCtx>> pool_;
        std::vector<QCtx *> free_;
        void publish(const QCtx &c);  // copy the query's statistics to the handle (last_* stats)

        // options
        bool collect_ = true;
        int tc_mode_ = -1;        // -1 auto, 0 off, 1 whenever eligible
        int tc_na_ = 0;
        int tc_variant_ = 0;      // image layout / descriptor variants (diagnostics)
        float *tc_dbg_ = nullptr; // device buffer for raw scan values (diagnostics), caller-owned
        int thr_mode_ = 0;        // -1 auto (index of >= 32768 slots), 0 never, 1 whenever the tensor-core scan runs
        int thr_step_ = 16;       // sample stride
        int sm_count_ = 148, scan_ctas_per_sm_ = 0;
        bool sample_thr_ = true;
        bool l2_window_ = false;
        std::atomic<long> tc_seen_{0}, tc_unproven_{0};  // auto mode: recent queries / unproven by the tensor-core filter
        std::atomic<bool> tc_demoted_{false};
        std::atomic<size_t> tc_demote_slots_{0};
        bool sort_queries_ = true;  // large batches: scan the queries in Morton order
        bool few_mode_ = true;    // auto mode, up to 8 queries: knn_few_kernel
        long last_few_ = 0, last_proven_ = 0;
        bool tc_align_ = true;    // tensor-core scan: work pieces end on query-tile boundaries where that is nearly free
        int tc_grid_ = 0;         // experiments: upper limit on the tensor-core scan's work pieces (0: one per SM)
        int tc_pair_ = 1;         // with tc_ts_: clusters of 2 CTAs share the point chunks (TMA multicast)
        int tc_ts_ = 1;           // one resident tile: query tile in tensor memory (0: shared memory)
        bool force_exact_ = false, profile_ = false;

        // published statistics (stat_mu_)
        long last_filter_ = 0;
        long last_collect_redo_ = 0, last_tc_ = 0, last_thr_ = 0, last_kept_ = 0, last_rank_ = 0;
        long last_sample_ns_ = 0, last_main_ns_ = 0, last_scan_ns_ = 0, last_fallbacks_ = 0, last_path_ = 0;
        long last_rescans_ = 0;
        long scan_ns_total_ = 0, scan_count_ = 0;
        std::atomic<long> launches_{0};
    };

    // roadmap sharded over several GPUs of one process (sharded.cu)
    NNApi *make_sharded_nn(const SpaceDesc &sp, const int *devices, int ndev);

    void merge_topk_device(int device, const double *d_dists, const uint64_t *d_ids, int nparts, size_t part_stride,
                           size_t nq, size_t k,
                           double *d_out_dists, uint64_t *d_out_ids, cudaStream_t s);
}  // namespace omplcu
