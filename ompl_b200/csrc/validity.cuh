// validity.cuh -- host-side classes for batched validity checking.
#pragma once
#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace omplcu
{
    enum
    {
        OMPLCU_CHK_ALLVALID = 0,
        OMPLCU_CHK_HYPERCUBE = 1,
        OMPLCU_CHK_CHAIN = 2,
        OMPLCU_CHK_SE3 = 3,
        OMPLCU_CHK_CIRCLES = 4
    };
    constexpr int MAX_CHAIN_LINKS = 128;

    // passed to kernels by value; pointers are device pointers owned by Checker
    struct CheckerDesc
    {
        int kind;
        int dim;
        double edge_width;
        int nlinks;
        int nenv;
        double link_length;
        const double *env;
        int nrobot, nsph, nbox, ncirc;
        const double *robot, *sph, *box, *circ;
        int stage_smem;
        int nfilt_pad;            // SE3 filter rows (multiple of 4)
        const float4 *filt;       // (o.x, o.y, o.z, |o|^2 - W) per obstacle, see se3_valid_fast
        float filt_amax;          // bound on |o_c| and sqrt(W) over the rows
        int _pad;
        // broad-phase grid over the bounding spheres (validity.cu, build_se3_grid)
        float grid_x0, grid_y0, grid_z0, grid_inv_h;
        int grid_nx, grid_ny, grid_nz, _pad2;
        const uint32_t *cell_start;  // [ncell + 1]
        const uint16_t *cell_items;  // obstacle rows per cell
        const float4 *robot32, *sph32, *box32;  // FP32 copies for the narrow phase (box: 2 float4 each)
    };

    class Checker
    {
    public:
        Checker(int device, const CheckerDesc &proto, const double *env, const double *robot, const double *sph,
                const double *box, const double *circ);
        ~Checker();

        void check_states_device(int stride, const double *d_states, size_t n, uint8_t *d_valid, cudaStream_t s);
        void check_motions_device(const SpaceDesc &sp, const double *d_s1, const double *d_s2, size_t n,
                                  uint8_t *d_valid, double *d_last_t, double *d_last_state, cudaStream_t s);
        // allocation-free variant: scratch in the library's static __device__ array (validity.cu)
        void check_motions_device_noalloc(const SpaceDesc &sp, const double *d_s1, const double *d_s2, size_t n,
                                          uint8_t *d_valid, double *d_last_t, double *d_last_state, cudaStream_t s);
        void check_states_host(int stride, const double *states, size_t n, uint8_t *valid);
        void check_motions_host(const SpaceDesc &sp, const double *s1, const double *s2, size_t n, uint8_t *valid,
                                double *last_t, double *last_state);
        long get_stat(const char *key) const;
        int device() const
        {
            return device_;
        }

    private:
        void build_se3_filter(const double *robot, const double *sph, const double *box);
        void build_se3_grid(const std::vector<double> &centre, const std::vector<double> &reach, int nf);
        int device_;
        int sms_ = 148;
        CheckerDesc desc_;
        cudaStream_t stream_ = nullptr, copy_stream_ = nullptr;
        std::mutex mu_, host_mu_;  // device-side scratch / host-pointer staging (host_mu_ is taken first)
        DevBuf<double> params_;
        DevBuf<float4> filt_, narrow32_;
        DevBuf<uint32_t> cell_start_;
        DevBuf<uint16_t> cell_items_;
        long grid_cells_ = 0, grid_items_ = 0;
        DevBuf<unsigned long long> ctr_;
        size_t smem_bytes_ = 0;
        // scratch
        DevBuf<uint32_t> nd_;
        DevBuf<double2> prep_;  // per-segment SO(3) interpolation constants (segment_count_kernel)
        DevBuf<unsigned long long> cnt_, offs_;
        DevBuf<unsigned> code_;
        DevBuf<unsigned char> cub_tmp_;
        DevBuf<double> st_a_, st_b_, st_t_, st_ls_;
        DevBuf<uint8_t> st_v_;
        long launches_ = 0;
        long last_states_ = 0;
        long last_evaluated_ = 0;
        long fb_states_ = 0, fb_evaluated_ = 0;  // the previous motion batch (feedback for the lane-group width)
        long last_group_ = 0;
        unsigned char *noalloc_scratch_ = nullptr;  // device address of the static scratch array on device_
    };

    // batched space math (host buffers) used by the parity tests and adapters
    void space_distance_host(const SpaceDesc &sp, int device, const double *a, const double *b, size_t n, double *out);
    void space_interpolate_host(const SpaceDesc &sp, int device, const double *a, const double *b, const double *t,
                                size_t n, double *out);
    void space_satisfies_bounds_host(const SpaceDesc &sp, int device, const double *low, const double *high,
                                     const double *states, size_t n, uint8_t *out);
    void space_segment_count_host(const SpaceDesc &sp, int device, const double *a, const double *b, size_t n,
                                  uint32_t *out);
}  // namespace omplcu
