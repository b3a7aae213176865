// peak.cu -- FP32 FMA pipe micro-benchmark: the roofline denominator for the
// kNN scan (MEASURED_PEAKS.json carries HBM and bf16 tensor peaks only).
// Each thread runs 16 independent FMA chains from registers; no memory traffic.
#include "common.cuh"
#include "knn_scan.cuh"

namespace omplcu
{
    namespace
    {
        template <bool PACKED>
        __global__ void __launch_bounds__(256) fma_peak_kernel(float *out, int iters, float m, float c)
        {
            float2 a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                a[j] = make_float2(threadIdx.x * 1e-3f + j, threadIdx.x * 2e-3f - j);
            const float2 mm = make_float2(m, m), cc = make_float2(c, c);
            for (int it = 0; it < iters; ++it)
            {
#pragma unroll
                for (int r = 0; r < 8; ++r)
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                    {
                        if (PACKED)
                            a[j] = ffma2(a[j], mm, cc);
                        else
                        {
                            a[j].x = fmaf(a[j].x, m, c);
                            a[j].y = fmaf(a[j].y, m, c);
                        }
                    }
            }
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                s += a[j].x + a[j].y;
            if (s == 123.456f)
                out[0] = s;
        }

        template <bool PACKED>
        double run_peak(int sms, cudaStream_t s, float *d_out)
        {
            const int iters = 4096, blocks = sms * 8, threads = 256;
            cudaEvent_t e0, e1;
            CUDA_CHECK(cudaEventCreate(&e0));
            CUDA_CHECK(cudaEventCreate(&e1));
            double best = 0.0;
            for (int rep = 0; rep < 5; ++rep)
            {
                CUDA_CHECK(cudaEventRecord(e0, s));
                fma_peak_kernel<PACKED><<<blocks, threads, 0, s>>>(d_out, iters, 0.999f, 0.001f);
                CUDA_CHECK(cudaEventRecord(e1, s));
                CUDA_CHECK(cudaEventSynchronize(e1));
                float ms = 0.f;
                CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
                // 128 FMAs (= 256 flop) per thread per iteration
                const double flop = (double)blocks * threads * (double)iters * 128.0 * 2.0;
                if (rep > 0)
                    best = std::max(best, flop / (ms * 1e-3) / 1e12);
            }
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            return best;
        }
    }  // namespace

    void measure_fp32_peak(int device, double *tflops_ffma, double *tflops_ffma2)
    {
        DeviceGuard g(device);
        cudaDeviceProp prop;
        CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        cudaStream_t s;
        CUDA_CHECK(cudaStreamCreate(&s));
        float *d_out = nullptr;
        CUDA_CHECK(cudaMalloc(&d_out, 16));
        double a = run_peak<false>(prop.multiProcessorCount, s, d_out);
        double b = run_peak<true>(prop.multiProcessorCount, s, d_out);
        cudaFree(d_out);
        cudaStreamDestroy(s);
        if (tflops_ffma)
            *tflops_ffma = a;
        if (tflops_ffma2)
            *tflops_ffma2 = b;
    }
}  // namespace omplcu
