// ubench_rf.cu -- register-only FMA loops: how fast can FFMA2 / FFMA issue when
// operands vary (no shared memory, no compares)?  Finds the operand pattern for
// the kNN scan's inner loop.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b),
                       rc = *reinterpret_cast<unsigned long long *>(&c), rd;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}
// J accumulators (packed), K broadcast scalars; inner order k-major: B reused across the J consecutive FFMA2
template <int J, int K>
__global__ void __launch_bounds__(128) p1(float *out, int iters, float s)
{
    float2 acc[J], qa[J];
    float b[K];
    for (int j = 0; j < J; ++j) acc[j] = make_float2(threadIdx.x * s + j, j), qa[j] = make_float2(s * j + 1e-3f, s - j * 1e-3f);
    for (int k = 0; k < K; ++k) b[k] = s * (k + 1) + threadIdx.x * 1e-6f;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int j = 0; j < J; ++j)
                acc[j] = ffma2(qa[j], make_float2(b[k], b[k]), acc[j]);
    float r = 0;
    for (int j = 0; j < J; ++j) r += acc[j].x + acc[j].y;
    if (r == 123.f) out[0] = r;
}
// A (packed) reused across J consecutive FFMA2, B scalar varies
template <int J, int K>
__global__ void __launch_bounds__(128) p2(float *out, int iters, float s)
{
    float2 acc[J], P[K];
    float q[J];
    for (int j = 0; j < J; ++j) acc[j] = make_float2(threadIdx.x * s + j, j), q[j] = s * (j + 1) + threadIdx.x * 1e-6f;
    for (int k = 0; k < K; ++k) P[k] = make_float2(s * k + 1e-3f, s - k * 1e-3f);
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int j = 0; j < J; ++j)
                acc[j] = ffma2(P[k], make_float2(q[j], q[j]), acc[j]);
    float r = 0;
    for (int j = 0; j < J; ++j) r += acc[j].x + acc[j].y;
    if (r == 123.f) out[0] = r;
}
// both A and B true packed registers (no broadcast form), A reused
template <int J, int K>
__global__ void __launch_bounds__(128) p5(float *out, int iters, float s)
{
    float2 acc[J], P[K], q[J];
    for (int j = 0; j < J; ++j) acc[j] = make_float2(threadIdx.x * s + j, j), q[j] = make_float2(s * (j + 1), s * j + threadIdx.x * 1e-6f);
    for (int k = 0; k < K; ++k) P[k] = make_float2(s * k + 1e-3f, s - k * 1e-3f);
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int j = 0; j < J; ++j)
                acc[j] = ffma2(P[k], q[j], acc[j]);
    float r = 0;
    for (int j = 0; j < J; ++j) r += acc[j].x + acc[j].y;
    if (r == 123.f) out[0] = r;
}
// plain FFMA, b reused across J
template <int J, int K>
__global__ void __launch_bounds__(128) p4(float *out, int iters, float s)
{
    float acc[J], q[J], b[K];
    for (int j = 0; j < J; ++j) acc[j] = threadIdx.x * s + j, q[j] = s * j + 1e-3f;
    for (int k = 0; k < K; ++k) b[k] = s * (k + 1) + threadIdx.x * 1e-6f;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int j = 0; j < J; ++j)
                acc[j] = fmaf(q[j], b[k], acc[j]);
    float r = 0;
    for (int j = 0; j < J; ++j) r += acc[j];
    if (r == 123.f) out[0] = r;
}
template <typename Kn>
void run(const char *name, Kn kern, float *out, int sms, double fma_per_thread_iter)
{
    const int iters = 20000, ctas = sms * 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        kern<<<ctas, 128>>>(out, iters, 0.999f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = ms < best ? ms : best;
    }
    printf("%-40s %.3f ms  %.1f TFLOP/s  %s\n", name, best, 2.0 * ctas * 128 * fma_per_thread_iter * iters / best / 1e9,
           cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    float *out;
    cudaMalloc(&out, 16);
    int sms = prop.multiProcessorCount;
    run("p1 A var, B bcast reuse x2  (J2,K8)", p1<2, 8>, out, sms, 2 * 8 * 2);
    run("p1 A var, B bcast reuse x4  (J4,K8)", p1<4, 8>, out, sms, 4 * 8 * 2);
    run("p1 A var, B bcast reuse x8  (J8,K8)", p1<8, 8>, out, sms, 8 * 8 * 2);
    run("p1 A var, B bcast reuse x16 (J16,K4)", p1<16, 4>, out, sms, 16 * 4 * 2);
    run("p2 A reuse x4, B bcast var  (J4,K8)", p2<4, 8>, out, sms, 4 * 8 * 2);
    run("p2 A reuse x8, B bcast var  (J8,K8)", p2<8, 8>, out, sms, 8 * 8 * 2);
    run("p2 A reuse x16, B bcast var (J16,K4)", p2<16, 4>, out, sms, 16 * 4 * 2);
    run("p5 A reuse x8, B packed var (J8,K8)", p5<8, 8>, out, sms, 8 * 8 * 2);
    run("p5 A reuse x16, B packed var(J16,K4)", p5<16, 4>, out, sms, 16 * 4 * 2);
    run("p4 FFMA b reuse x8          (J8,K8)", p4<8, 8>, out, sms, 8 * 8);
    run("p4 FFMA b reuse x16         (J16,K4)", p4<16, 4>, out, sms, 16 * 4);
    return 0;
}
