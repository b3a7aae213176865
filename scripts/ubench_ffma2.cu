// ubench_ffma2.cu -- which operand pattern lets the scan's inner loop run the
// FP32 FMA pipe at full rate on sm_100a?  Standalone micro-benchmark (not part
// of the library).  Build + run on the GPU box:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/ub scripts/ubench_ffma2.cu && /tmp/ub
// Each variant computes, for RQ=4 queries per thread and a shared-memory tile
// of 256 points x 8 floats, s = pn + sum_c q_c * p_c and keeps the minimum.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long *>(&a);
    rb = *reinterpret_cast<unsigned long long *>(&b);
    rc = *reinterpret_cast<unsigned long long *>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}

constexpr int PT = 256, D = 7, ROWF = 8;

// V1: packed queries (2 per register pair), point coordinate broadcast (the scan's form)
template <int GP>
__global__ void __launch_bounds__(128) v1(const float *g, float *out, int iters)
{
    __shared__ __align__(16) float tile[PT * ROWF];
    for (int i = threadIdx.x; i < PT * ROWF; i += 128)
        tile[i] = g[i];
    __syncthreads();
    float2 qq[2][D];
    for (int c = 0; c < D; ++c)
    {
        qq[0][c] = make_float2(g[c] + threadIdx.x, g[c + 8] - threadIdx.x);
        qq[1][c] = make_float2(g[c + 16] + threadIdx.x, g[c + 24] - threadIdx.x);
    }
    float m = 1e30f;
    const float4 *tp = reinterpret_cast<const float4 *>(tile);
    for (int it = 0; it < iters; ++it)
#pragma unroll 1
        for (int i = 0; i < PT; i += GP)
        {
            float pv[GP][ROWF];
#pragma unroll
            for (int p = 0; p < GP; ++p)
            {
                float4 a = tp[(i + p) * 2], b = tp[(i + p) * 2 + 1];
                pv[p][0] = a.x, pv[p][1] = a.y, pv[p][2] = a.z, pv[p][3] = a.w;
                pv[p][4] = b.x, pv[p][5] = b.y, pv[p][6] = b.z, pv[p][7] = b.w;
            }
            float2 acc[GP][2];
#pragma unroll
            for (int p = 0; p < GP; ++p)
                acc[p][0] = acc[p][1] = make_float2(pv[p][D], pv[p][D]);
#pragma unroll
            for (int c = 0; c < D; ++c)
#pragma unroll
                for (int p = 0; p < GP; ++p)
#pragma unroll
                    for (int jj = 0; jj < 2; ++jj)
                        acc[p][jj] = ffma2(qq[jj][c], make_float2(pv[p][c], pv[p][c]), acc[p][jj]);
#pragma unroll
            for (int p = 0; p < GP; ++p)
                m = fminf(m, fminf(fminf(acc[p][0].x, acc[p][0].y), fminf(acc[p][1].x, acc[p][1].y)));
        }
    if (m == 123.f)
        out[0] = m;
}

// V3: plain FFMA, 4 queries per thread
template <int GP>
__global__ void __launch_bounds__(128) v3(const float *g, float *out, int iters)
{
    __shared__ __align__(16) float tile[PT * ROWF];
    for (int i = threadIdx.x; i < PT * ROWF; i += 128)
        tile[i] = g[i];
    __syncthreads();
    float q[4][D];
    for (int c = 0; c < D; ++c)
        for (int j = 0; j < 4; ++j)
            q[j][c] = g[c + 8 * j] + threadIdx.x * (j + 1);
    float m = 1e30f;
    const float4 *tp = reinterpret_cast<const float4 *>(tile);
    for (int it = 0; it < iters; ++it)
#pragma unroll 1
        for (int i = 0; i < PT; i += GP)
        {
            float pv[GP][ROWF];
#pragma unroll
            for (int p = 0; p < GP; ++p)
            {
                float4 a = tp[(i + p) * 2], b = tp[(i + p) * 2 + 1];
                pv[p][0] = a.x, pv[p][1] = a.y, pv[p][2] = a.z, pv[p][3] = a.w;
                pv[p][4] = b.x, pv[p][5] = b.y, pv[p][6] = b.z, pv[p][7] = b.w;
            }
            float acc[GP][4];
#pragma unroll
            for (int p = 0; p < GP; ++p)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    acc[p][j] = pv[p][D];
#pragma unroll
            for (int c = 0; c < D; ++c)
#pragma unroll
                for (int p = 0; p < GP; ++p)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        acc[p][j] = fmaf(q[j][c], pv[p][c], acc[p][j]);
#pragma unroll
            for (int p = 0; p < GP; ++p)
                m = fminf(m, fminf(fminf(acc[p][0], acc[p][1]), fminf(acc[p][2], acc[p][3])));
        }
    if (m == 123.f)
        out[0] = m;
}

// V4: points packed in pairs (tile layout [pair][c][2]), query coordinate is the broadcast scalar
template <int GP2>  // point pairs per group
__global__ void __launch_bounds__(128) v4(const float *g, float *out, int iters)
{
    __shared__ __align__(16) float tile[PT * ROWF];
    for (int i = threadIdx.x; i < PT * ROWF; i += 128)
        tile[i] = g[i];
    __syncthreads();
    float q[4][D];
    for (int c = 0; c < D; ++c)
        for (int j = 0; j < 4; ++j)
            q[j][c] = g[c + 8 * j] + threadIdx.x * (j + 1);
    float m = 1e30f;
    const float4 *tp = reinterpret_cast<const float4 *>(tile);
    for (int it = 0; it < iters; ++it)
#pragma unroll 1
        for (int i = 0; i < PT / 2; i += GP2)
        {
            float2 pv[GP2][ROWF];  // pv[pair][c] = (x_c of point 2i, x_c of point 2i+1)
#pragma unroll
            for (int p = 0; p < GP2; ++p)
#pragma unroll
                for (int w = 0; w < 4; ++w)
                {
                    float4 a = tp[(i + p) * 4 + w];
                    pv[p][2 * w] = make_float2(a.x, a.y);
                    pv[p][2 * w + 1] = make_float2(a.z, a.w);
                }
            float2 acc[GP2][4];
#pragma unroll
            for (int p = 0; p < GP2; ++p)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    acc[p][j] = pv[p][D];
#pragma unroll
            for (int c = 0; c < D; ++c)
#pragma unroll
                for (int p = 0; p < GP2; ++p)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        acc[p][j] = ffma2(pv[p][c], make_float2(q[j][c], q[j][c]), acc[p][j]);
#pragma unroll
            for (int p = 0; p < GP2; ++p)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    m = fminf(m, fminf(acc[p][j].x, acc[p][j].y));
        }
    if (m == 123.f)
        out[0] = m;
}

template <typename K>
void run(const char *name, K kern, const float *g, float *out, int sms, int ctas_per_sm)
{
    const int iters = 400;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        kern<<<sms * ctas_per_sm, 128>>>(g, out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep)
            best = ms < best ? ms : best;
    }
    const double fma = (double)sms * ctas_per_sm * 128 * 4.0 * D * PT * iters;  // lane FMAs
    printf("%-28s ctas/sm %d  %.3f ms  %.1f TFLOP/s (FMA=2)  err=%s\n", name, ctas_per_sm, best, 2 * fma / best / 1e9,
           cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    float *g, *out;
    cudaMalloc(&g, PT * ROWF * 4);
    cudaMalloc(&out, 16);
    float h[PT * ROWF];
    for (int i = 0; i < PT * ROWF; ++i)
        h[i] = (float)((i * 2654435761u) >> 8) / 16777216.f;
    cudaMemcpy(g, h, sizeof(h), cudaMemcpyHostToDevice);
    const int sms = prop.multiProcessorCount;
    for (int c : {3, 4, 8})
    {
        run("v1 q-packed GP=1", v1<1>, g, out, sms, c);
        run("v1 q-packed GP=2", v1<2>, g, out, sms, c);
        run("v1 q-packed GP=4", v1<4>, g, out, sms, c);
        run("v3 plain FFMA GP=2", v3<2>, g, out, sms, c);
        run("v3 plain FFMA GP=4", v3<4>, g, out, sms, c);
        run("v4 p-packed GP2=1", v4<1>, g, out, sms, c);
        run("v4 p-packed GP2=2", v4<2>, g, out, sms, c);
        run("v4 p-packed GP2=4", v4<4>, g, out, sms, c);
    }
    return 0;
}
