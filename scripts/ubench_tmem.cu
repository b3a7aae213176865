// Micro-benchmark: TMEM read throughput of tcgen05.ld (32x32b, x8..x128) with 1..4 warps reading
// their own lane quadrants, one CTA per SM.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int X> struct Ld;
#define REGS8(b) "%" #b "0"
template <> struct Ld<32>
{
    __device__ static void go(uint32_t a, uint32_t (&r)[128])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(a) : "memory");
    }
};
template <> struct Ld<8>
{
    __device__ static void go(uint32_t a, uint32_t (&r)[128])
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "r"(a) : "memory");
    }
};

template <int X, int INFLIGHT>
__global__ void k(int iters, int nwarps, long long *out, uint32_t *sink)
{
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = slot + ((uint32_t)(warp * 32) << 16);
    uint32_t r[128];
    uint32_t acc = 0;
    long long t0 = clock64();
    if (warp < nwarps)
    {
        for (int i = 0; i < iters; ++i)
        {
#pragma unroll
            for (int j = 0; j < INFLIGHT; ++j)
                Ld<X>::go(base + ((i * INFLIGHT + j) * X) % 512, r);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            acc += r[0];
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0)
        out[0] = t1 - t0;
    if (acc == 0x12345)
        sink[0] = acc;
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
}

template <int X, int INF>
void run(int nwarps)
{
    long long *d; uint32_t *s;
    cudaMalloc(&d, 8); cudaMalloc(&s, 4);
    const int iters = 2000;
    k<X, INF><<<148, 128>>>(iters, nwarps, d, s);
    k<X, INF><<<148, 128>>>(iters, nwarps, d, s);
    long long h = 0;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaDeviceSynchronize();
    const double per = (double)h / (iters * INF);
    printf("x%-3d inflight %d warps %d: %.1f clk per ld (%.1f B/clk per warp, %.1f B/clk per SM) %s\n", X, INF, nwarps, per,
           X * 128.0 / per, X * 128.0 * nwarps / per, cudaGetErrorString(e));
}

int main()
{
    run<32, 1>(1); run<32, 1>(4); run<32, 2>(1); run<32, 2>(4); run<32, 4>(1); run<32, 4>(4);
    run<8, 1>(1); run<8, 4>(4); run<8, 8>(4);
    return 0;
}
