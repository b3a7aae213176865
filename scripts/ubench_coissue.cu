// ubench_coissue.cu -- which instruction classes issue "for free" between FFMA2s?
// Base loop: 64 FFMA2 (A reused x8, B broadcast) per iteration = 128 FMA-pipe cycles;
// variants add 16 independent extra instructions of one class per iteration.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b),
                       rc = *reinterpret_cast<unsigned long long *>(&c), rd;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}
template <int MODE, int NX>
__global__ void __launch_bounds__(128) kern(float *out, int iters, float s, int si)
{
    __shared__ float sm[1024];
    for (int i = threadIdx.x; i < 1024; i += 128) sm[i] = i * s;
    __syncthreads();
    float2 acc[8], P[8];
    float q[8];
    for (int j = 0; j < 8; ++j) acc[j] = make_float2(threadIdx.x * s + j, j), q[j] = s * (j + 1) + threadIdx.x * 1e-6f;
    for (int k = 0; k < 8; ++k) P[k] = make_float2(s * k + 1e-3f, s - k * 1e-3f);
    int xi[16];
    float xf[16];
    for (int e = 0; e < 16; ++e) xi[e] = threadIdx.x + e * si, xf[e] = threadIdx.x * s + e;
    unsigned pred = 0;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
        {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                acc[j] = ffma2(P[k], make_float2(q[j], q[j]), acc[j]);
#pragma unroll
            for (int e = 2 * k; e < 2 * k + 2 && e < NX; ++e)
            {
                if (MODE == 1) asm volatile("add.s32 %0, %0, %1;" : "+r"(xi[e]) : "r"(si));
                if (MODE == 2) asm volatile("xor.b32 %0, %0, %1;" : "+r"(xi[e]) : "r"(si));
                if (MODE == 3) asm volatile("min.f32 %0, %0, %1;" : "+f"(xf[e]) : "f"(s));
                if (MODE == 4) { unsigned p; asm volatile("{.reg .pred t; setp.lt.f32 t, %1, %2; selp.u32 %0, 1, 0, t;}" : "=r"(p) : "f"(xf[e]), "f"(s)); pred |= p; }
                if (MODE == 5) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(&sm[(xi[e] & 1023)]))); xf[e] = v; }
                if (MODE == 6) asm volatile("min.s32 %0, %0, %1;" : "+r"(xi[e]) : "r"(si));
                if (MODE == 7) asm volatile("mul.f32 %0, %0, %1;" : "+f"(xf[e]) : "f"(s));
                if (MODE == 8) asm volatile("mov.b32 %0, %1;" : "=r"(xi[e]) : "r"(xi[(e + 1) & 15]));
                if (MODE == 9) { float a, b, c, d; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "r"((unsigned)__cvta_generic_to_shared(&sm[(e * 4) & 1023]))); xf[e] = a; xf[(e + 1) & 15] = b; xf[(e + 2) & 15] = c; xf[(e + 3) & 15] = d; }
                if (MODE == 10) { float a, b; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "r"((unsigned)__cvta_generic_to_shared(&sm[(e * 2) & 1023]))); xf[e] = a; xf[(e + 1) & 15] = b; }
                if (MODE == 11) asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(xf[e]) : "f"(s), "f"(xf[(e + 5) & 15]));
                if (MODE == 12) asm volatile("{.reg .pred t; setp.ne.u32 t, %0, 0; setp.lt.or.f32 t, %1, %2, t; selp.u32 %0, 1, 0, t;}" : "+r"(pred) : "f"(xf[e]), "f"(s));
                if (MODE == 13) asm volatile("{.reg .pred t; setp.lt.s32 t, %1, %2; @t add.s32 %0, %0, 1;}" : "+r"(pred) : "r"(xi[e]), "r"(si));
            }
        }
    }
    float r = pred;
    for (int j = 0; j < 8; ++j) r += acc[j].x + acc[j].y;
    for (int e = 0; e < 16; ++e) r += xi[e] + xf[e];
    if (r == 123.f) out[0] = r;
}
template <typename Kn>
void run(const char *name, Kn k, float *out, int sms)
{
    const int iters = 20000, ctas = sms * 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        k<<<ctas, 128>>>(out, iters, 0.999f, 3);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = ms < best ? ms : best;
    }
    printf("%-30s %.3f ms  %.1f TFLOP/s\n", name, best, 2.0 * ctas * 128 * 128.0 * iters / best / 1e9);
}
int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    float *out;
    cudaMalloc(&out, 16);
    int sms = prop.multiProcessorCount;
    run("base (64 FFMA2)", kern<0, 0>, out, sms);
    run("+16 IADD", kern<1, 16>, out, sms);
    run("+16 LOP(xor)", kern<2, 16>, out, sms);
    run("+16 FMNMX", kern<3, 16>, out, sms);
    run("+16 FSETP+SEL", kern<4, 16>, out, sms);
    run("+16 LDS.32", kern<5, 16>, out, sms);
    run("+16 IMNMX", kern<6, 16>, out, sms);
    run("+16 FMUL", kern<7, 16>, out, sms);
    run("+16 MOV", kern<8, 16>, out, sms);
    run("+16 LDS.128", kern<9, 16>, out, sms);
    run("+8 LDS.128", kern<9, 8>, out, sms);
    run("+16 LDS.64", kern<10, 16>, out, sms);
    run("+16 FMNMX3", kern<11, 16>, out, sms);
    run("+16 (ISETP.NE,FSETP.OR,SEL)", kern<12, 16>, out, sms);
    run("+16 (ISETP,@IADD)", kern<13, 16>, out, sms);
    run("+8 FMNMX", kern<3, 8>, out, sms);
    run("+8 LDS.32", kern<5, 8>, out, sms);
    return 0;
}
