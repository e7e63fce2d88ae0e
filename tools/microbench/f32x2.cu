// Issue-rate microbenchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2), alone and mixed with integer ALU work.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2 f32x2.cu && ./f32x2
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b) { unsigned d; asm volatile("xor.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, float c)
{
    float s[16]; u64 p[8]; unsigned q[8];
    for (int i = 0; i < 16; ++i) s[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 8; ++i) { float2 t = make_float2(s[2 * i], s[2 * i + 1]); p[i] = *reinterpret_cast<u64*>(&t); q[i] = threadIdx.x + i; }
    float2 cc2 = make_float2(c, c); u64 C = *reinterpret_cast<u64*>(&cc2);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {                 // 16 independent scalar FFMA
#pragma unroll
            for (int i = 0; i < 16; ++i) s[i] = fma1(s[i], c, s[(i + 1) & 15]);
        } else if (MODE == 1) {          // 8 packed FFMA2 (= 16 FMAs)
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], C, p[(i + 1) & 7]);
        } else if (MODE == 2) {          // 16 FFMA + 16 XOR
#pragma unroll
            for (int i = 0; i < 16; ++i) { s[i] = fma1(s[i], c, s[(i + 1) & 15]); q[i & 7] = lop(q[i & 7], q[(i + 3) & 7]); }
        } else if (MODE == 3) {          // 8 FFMA2 + 16 XOR
#pragma unroll
            for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], C, p[(i + 1) & 7]); q[i] = lop(q[i], q[(i + 3) & 7]); q[(i + 1) & 7] = lop(q[(i + 1) & 7], q[(i + 5) & 7]); }
        }
    }
    long long t1 = clock64();
    float acc = 0; for (int i = 0; i < 16; ++i) acc += s[i];
    for (int i = 0; i < 8; ++i) { float2 t = *reinterpret_cast<float2*>(&p[i]); acc += t.x + t.y + q[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0);
}
template <int MODE> void run(const char* name, float* d, int fmas, int others)
{
    const int iters = 20000;
    k<MODE><<<148, 512>>>(d, iters, 1.0001f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<148, 512>>>(d, iters, 1.0001f); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); float cyc; cudaMemcpy(&cyc, d, 4, cudaMemcpyDeviceToHost);
    // per SMSP: 4 warps; cycles per iteration per SMSP
    double cpi = cyc / iters;
    printf("%-22s %8.3f ms  %7.2f cycles/iter/SMSP(4 warps)  -> %.2f FMA-lanes/clk/SM, %.2f warp-inst/clk/SMSP\n", name, ms, cpi,
           4.0 * 4 * fmas * 32 / cpi, 4.0 * ((MODE & 1 ? fmas / 2 : fmas) + others) / cpi);
}
int main()
{
    float* d; cudaMalloc(&d, 148 * 512 * 4);
    run<0>("16 FFMA", d, 16, 0);
    run<1>("8 FFMA2", d, 16, 0);
    run<2>("16 FFMA + 16 LOP", d, 16, 16);
    run<3>("8 FFMA2 + 16 LOP", d, 16, 16);
    return 0;
}
