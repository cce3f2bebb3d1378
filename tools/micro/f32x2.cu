// Throughput of packed fp32 (FFMA2) against scalar FFMA on sm_100a: independent chains per thread, all SMs busy.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/f32x2 tools/micro/f32x2.cu && /tmp/f32x2
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up(u64 v, float &a, float &b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

template <int CH>
__global__ void k_scalar(float *out, int iters, float a, float b)
{
    float x[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) x[c] = threadIdx.x * 1e-3f + c;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CH; c++) x[c] = fma1(x[c], a, b);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CH; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int CH>
__global__ void k_packed(float *out, int iters, float a, float b)
{
    u64 x[CH];
    const u64 A = pk(a, a), B = pk(b, b);
#pragma unroll
    for (int c = 0; c < CH; c++) x[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CH; c++) x[c] = fma2(x[c], A, B);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CH; c++) { float p, q; up(x[c], p, q); s += p + q; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// mixed: per iteration CH packed FMAs and CH integer adds (does the packed op leave issue slots for the ALU pipe?)
template <int CH>
__global__ void k_mixed(float *out, int iters, float a, float b, int packed)
{
    u64 x[CH];
    float y[2 * CH];
    int z[2 * CH];
    const u64 A = pk(a, a), B = pk(b, b);
#pragma unroll
    for (int c = 0; c < CH; c++) { x[c] = pk(threadIdx.x * 1e-3f + c, c); y[2 * c] = c; y[2 * c + 1] = -c; z[2 * c] = c; z[2 * c + 1] = threadIdx.x; }
    for (int i = 0; i < iters; i++) {
        if (packed) {
#pragma unroll
            for (int c = 0; c < CH; c++) x[c] = fma2(x[c], A, B);
        } else {
#pragma unroll
            for (int c = 0; c < 2 * CH; c++) y[c] = fma1(y[c], a, b);
        }
#pragma unroll
        for (int c = 0; c < 2 * CH; c++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[c]) : "r"(i), "r"(c));
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CH; c++) { float p, q; up(x[c], p, q); s += p + q + y[2 * c] + y[2 * c + 1] + z[2 * c] + z[2 * c + 1]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F> static float time_ms(F f)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 4, threads = 512, iters = 20000;
    float *out; cudaMalloc(&out, sizeof(float) * blocks * threads);
    constexpr int CH = 8;
    const double lanes = (double)blocks * threads * iters * CH;
    float ms = time_ms([&] { k_scalar<CH><<<blocks, threads>>>(out, iters, 0.999f, 1e-3f); });
    printf("scalar FFMA : %.3f ms  %.1f Gfma/s  (%.1f fma/clk/SM at %d MHz)\n", ms, lanes / ms * 1e-6, lanes / ms * 1e-3 / sms / (p.clockRate * 1e-3) , p.clockRate / 1000);
    ms = time_ms([&] { k_packed<CH><<<blocks, threads>>>(out, iters, 0.999f, 1e-3f); });
    printf("packed FFMA2: %.3f ms  %.1f Gfma/s  (%.1f fma/clk/SM)\n", ms, 2 * lanes / ms * 1e-6, 2 * lanes / ms * 1e-3 / sms / (p.clockRate * 1e-3));
    ms = time_ms([&] { k_mixed<CH><<<blocks, threads>>>(out, iters, 0.999f, 1e-3f, 0); });
    printf("mixed scalar (2CH FFMA + 2CH LOP3): %.3f ms\n", ms);
    ms = time_ms([&] { k_mixed<CH><<<blocks, threads>>>(out, iters, 0.999f, 1e-3f, 1); });
    printf("mixed packed (CH FFMA2 + 2CH LOP3): %.3f ms\n", ms);
    return 0;
}
