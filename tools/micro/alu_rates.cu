// Issue rates of the integer / select / compare instructions the list-build loop is made of (sm_100a):
// 8 independent chains per thread, 16 warps per scheduler.  Prints warp-instructions per clock per scheduler.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/alu_rates tools/micro/alu_rates.cu && /tmp/alu_rates
#include <cstdio>
#include <cuda_runtime.h>
#define CH 8
#define BODY(NAME, STMT)                                                        \
    __global__ void NAME(float *out, int iters, float a, float b, int ia)       \
    {                                                                           \
        float x[CH]; int z[CH];                                                 \
        _Pragma("unroll") for (int c = 0; c < CH; c++) { x[c] = threadIdx.x * 1e-3f + c; z[c] = threadIdx.x + c; } \
        for (int i = 0; i < iters; i++) {                                       \
            _Pragma("unroll") for (int c = 0; c < CH; c++) { STMT }             \
        }                                                                       \
        float s = 0;                                                            \
        _Pragma("unroll") for (int c = 0; c < CH; c++) s += x[c] + z[c];        \
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;                         \
    }
BODY(k_ffma,  asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[c]) : "f"(a), "f"(b));)
BODY(k_fadd,  asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(a));)
BODY(k_fmnmx, asm volatile("max.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(a));)
BODY(k_imnmx, asm volatile("min.s32 %0, %0, %1;" : "+r"(z[c]) : "r"(ia));)
BODY(k_iadd,  asm volatile("add.s32 %0, %0, %1;" : "+r"(z[c]) : "r"(ia));)
BODY(k_lop3,  asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[c]) : "r"(ia), "r"(i));)
BODY(k_imad,  asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(z[c]) : "r"(ia), "r"(i));)
BODY(k_fsel,  asm volatile("{.reg .pred p; setp.lt.f32 p, %0, %1; selp.f32 %0, %2, %0, p;}" : "+f"(x[c]) : "f"(a), "f"(b));)
BODY(k_isetp_sel, asm volatile("{.reg .pred p; setp.ne.s32 p, %0, %1; selp.s32 %0, %2, %0, p;}" : "+r"(z[c]) : "r"(ia), "r"(i));)
BODY(k_padd,  asm volatile("{.reg .pred p; setp.lt.f32 p, %1, %2; @p add.s32 %0, %0, 1;}" : "+r"(z[c]) : "f"(x[c]), "f"(a));)
BODY(k_mix_fma_alu, asm volatile("fma.rn.f32 %0, %0, %2, %3; min.s32 %1, %1, %4;" : "+f"(x[c]), "+r"(z[c]) : "f"(a), "f"(b), "r"(ia));)
BODY(k_mix_2fma_alu, asm volatile("fma.rn.f32 %0, %0, %2, %3; fma.rn.f32 %0, %0, %2, %3; min.s32 %1, %1, %4;" : "+f"(x[c]), "+r"(z[c]) : "f"(a), "f"(b), "r"(ia));)

// does a packed FFMA2 take one issue slot?  (x2: two fp32 FMAs per instruction)
BODY(k_2ffma_2imnmx, asm volatile("fma.rn.f32 %0, %0, %2, %3; fma.rn.f32 %0, %0, %3, %2; min.s32 %1, %1, %4; min.s32 %1, %1, %5;" : "+f"(x[c]), "+r"(z[c]) : "f"(a), "f"(b), "r"(ia), "r"(i));)
BODY(k_ffma2_2imnmx, asm volatile("{.reg .b64 p, q, r; mov.b64 p, {%0, %0}; mov.b64 q, {%2, %2}; mov.b64 r, {%3, %3}; fma.rn.f32x2 p, p, q, r; mov.b64 {%0, _}, p;} min.s32 %1, %1, %4; min.s32 %1, %1, %5;" : "+f"(x[c]), "+r"(z[c]) : "f"(a), "f"(b), "r"(ia), "r"(i));)
BODY(k_ffma2_only, asm volatile("{.reg .b64 p, q, r; mov.b64 p, {%0, %0}; mov.b64 q, {%1, %1}; mov.b64 r, {%2, %2}; fma.rn.f32x2 p, p, q, r; mov.b64 {%0, _}, p;}" : "+f"(x[c]) : "f"(a), "f"(b));)
BODY(k_2imnmx, asm volatile("min.s32 %0, %0, %1; min.s32 %0, %0, %2;" : "+r"(z[c]) : "r"(ia), "r"(i));)
typedef void (*kern_t)(float *, int, float, float, int);
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 4, threads = 512, iters = 20000;
    float *out; cudaMalloc(&out, sizeof(float) * blocks * threads);
    struct { const char *name; kern_t k; int per; } tests[] = {
        {"FFMA", k_ffma, 1}, {"FADD", k_fadd, 1}, {"FMNMX", k_fmnmx, 1}, {"IMNMX", k_imnmx, 1}, {"IADD", k_iadd, 1},
        {"LOP3", k_lop3, 1}, {"IMAD", k_imad, 1}, {"FSETP+SEL", k_fsel, 2}, {"ISETP+SEL", k_isetp_sel, 2},
        {"FSETP+@IADD", k_padd, 2}, {"FFMA+IMNMX", k_mix_fma_alu, 2}, {"2FFMA+IMNMX", k_mix_2fma_alu, 3},
        {"2FFMA+2IMNMX", k_2ffma_2imnmx, 4}, {"FFMA2+2IMNMX", k_ffma2_2imnmx, 3}, {"FFMA2", k_ffma2_only, 1}, {"2IMNMX", k_2imnmx, 2}};
    for (auto &t : tests) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        t.k<<<blocks, threads>>>(out, iters, 0.999f, 1e-3f, 12345);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        t.k<<<blocks, threads>>>(out, iters, 0.999f, 1e-3f, 12345);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double winst = (double)blocks * (threads / 32) * iters * CH * t.per;     // warp-instructions (PTX count)
        const double clk = ms * 1e-3 * p.clockRate * 1e3;
        printf("%-12s %.3f ms  %.2f warp-inst/clk/scheduler\n", t.name, ms, winst / clk / (sms * 4));
    }
    return 0;
}
