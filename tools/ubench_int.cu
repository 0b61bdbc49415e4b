// Integer / FP64 pipe micro-benchmark for B200: thread-ops / clk / SM of the instructions a 64-bit modular butterfly is
// built from (results: profiles/r1_ubench_int.txt).  Every kernel runs ITERS iterations of 64 "bodies" over 8 independent
// dependency chains per thread (8 chains x 8 repeats, fully unrolled), 2 CTAs of 1024 threads per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_int tools/ubench_int.cu && tools/ubench_int
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
constexpr int ITERS = 1024;

#define WIDE(i) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %1, %0;}" : "+l"(x[i]) : "r"(b));
#define IADD2(i) asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(z[i])); asm volatile("add.u32 %0, %0, %1;" : "+r"(z[i]) : "r"(y[i]));
#define LO(v, i) asm volatile("mad.lo.u32 %0, %1, %0, %2;" : "+r"(v[i]) : "r"(a), "r"(b));

// XT: type of the x chain; BODY(i): the instruction(s) of one body on chain i
#define UBENCH(NAME, XT, BODY)                                                                                           \
    __global__ void __launch_bounds__(1024) NAME(u64* out, u32 a0, u32 b0) {                                             \
        u32 a = a0 + threadIdx.x, b = b0 | 1; u64 a64 = ((u64)a << 29) | b;                                              \
        double ad = 1.0 + 1e-9 * a; float af = 1.0f + 1e-6f * a; (void)a64; (void)ad; (void)af; (void)b;                 \
        XT x[8]; u32 y[8], z[8]; double d[8];                                                                            \
        _Pragma("unroll") for (int i = 0; i < 8; ++i) { x[i] = (XT)(b0 + i + threadIdx.x); y[i] = i + a0; z[i] = i + b0; d[i] = 1.0 + i; } \
        _Pragma("unroll 1") for (int it = 0; it < ITERS; ++it) {                                                         \
            _Pragma("unroll") for (int r = 0; r < 8; ++r) {                                                              \
                _Pragma("unroll") for (int i = 0; i < 8; ++i) { BODY(i) }                                                \
            }                                                                                                            \
        }                                                                                                                \
        u64 s = 0;                                                                                                       \
        _Pragma("unroll") for (int i = 0; i < 8; ++i) s += (u64)x[i] + y[i] + z[i] + (u64)d[i];                          \
        if (s == 0x123456789ull) out[threadIdx.x] = s;                                                                   \
    }

#define B_WIDE(i) WIDE(i)
#define B_LO(i) LO(x, i)
#define B_HI(i) asm volatile("mad.hi.u32 %0, %1, %0, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
#define B_IADD3(i) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(y[i])); asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
#define B_ADD64(i) asm volatile("add.u64 %0, %0, %1;" : "+l"(x[i]) : "l"(a64));
#define B_LOP3(i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
#define B_SETP(i) asm volatile("{.reg .pred p; setp.ge.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(x[i]) : "r"(a), "r"(b));
#define B_PSUB(i) asm volatile("{.reg .pred p; setp.ge.u64 p, %0, %1; @p sub.u64 %0, %0, %1;}" : "+l"(x[i]) : "l"(a64));
#define B_MINSUB(i) asm volatile("{.reg .u64 t; sub.u64 t, %0, %1; min.u64 %0, %0, t;}" : "+l"(x[i]) : "l"(a64));
#define B_MULHI(i) asm volatile("mul.hi.u64 %0, %0, %1;" : "+l"(x[i]) : "l"(a64));
#define B_MULLO(i) asm volatile("mul.lo.u64 %0, %0, %1;" : "+l"(x[i]) : "l"(a64));
#define B_DFMA(i) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(x[i]) : "d"(ad));
#define B_FFMA(i) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x[i]) : "f"(af));
#define B_WIDE_IADD(i) WIDE(i) IADD2(i)
#define B_WIDE_2ALU(i) WIDE(i) IADD2(i) IADD2(i)
#define B_LO_IADD(i) LO(x, i) IADD2(i)
#define B_WIDE_LO(i) WIDE(i) LO(y, i)
#define B_DFMA_WIDE(i) WIDE(i) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(d[i]) : "d"(ad));

UBENCH(k_imad_wide, u64, B_WIDE)
UBENCH(k_imad_lo, u32, B_LO)
UBENCH(k_imad_hi, u32, B_HI)
UBENCH(k_iadd3, u32, B_IADD3)
UBENCH(k_add_u64, u64, B_ADD64)
UBENCH(k_lop3, u32, B_LOP3)
UBENCH(k_setp_selp, u32, B_SETP)
UBENCH(k_pred_sub64, u64, B_PSUB)
UBENCH(k_min_sub64, u64, B_MINSUB)
UBENCH(k_mulhi64, u64, B_MULHI)
UBENCH(k_mullo64, u64, B_MULLO)
UBENCH(k_dfma, double, B_DFMA)
UBENCH(k_ffma, float, B_FFMA)
UBENCH(k_wide_plus_iadd, u64, B_WIDE_IADD)
UBENCH(k_wide_plus_2alu, u64, B_WIDE_2ALU)
UBENCH(k_lo_plus_iadd, u32, B_LO_IADD)
UBENCH(k_wide_plus_lo, u64, B_WIDE_LO)
UBENCH(k_dfma_plus_wide, u64, B_DFMA_WIDE)

template <class K> void run(const char* name, K k, u64* d) {
    int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    const int blocks = p.multiProcessorCount * 2;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<blocks, 1024>>>(d, 3, 5); k<<<blocks, 1024>>>(d, 3, 5); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<<<blocks, 1024>>>(d, 3, 5); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bodies = (double)blocks * 1024 * ITERS * 64;
    printf("%-18s %8.3f ms  %7.2f bodies/clk/SM (clock %d MHz)  %s\n", name, ms,
           bodies / (ms * 1e-3) / (clk_khz * 1e3) / p.multiProcessorCount, clk_khz / 1000, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    u64* d; cudaMalloc(&d, 1024 * 8);
#define RUN(n) run(#n, k_##n, d);
    RUN(imad_wide) RUN(imad_lo) RUN(imad_hi) RUN(iadd3) RUN(add_u64) RUN(lop3) RUN(setp_selp) RUN(pred_sub64) RUN(min_sub64)
    RUN(mulhi64) RUN(mullo64) RUN(dfma) RUN(ffma) RUN(wide_plus_iadd) RUN(wide_plus_2alu) RUN(lo_plus_iadd) RUN(wide_plus_lo)
    RUN(dfma_plus_wide)
    return 0;
}
