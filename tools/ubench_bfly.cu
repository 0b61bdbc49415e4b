// Butterfly micro-benchmark: Harvey/Shoup 64-bit integer butterfly vs FP64 (DFMA) butterfly for primes < 2^49.
// Reports butterflies / clk / SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o ubench_bfly ubench_bfly.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
constexpr int ITERS = 512;

__device__ __forceinline__ u64 shoup_lazy(u64 x, u64 w, u64 wp, u64 q) { return x * w - __umul64hi(x, wp) * q; }
__device__ __forceinline__ void ct_int(u64& a, u64& b, u64 w, u64 wp, u64 q, u64 q2) {
    u64 x = a >= q2 ? a - q2 : a; u64 v = shoup_lazy(b, w, wp, q); a = x + v; b = x - v + q2;
}
__device__ __forceinline__ void ct_f64(double& a, double& b, double w, double qinv, double q) {
    const double M = 6755399441055744.0;   // 1.5 * 2^52
    double h = __dmul_rn(b, w), l = __fma_rn(b, w, -h);
    double t = __dsub_rn(__fma_rn(h, qinv, M), M);
    double r = __dadd_rn(__fma_rn(-t, q, h), l);
    b = __dsub_rn(a, r); a = __dadd_rn(a, r);
}
__global__ void __launch_bounds__(256) k_int(u64* out, u64 q, u64 w0, u64 wp0) {
    u64 v[8]; for (int i = 0; i < 8; ++i) v[i] = (threadIdx.x * 8 + i + blockIdx.x) % q;
    u64 w = w0 + threadIdx.x, wp = wp0 + threadIdx.x; const u64 q2 = 2 * q;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int half = 4 >> s;
#pragma unroll
            for (int k = 0; k < 4; ++k) { int a = (k / half) * 2 * half + (k % half); ct_int(v[a], v[a + half], w, wp, q, q2); }
        }
    }
    u64 s = 0; for (int i = 0; i < 8; ++i) s += v[i];
    if (s == 0x123456789ull) out[threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_f64(u64* out, double q, double w0) {
    double v[8]; for (int i = 0; i < 8; ++i) v[i] = (double)((threadIdx.x * 8 + i + blockIdx.x) % 1000);
    double w = w0 + threadIdx.x; const double qinv = 1.0 / q;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int half = 4 >> s;
#pragma unroll
            for (int k = 0; k < 4; ++k) { int a = (k / half) * 2 * half + (k % half); ct_f64(v[a], v[a + half], w, qinv, q); }
        }
        // keep magnitudes bounded like a real NTT would after a pass (one reduction of a per 3 stages is NOT in a real kernel;
        // here it only stops the synthetic loop from overflowing): 8 extra cheap ops per 12 butterflies
#pragma unroll
        for (int i = 0; i < 8; ++i) { const double M = 6755399441055744.0; double t = __dsub_rn(__fma_rn(v[i], qinv, M), M); v[i] = __fma_rn(-t, q, v[i]); }
    }
    double s = 0; for (int i = 0; i < 8; ++i) s += v[i];
    if (s == 12345.678) out[threadIdx.x] = (u64)s;
}
int main() {
    int dev = 0; cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    u64* d; cudaMalloc(&d, 8192);
    const u64 q = (1ull << 45) - 229; 
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int which = 0; which < 2; ++which)
        for (int bpsm = 2; bpsm <= 8; bpsm *= 2) {
            const int blocks = p.multiProcessorCount * bpsm * 4;
            float ms = 0;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                if (which == 0) k_int<<<blocks, 256>>>(d, q, 12345678901ull, 987654321012345ull);
                else k_f64<<<blocks, 256>>>(d, (double)q, 1234567890123.0);
                cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
            }
            const double bf = (double)blocks * 256 * ITERS * 12;
            printf("%s blocks/SM=%d  %.3f ms  %.2f butterflies/clk/SM  %s\n", which ? "f64" : "int", bpsm * 4, ms,
                   bf / (ms * 1e-3) / (clk_khz * 1e3) / p.multiProcessorCount, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
