// modarith.cuh -- 64-bit modular arithmetic for sm_100a kernels (and the host code that builds tables).
//
// All moduli are odd primes q < 2^62 so that the lazy ranges [0,4q) fit a word.
//   * Shoup   : x*w mod q for a known multiplicand w with companion w' = floor(w*2^64/q).
//               1 mul.hi.u64 + 2 mul.lo.u64.  Used for twiddles and every per-modulus constant.
//   * Barrett : data x data.  One-word variant keyed on the bit length b of q:
//               qhat = mulhi(x >> (b-2), mu), mu = floor(2^(62+b)/q); x < 2^(2b) -> r in [0,3q).
//   * Montgomery REDC for multiply-accumulate against operands stored pre-multiplied by 2^64
//               (evaluation keys, linear-transform diagonals): sum of products reduced once.
// The reference's GPU kernels use a two-shift Barrett for everything
// (GPU-NTT src/include/common/modular_arith.cuh:314-341); the split above is this backend's own.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ob {

typedef unsigned long long u64;
typedef unsigned __int128 u128;

// per-modulus constants, one entry per prime of the chain (device array, indexed by chain index)
struct ModConst {
    u64 q;         // modulus
    u64 mu;        // floor(2^(62+bits)/q)          (Barrett)
    u64 qinv_neg;  // -q^{-1} mod 2^64              (Montgomery)
    u64 r2;        // 2^128 mod q                   (to Montgomery form)
    u64 r2_sh;     // Shoup companion of r2
    u64 r1;        // 2^64 mod q
    u64 r1_sh;
    int bits;      // bit length of q
    int pad;
    double qd;     // (double) q
    u64 one_sh;    // floor(2^64/q): Shoup companion of 1 (reduces any 64-bit value to [0,2q))
    double qinv_d; // 1.0 / (double) q, correctly rounded (FP64 butterfly path)
    double rinv_d; // (double)(2^-64 mod q): takes a Montgomery factor out on the FP64 path (exact for q < 2^53)
};

#ifdef __CUDACC__
#define OB_HD __host__ __device__ __forceinline__
#define OB_D __device__ __forceinline__
#else
#define OB_HD inline
#define OB_D inline
#endif

OB_HD u64 mulhi64(u64 a, u64 b) {
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (u64)(((u128)a * b) >> 64);
#endif
}

// ---- Shoup ----
OB_HD u64 shoup_lazy(u64 x, u64 w, u64 wp, u64 q) {       // [0,2q), any x < 2^64
    u64 hi = mulhi64(x, wp);
    return x * w - hi * q;
}
OB_HD u64 shoup_mul(u64 x, u64 w, u64 wp, u64 q) {        // [0,q)
    u64 r = shoup_lazy(x, w, wp, q);
    return r >= q ? r - q : r;
}

// ---- helpers ----
OB_HD u64 add_mod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
OB_HD u64 sub_mod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
OB_HD u64 csub(u64 a, u64 q) { return a >= q ? a - q : a; }

// ---- Barrett on a 128-bit value x = hi:lo, x < 2^(bits+62) ----
OB_HD u64 barrett_red(u64 hi, u64 lo, const ModConst& m) {
    int s = m.bits - 2;
    u64 t = (s == 0) ? lo : ((lo >> s) | (hi << (64 - s)));
    u64 qh = mulhi64(t, m.mu);
    u64 r = lo - qh * m.q;
    r = r >= 2 * m.q ? r - 2 * m.q : r;
    return r >= m.q ? r - m.q : r;
}
OB_HD u64 barrett_mul(u64 a, u64 b, const ModConst& m) {
    return barrett_red(mulhi64(a, b), a * b, m);
}

// ---- Montgomery: REDC(hi:lo) = (hi:lo) * 2^-64 mod q, requires hi:lo < q*2^64; result in [0,q) ----
OB_HD u64 mont_redc(u64 hi, u64 lo, u64 q, u64 qinv_neg) {
    u64 mfac = lo * qinv_neg;
    u64 t = mulhi64(mfac, q);
    // lo + low64(mfac*q) == 0 mod 2^64 ; carry is 1 iff lo != 0
    u64 r = hi + t + (lo != 0);
    return r >= q ? r - q : r;
}

// 128-bit multiply-accumulate with the high word kept below q (acc stays < q*2^64)
OB_HD void mac128(u64& hi, u64& lo, u64 a, u64 b, u64 q) {
    u64 pl = a * b, ph = mulhi64(a, b);
    lo += pl;
    hi += ph + (lo < pl);
    hi = hi >= q ? hi - q : hi;
}

}  // namespace ob
