/*
 * ckks_oracle_ckks.c -- CPU oracle, scheme-level pieces: canonical-embedding FFT, encode rounding,
 * decode CRT.  TEST INFRASTRUCTURE ONLY (see ckks_oracle.h).  "parity unpinned" versus Lattigo: restated from
 * memory of Lattigo v6's ckks.Encoder (SpecialFFTDouble / SpecialIFFTDouble, Complex128ToFixedPointCRT), reached
 * from orion/backend/lattigo/encoder.go:16-42.  CPU FFT structure cross-checked against the reference's
 * GPU-FFT CPU code (thirdparty/GPU-FFT/src/lib/fft_cpu.cu) only for the butterfly shape.
 * Compile with -ffp-contract=off: every float64 operation is a single IEEE op in a fixed order.
 */
#include "ckks_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* roots[k] = exp(2*pi*i*k/M): first quadrant from cos(), the rest by symmetry */
void orc_roots(int M, double *re, double *im) {
    int quar = M >> 2;
    double angle = 2.0 * 3.141592653589793238462643383279502884 / (double)M;
    double *cs = (double *)malloc((quar + 1) * sizeof(double));
    for (int k = 0; k <= quar; k++) cs[k] = cos(angle * (double)k);
    for (int k = 0; k <= quar; k++) {
        double co = cs[k], si = cs[quar - k];
        re[k] = co; im[k] = si;
        re[2 * quar - k] = -co; im[2 * quar - k] = si;
        re[2 * quar + k] = -co; im[2 * quar + k] = -si;
        re[4 * quar - k] = co; im[4 * quar - k] = -si;
    }
    free(cs);
}

static void bitrev_perm(double *re, double *im, int n, int logn) {
    for (int i = 0; i < n; i++) {
        int j = 0, x = i;
        for (int b = 0; b < logn; b++) { j = (j << 1) | (x & 1); x >>= 1; }
        if (j > i) { double t = re[i]; re[i] = re[j]; re[j] = t; t = im[i]; im[i] = im[j]; im[j] = t; }
    }
}

/* Lattigo SpecialIFFTDouble: DIF stages len = n..2 with root index (lenq - 5^j mod lenq) * (M/lenq), then /n, bit-reverse */
void orc_special_ifft(double *re, double *im, int log_slots, int M) {
    int n = 1 << log_slots;
    double *rr = (double *)malloc((M + 1) * sizeof(double)), *ri = (double *)malloc((M + 1) * sizeof(double));
    int *rot = (int *)malloc(n * sizeof(int));
    orc_roots(M, rr, ri);
    long long p = 1;
    for (int j = 0; j < n; j++) { rot[j] = (int)p; p = (p * 5) % M; }
    for (int len = n; len >= 2; len >>= 1) {
        int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (int i = 0; i < n; i += len)
            for (int j = 0; j < lenh; j++) {
                int k = i + j, idx = (lenq - (rot[j] % lenq)) * gap;
                double ur = re[k] + re[k + lenh], ui = im[k] + im[k + lenh];
                double vr = re[k] - re[k + lenh], vi = im[k] - im[k + lenh];
                re[k] = ur; im[k] = ui;
                re[k + lenh] = vr * rr[idx] - vi * ri[idx];
                im[k + lenh] = vr * ri[idx] + vi * rr[idx];
            }
    }
    for (int i = 0; i < n; i++) { re[i] = re[i] / (double)n; im[i] = im[i] / (double)n; }
    bitrev_perm(re, im, n, log_slots);
    free(rr); free(ri); free(rot);
}

/* Lattigo SpecialFFTDouble: bit-reverse, then DIT stages len = 2..n with root index (5^j mod lenq) * (M/lenq) */
void orc_special_fft(double *re, double *im, int log_slots, int M) {
    int n = 1 << log_slots;
    double *rr = (double *)malloc((M + 1) * sizeof(double)), *ri = (double *)malloc((M + 1) * sizeof(double));
    int *rot = (int *)malloc(n * sizeof(int));
    orc_roots(M, rr, ri);
    long long p = 1;
    for (int j = 0; j < n; j++) { rot[j] = (int)p; p = (p * 5) % M; }
    bitrev_perm(re, im, n, log_slots);
    for (int len = 2; len <= n; len <<= 1) {
        int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (int i = 0; i < n; i += len)
            for (int j = 0; j < lenh; j++) {
                int k = i + j, idx = (rot[j] % lenq) * gap;
                double br = re[k + lenh], bi = im[k + lenh];
                double vr = br * rr[idx] - bi * ri[idx], vi = br * ri[idx] + bi * rr[idx];
                double ar = re[k], ai = im[k];
                re[k] = ar + vr; im[k] = ai + vi;
                re[k + lenh] = ar - vr; im[k + lenh] = ai - vi;
            }
    }
    free(rr); free(ri); free(rot);
}

static u64 double_to_mod(double x, u64 q) {          /* x >= 0, integer valued */
    if (x < 9223372036854775808.0) return (u64)x % q;
    union { double d; unsigned long long b; } u; u.d = x;
    int e = (int)((u.b >> 52) & 0x7ff) - 1075;
    u64 mant = (u.b & ((1ULL << 52) - 1)) | (1ULL << 52);
    u64 r = mant % q;
    for (; e > 0; e--) { r += r; if (r >= q) r -= q; }
    return r;
}

/* slots (after the special IFFT) -> coefficient-domain residues: real part of slot i at coefficient i*gap, imaginary
 * part at N/2 + i*gap; value = floor(|x*scale| + 0.5) with the sign restored (Lattigo scaleUpExact, recalled). */
void orc_encode_round(const orc_ctx *c, const double *re, const double *im, int log_slots, double scale,
                      const int *mods, int nmods, u64 *out) {
    int N = orc_ctx_N(c), half = N >> 1, gap = half >> log_slots;
    for (int t = 0; t < nmods; t++) {
        u64 q = orc_ctx_modulus(c, mods[t]);
        u64 *o = out + (size_t)t * N;
        for (int i = 0; i < N; i++) {
            int cc = i < half ? i : i - half;
            if (cc % gap) { o[i] = 0; continue; }
            double x = (i < half ? re[cc / gap] : im[cc / gap]) * scale;
            int neg = x < 0.0;
            double ax = floor((neg ? -x : x) + 0.5);
            u64 r = double_to_mod(ax, q);
            o[i] = (neg && r) ? q - r : r;
        }
    }
}

/* coefficient-domain residues over Q_0..Q_level -> centred value / scale per slot (before the special FFT).
 * Garner mixed-radix digits a_i (x = sum a_i * prod_{j<i} q_j), sign from the comparison with (Q-1)/2 whose
 * digits are (q_i-1)/2, magnitude evaluated in float64 from the top digit down. */
void orc_decode_crt(const orc_ctx *c, const u64 *coeffs, int level, double scale, int log_slots, double *re, double *im) {
    int N = orc_ctx_N(c), L = level + 1, half = N >> 1, slots = 1 << log_slots, gap = half >> log_slots;
    u64 q[64], inv[64][64]; double pd[64];
    double acc0 = 1.0;
    for (int i = 0; i < L; i++) { q[i] = orc_ctx_modulus(c, i); pd[i] = acc0; acc0 = acc0 * (double)q[i]; }
    for (int i = 0; i < L; i++) for (int j = 0; j < i; j++) inv[j][i] = orc_invmod(q[j] % q[i], q[i]);
    for (int s = 0; s < 2 * slots; s++) {
        int ci = s < slots ? s * gap : half + (s - slots) * gap;
        u64 a[64];
        for (int i = 0; i < L; i++) {
            u64 t = coeffs[(size_t)i * N + ci];
            for (int j = 0; j < i; j++) {
                u64 aj = a[j] % q[i];
                u64 d = t >= aj ? t - aj : t + q[i] - aj;
                t = (u64)(((u128)d * inv[j][i]) % q[i]);
            }
            a[i] = t;
        }
        int neg = 0;
        for (int i = L - 1; i >= 0; i--) { u64 h = (q[i] - 1) >> 1; if (a[i] != h) { neg = a[i] > h; break; } }
        if (neg) {
            int carry = 1;
            for (int i = 0; i < L; i++) {
                u64 b = q[i] - 1 - a[i];
                if (carry) { b += 1; if (b == q[i]) b = 0; else carry = 0; }
                a[i] = b;
            }
        }
        double acc = 0.0;
        for (int i = L - 1; i >= 0; i--) if (a[i]) acc = acc + (double)a[i] * pd[i];
        acc = (neg ? -acc : acc) / scale;
        if (s < slots) re[s] = acc; else im[s - slots] = acc;
    }
}

/* residues of sign(c) * roundl(|c * cscale|) for the listed moduli (scalar add / mul: LG/evaluator.go:102-179) */
void orc_scaled_const_rns(const orc_ctx *c, double v, double cscale_hi, double cscale_lo, const int *mods, int n, u64 *out) {
    /* the scale is a long double (64-bit mantissa) on both sides of the parity tests; it crosses the ctypes boundary as
     * hi + lo doubles because ctypes.c_longdouble silently narrows to a double */
    long double cscale = (long double)cscale_hi + (long double)cscale_lo;
    long double mag = roundl(fabsl((long double)v * cscale));
    for (int t = 0; t < n; t++) {
        u64 q = orc_ctx_modulus(c, mods[t]);
        u64 r = (u64)fmodl(mag, (long double)q);
        if (r >= q) r %= q;
        out[t] = (v < 0 && r) ? q - r : r;
    }
}
