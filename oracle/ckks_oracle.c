/*
 * ckks_oracle.c -- CPU oracle for the orion-b200 CKKS hot path.  TEST INFRASTRUCTURE ONLY
 * (see ckks_oracle.h for the rules and the parity status: NTT pinned, the rest "parity unpinned").
 *
 * Path citations are relative to /root/reference.  LG/ = orion/backend/lattigo/,
 * NT/ = orion/backend/heongpu/adrianHEonGPU/thirdparty/GPU-NTT/.
 */
#include "ckks_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

static int g_threads = 1;
void orc_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
int  orc_get_threads(void) { return g_threads; }

/* ------------------------------------------------------------------------------------------
 * scalar modular arithmetic.  mulmod is the obviously-correct 128-bit remainder; the fast paths
 * (Shoup for known multiplicands, Barrett for data x data) are checked against it in
 * tests/test_oracle.py.  Host Barrett follows NT/src/include/common/modular_arith.cuh:58-152.
 * ---------------------------------------------------------------------------------------- */
u64 orc_mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }

u64 orc_powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q; a %= q;
    while (e) { if (e & 1) r = orc_mulmod(r, a, q); a = orc_mulmod(a, a, q); e >>= 1; }
    return r;
}
u64 orc_invmod(u64 a, u64 q) { return orc_powmod(a, q - 2, q); } /* q prime */

static inline u64 shoup_pre(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
/* w*x mod q for w < q with companion w' = floor(w*2^64/q); result in [0,q) */
static inline u64 shoup_mul(u64 x, u64 w, u64 wp, u64 q) {
    u64 hi = (u64)(((u128)x * wp) >> 64);
    u64 r = x * w - hi * q;
    return r >= q ? r - q : r;
}
typedef struct { u64 q; u64 mu_hi, mu_lo; } barrett_t;   /* mu = floor(2^128 / q) */
static barrett_t barrett_pre(u64 q) {
    barrett_t b; b.q = q;
    /* floor(2^128/q): long division of 2^128 by q in two 64-bit steps */
    u128 num = ((u128)1 << 64);
    u64 hi = (u64)(num / q);             /* floor(2^64/q) */
    u128 rem = num % q;
    u64 lo = (u64)((rem << 64) / q);
    b.mu_hi = hi; b.mu_lo = lo;
    return b;
}
/* reduce a 128-bit value x < q*2^64 to [0,q) */
static inline u64 barrett_red128(u128 x, const barrett_t *b) {
    u64 xl = (u64)x, xh = (u64)(x >> 64);
    /* qhat = floor(x * mu / 2^128), using the top three partial products */
    u128 t1 = (u128)xl * b->mu_hi;
    u128 t2 = (u128)xh * b->mu_lo;
    u128 t0 = (u128)xl * b->mu_lo;
    u128 mid = (t1 & (u128)0xFFFFFFFFFFFFFFFFULL) + (t2 & (u128)0xFFFFFFFFFFFFFFFFULL) + (t0 >> 64);
    u128 qhat = (u128)xh * b->mu_hi + (t1 >> 64) + (t2 >> 64) + (mid >> 64);
    u64 r = xl - (u64)qhat * b->q;
    while (r >= b->q) r -= b->q;
    return r;
}
static inline u64 barrett_mul(u64 a, u64 c, const barrett_t *b) { return barrett_red128((u128)a * c, b); }

/* ------------------------------------------------------------------------------------------
 * primes and roots
 * ---------------------------------------------------------------------------------------- */
int orc_is_prime(u64 n) {
    if (n < 2) return 0;
    static const u64 small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (int i = 0; i < 12; i++) { if (n == small[i]) return 1; if (n % small[i] == 0) return 0; }
    u64 d = n - 1; int s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (int i = 0; i < 12; i++) {                 /* deterministic for n < 2^64 */
        u64 x = orc_powmod(small[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) { x = orc_mulmod(x, x, n); if (x == n - 1) { comp = 0; break; } }
        if (comp) return 0;
    }
    return 1;
}

/* Lattigo ring.NTTFriendlyPrimesGenerator.NextAlternatingPrime (recalled): each round tests one
 * candidate above 2^bits (2^bits+1, +nthroot, ...) and then one below (2^bits+1-nthroot, ...). */
int orc_gen_ntt_primes(int bits, u64 nthroot, int count, const u64 *skip, int nskip, u64 *out) {
    u64 next = ((u64)1 << bits) + 1, prev = ((u64)1 << bits) + 1 - nthroot;
    int check_next = 1, check_prev = 1, found = 0;
    while (found < count) {
        if (!check_next && !check_prev) return -1;
        for (int dir = 0; dir < 2 && found < count; dir++) {
            u64 cand;
            if (dir == 0) {
                if (!check_next) continue;
                if ((next >> (bits + 1)) != 0) { check_next = 0; continue; }   /* left the bit-size window */
                cand = next; next += nthroot;
            } else {
                if (!check_prev) continue;
                if (prev < nthroot || (prev >> (bits - 1)) == 0) { check_prev = 0; continue; }
                cand = prev; prev -= nthroot;
            }
            if (!orc_is_prime(cand)) continue;
            int used = 0;
            for (int k = 0; k < nskip; k++) if (skip[k] == cand) used = 1;
            for (int k = 0; k < found; k++) if (out[k] == cand) used = 1;
            if (!used) out[found++] = cand;
        }
    }
    return 0;
}

/* Lattigo rlwe.GenModuli (recalled): count primes per bit-size over LogQ and LogP, generate that
 * many alternating primes per size, hand them out to Q in order, then to P. */
int orc_gen_moduli(const int *logQ, int nQ, const int *logP, int nP, int log_nthroot, u64 *Q, u64 *P) {
    int cnt[65] = {0}, taken[65] = {0};
    u64 *pool[65] = {0};
    for (int i = 0; i < nQ; i++) cnt[logQ[i]]++;
    for (int i = 0; i < nP; i++) cnt[logP[i]]++;
    int rc = 0;
    for (int b = 0; b < 65; b++) if (cnt[b]) {
        pool[b] = (u64 *)malloc(sizeof(u64) * cnt[b]);
        if (orc_gen_ntt_primes(b, (u64)1 << log_nthroot, cnt[b], NULL, 0, pool[b])) rc = -1;
    }
    if (!rc) {
        for (int i = 0; i < nQ; i++) Q[i] = pool[logQ[i]][taken[logQ[i]]++];
        for (int i = 0; i < nP; i++) P[i] = pool[logP[i]][taken[logP[i]]++];
    }
    for (int b = 0; b < 65; b++) free(pool[b]);
    return rc;
}

static u64 gcd64(u64 a, u64 b) { while (b) { u64 t = a % b; a = b; b = t; } return a; }
static u64 pollard_rho(u64 n) {
    if (!(n & 1)) return 2;
    for (u64 c = 1;; c++) {
        u64 x = 2, y = 2, d = 1;
        while (d == 1) {
            x = (orc_mulmod(x, x, n) + c) % n;
            y = (orc_mulmod(y, y, n) + c) % n; y = (orc_mulmod(y, y, n) + c) % n;
            d = gcd64(x > y ? x - y : y - x, n);
        }
        if (d != n) return d;
    }
}
static void factor_rec(u64 n, u64 *f, int *nf) {
    if (n == 1) return;
    if (orc_is_prime(n)) { for (int i = 0; i < *nf; i++) if (f[i] == n) return; f[(*nf)++] = n; return; }
    for (u64 p = 2; p < 1000; p++) if (n % p == 0) { factor_rec(p, f, nf); while (n % p == 0) n /= p; factor_rec(n, f, nf); return; }
    u64 d = pollard_rho(n);
    factor_rec(d, f, nf); factor_rec(n / d, f, nf);
}
/* Lattigo ring.PrimitiveRoot (recalled): g starts at 2 and is incremented before each test */
u64 orc_primitive_root(u64 q) {
    u64 f[64]; int nf = 0;
    factor_rec(q - 1, f, &nf);
    for (u64 g = 3;; g++) {
        int ok = 1;
        for (int i = 0; i < nf && ok; i++) if (orc_powmod(g, (q - 1) / f[i], q) == 1) ok = 0;
        if (ok) return g;
    }
}
u64 orc_psi(u64 q, u64 nthroot) { return orc_powmod(orc_primitive_root(q), (q - 1) / nthroot, q); }

/* ------------------------------------------------------------------------------------------
 * context
 * ---------------------------------------------------------------------------------------- */
struct orc_ctx {
    int logN, N, nQ, nP, nmod;
    u64 *mod, *psi;
    barrett_t *br;
    u64 **tw_f, **tw_f_sh, **tw_i, **tw_i_sh;   /* [nmod][N]: psi^bitrev(i), psi^-bitrev(i) */
    u64 *ninv, *ninv_sh;
};

static uint32_t bitrev32(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

orc_ctx *orc_ctx_new_with_psi(int logN, const u64 *moduli, const u64 *psi, int nQ, int nP) {
    orc_ctx *c = (orc_ctx *)calloc(1, sizeof(orc_ctx));
    c->logN = logN; c->N = 1 << logN; c->nQ = nQ; c->nP = nP; c->nmod = nQ + nP;
    int n = c->nmod, N = c->N;
    c->mod = (u64 *)malloc(n * 8); c->psi = (u64 *)malloc(n * 8);
    c->br = (barrett_t *)malloc(n * sizeof(barrett_t));
    c->tw_f = (u64 **)malloc(n * sizeof(u64 *)); c->tw_f_sh = (u64 **)malloc(n * sizeof(u64 *));
    c->tw_i = (u64 **)malloc(n * sizeof(u64 *)); c->tw_i_sh = (u64 **)malloc(n * sizeof(u64 *));
    c->ninv = (u64 *)malloc(n * 8); c->ninv_sh = (u64 *)malloc(n * 8);
    for (int m = 0; m < n; m++) {
        u64 q = moduli[m];
        c->mod[m] = q; c->br[m] = barrett_pre(q);
        c->psi[m] = psi ? psi[m] : orc_psi(q, (u64)2 * N);
        u64 psi_inv = orc_invmod(c->psi[m], q);
        c->tw_f[m] = (u64 *)malloc(N * 8); c->tw_f_sh[m] = (u64 *)malloc(N * 8);
        c->tw_i[m] = (u64 *)malloc(N * 8); c->tw_i_sh[m] = (u64 *)malloc(N * 8);
        u64 pf = 1, pi = 1;
        for (int i = 0; i < N; i++) {          /* table[bitrev(i)] = psi^i  (NT/src/lib/common/nttparameters.cu:168-212) */
            uint32_t j = bitrev32((uint32_t)i, logN);
            c->tw_f[m][j] = pf; c->tw_f_sh[m][j] = shoup_pre(pf, q);
            c->tw_i[m][j] = pi; c->tw_i_sh[m][j] = shoup_pre(pi, q);
            pf = orc_mulmod(pf, c->psi[m], q); pi = orc_mulmod(pi, psi_inv, q);
        }
        c->ninv[m] = orc_invmod((u64)N % q, q); c->ninv_sh[m] = shoup_pre(c->ninv[m], q);
    }
    return c;
}
orc_ctx *orc_ctx_new(int logN, const u64 *moduli, int nQ, int nP) {
    return orc_ctx_new_with_psi(logN, moduli, NULL, nQ, nP);
}
void orc_ctx_free(orc_ctx *c) {
    if (!c) return;
    for (int m = 0; m < c->nmod; m++) { free(c->tw_f[m]); free(c->tw_f_sh[m]); free(c->tw_i[m]); free(c->tw_i_sh[m]); }
    free(c->tw_f); free(c->tw_f_sh); free(c->tw_i); free(c->tw_i_sh);
    free(c->mod); free(c->psi); free(c->br); free(c->ninv); free(c->ninv_sh); free(c);
}
int orc_ctx_N(const orc_ctx *c) { return c->N; }
u64 orc_ctx_modulus(const orc_ctx *c, int idx) { return c->mod[idx]; }
u64 orc_ctx_psi(const orc_ctx *c, int idx) { return c->psi[idx]; }

/* ------------------------------------------------------------------------------------------
 * NTT / INTT.  Cooley-Tukey forward (natural in, bit-reversed out), Gentleman-Sande inverse,
 * twiddle psi^bitrev(m+i) at stage m, group i -- the loop structure of
 * NT/src/lib/ntt_merge/ntt_cpu.cu:81-129 (forward, X_N_plus branch) and :131-185 (inverse),
 * with Shoup multiplication for the table entries.
 * ---------------------------------------------------------------------------------------- */
void orc_ntt(const orc_ctx *c, int mi, u64 *a) {
    const u64 q = c->mod[mi]; const u64 *w = c->tw_f[mi], *ws = c->tw_f_sh[mi];
    int N = c->N, t = N;
    for (int m = 1; m < N; m <<= 1) {
        t >>= 1;
        for (int i = 0; i < m; i++) {
            u64 S = w[m + i], Sp = ws[m + i];
            u64 *x = a + 2 * i * t, *y = x + t;
            for (int j = 0; j < t; j++) {
                u64 U = x[j], V = shoup_mul(y[j], S, Sp, q);
                u64 s = U + V; x[j] = s >= q ? s - q : s;
                y[j] = U >= V ? U - V : U + q - V;
            }
        }
    }
}
void orc_intt(const orc_ctx *c, int mi, u64 *a) {
    const u64 q = c->mod[mi]; const u64 *w = c->tw_i[mi], *ws = c->tw_i_sh[mi];
    int N = c->N, t = 1;
    for (int m = N; m > 1; m >>= 1) {
        int h = m >> 1;
        for (int i = 0; i < h; i++) {
            u64 S = w[h + i], Sp = ws[h + i];
            u64 *x = a + 2 * i * t, *y = x + t;
            for (int j = 0; j < t; j++) {
                u64 U = x[j], V = y[j];
                u64 s = U + V; x[j] = s >= q ? s - q : s;
                u64 d = U >= V ? U - V : U + q - V;
                y[j] = shoup_mul(d, S, Sp, q);
            }
        }
        t <<= 1;
    }
    for (int j = 0; j < N; j++) a[j] = shoup_mul(a[j], c->ninv[mi], c->ninv_sh[mi], q);
}
void orc_ntt_many(const orc_ctx *c, const int *mod_idx, int n, u64 *a) {
    #pragma omp parallel for num_threads(g_threads) schedule(dynamic)
    for (int i = 0; i < n; i++) orc_ntt(c, mod_idx[i], a + (size_t)i * c->N);
}
void orc_intt_many(const orc_ctx *c, const int *mod_idx, int n, u64 *a) {
    #pragma omp parallel for num_threads(g_threads) schedule(dynamic)
    for (int i = 0; i < n; i++) orc_intt(c, mod_idx[i], a + (size_t)i * c->N);
}
/* method of NT/example/ntt_merge/test_cpu_merge_ntt.cu:69-101 / NT/src/lib/ntt_merge/ntt_cpu.cu:10-52 */
void orc_schoolbook_negacyclic(const u64 *a, const u64 *b, u64 *out, int N, u64 q) {
    u64 *t = (u64 *)calloc(2 * N, 8);
    for (int i = 0; i < N; i++) for (int j = 0; j < N; j++)
        t[i + j] = (t[i + j] + orc_mulmod(a[i], b[j], q)) % q;
    for (int i = 0; i < N; i++) out[i] = (t[i] + q - t[i + N]) % q;
    free(t);
}

/* ------------------------------------------------------------------------------------------
 * element-wise (LG/evaluator.go:48-294 reach Lattigo ring.Add/Sub/MulCoeffs per limb)
 * ---------------------------------------------------------------------------------------- */
void orc_add(const orc_ctx *c, int mi, const u64 *a, const u64 *b, u64 *o) {
    u64 q = c->mod[mi]; for (int j = 0; j < c->N; j++) { u64 s = a[j] + b[j]; o[j] = s >= q ? s - q : s; }
}
void orc_sub(const orc_ctx *c, int mi, const u64 *a, const u64 *b, u64 *o) {
    u64 q = c->mod[mi]; for (int j = 0; j < c->N; j++) o[j] = a[j] >= b[j] ? a[j] - b[j] : a[j] + q - b[j];
}
void orc_neg(const orc_ctx *c, int mi, const u64 *a, u64 *o) {
    u64 q = c->mod[mi]; for (int j = 0; j < c->N; j++) o[j] = a[j] ? q - a[j] : 0;
}
void orc_mul(const orc_ctx *c, int mi, const u64 *a, const u64 *b, u64 *o) {
    const barrett_t *br = &c->br[mi]; for (int j = 0; j < c->N; j++) o[j] = barrett_mul(a[j], b[j], br);
}
void orc_mul_scalar(const orc_ctx *c, int mi, const u64 *a, u64 s, u64 *o) {
    u64 q = c->mod[mi]; s %= q; u64 sp = shoup_pre(s, q);
    for (int j = 0; j < c->N; j++) o[j] = shoup_mul(a[j], s, sp, q);
}
void orc_add_scalar(const orc_ctx *c, int mi, const u64 *a, u64 s, u64 *o) {
    u64 q = c->mod[mi]; s %= q; for (int j = 0; j < c->N; j++) { u64 t = a[j] + s; o[j] = t >= q ? t - q : t; }
}

/* ------------------------------------------------------------------------------------------
 * automorphisms.  NTT-domain index (Lattigo ring.AutomorphismNTTIndex, recalled): slot i of the
 * bit-reversed NTT holds f(psi^(2*bitrev(i)+1)), so out[i] = in[i'] with
 * 2*bitrev(i')+1 = galEl*(2*bitrev(i)+1) mod 2N.  Reached from LG/evaluator.go:61-81 (Rotate).
 * ---------------------------------------------------------------------------------------- */
void orc_automorphism_index_ntt(const orc_ctx *c, u64 galEl, uint32_t *index) {
    int N = c->N, logN = c->logN; u64 mask = 2 * (u64)N - 1;
    for (int i = 0; i < N; i++) {
        u64 e = (galEl * (2 * (u64)bitrev32(i, logN) + 1)) & mask;
        index[i] = bitrev32((uint32_t)((e - 1) >> 1), logN);
    }
}
void orc_automorphism_ntt(const orc_ctx *c, u64 galEl, const u64 *in, u64 *out) {
    uint32_t *idx = (uint32_t *)malloc(c->N * 4);
    orc_automorphism_index_ntt(c, galEl, idx);
    for (int i = 0; i < c->N; i++) out[i] = in[idx[i]];
    free(idx);
}
/* coefficient domain: X^i -> X^(i*galEl mod 2N), sign flip past N (HE/.../switchkey.cu:1438-1446 analogue) */
void orc_automorphism_coeff(const orc_ctx *c, int mi, u64 galEl, const u64 *in, u64 *out) {
    int N = c->N; u64 q = c->mod[mi], mask = 2 * (u64)N - 1;
    for (int i = 0; i < N; i++) {
        u64 e = ((u64)i * galEl) & mask;
        if (e < (u64)N) out[e] = in[i]; else out[e - N] = in[i] ? q - in[i] : 0;
    }
}
/* Lattigo rlwe.Parameters.GaloisElement(k) (recalled): GaloisGen^k mod 2N with GaloisGen = 5,
 * negative k taken modulo the order N/2 of 5.  LG/evaluator.go:34-36. */
u64 orc_galois_element(const orc_ctx *c, int rot) {
    u64 twoN = 2 * (u64)c->N; int ord = c->N / 2;
    int k = ((rot % ord) + ord) % ord;
    u64 g = 1; for (int i = 0; i < k; i++) g = (g * 5) & (twoN - 1);
    return g;
}
u64 orc_galois_element_conj(const orc_ctx *c) { return 2 * (u64)c->N - 1; }

/* ------------------------------------------------------------------------------------------
 * RNS base conversion.  Lattigo ring.ModUpExact + reconstructRNS + multSum (recalled):
 *   y_i = x_i * (Q/q_i)^-1 mod q_i ;  v = (uint64) sum_i float64(y_i)/float64(q_i)   (in order i)
 *   out_t = ( sum_i y_i * (Q/q_i mod p_t)  -  v * (Q mod p_t) ) mod p_t
 * centered: x += floor(Q/2) first, out -= floor(Q/2) after (ring.BasisExtender.ModUpQtoP/ModUpPtoQ).
 * A single source modulus is a plain copy + reduction (DecomposeAndSplit's first branch).
 * ---------------------------------------------------------------------------------------- */
typedef struct { int ns, nd; u64 *inv, *inv_sh, *mat /* [ns][nd] */, *qmod /* [nd] Q mod p_t */,
                 *half_src /* [ns] */, *half_dst /* [nd] */; } bc_tab;

static void bc_build(const orc_ctx *c, const int *src, int ns, const int *dst, int nd, bc_tab *t) {
    t->ns = ns; t->nd = nd;
    t->inv = (u64 *)malloc(ns * 8); t->inv_sh = (u64 *)malloc(ns * 8);
    t->mat = (u64 *)malloc((size_t)ns * nd * 8); t->qmod = (u64 *)malloc(nd * 8);
    t->half_src = (u64 *)malloc(ns * 8); t->half_dst = (u64 *)malloc(nd * 8);
    for (int i = 0; i < ns; i++) {
        u64 qi = c->mod[src[i]], prod = 1;
        for (int k = 0; k < ns; k++) if (k != i) prod = orc_mulmod(prod, c->mod[src[k]] % qi, qi);
        t->inv[i] = orc_invmod(prod, qi); t->inv_sh[i] = shoup_pre(t->inv[i], qi);
    }
    for (int d = 0; d < nd; d++) {
        u64 p = c->mod[dst[d]], all = 1;
        for (int i = 0; i < ns; i++) {
            u64 prod = 1;
            for (int k = 0; k < ns; k++) if (k != i) prod = orc_mulmod(prod, c->mod[src[k]] % p, p);
            t->mat[(size_t)i * nd + d] = prod;
            all = orc_mulmod(all, c->mod[src[i]] % p, p);
        }
        t->qmod[d] = all;
    }
    /* floor(Q/2) mod m for every source and destination modulus: Q odd, so floor(Q/2) = (Q-1)/2
     * and ((Q-1)/2) mod m = ((Q mod m) - 1) * inv2 mod m  (m odd). */
    for (int i = 0; i < ns; i++) {
        u64 m = c->mod[src[i]];       /* Q = 0 mod m  ->  (Q-1)/2 = -1 * inv2 */
        u64 inv2 = (m + 1) / 2;
        t->half_src[i] = orc_mulmod(m - 1, inv2, m);
    }
    for (int d = 0; d < nd; d++) {
        u64 m = c->mod[dst[d]], inv2 = (m + 1) / 2;
        t->half_dst[d] = orc_mulmod((t->qmod[d] + m - 1) % m, inv2, m);
    }
}
static void bc_free(bc_tab *t) { free(t->inv); free(t->inv_sh); free(t->mat); free(t->qmod); free(t->half_src); free(t->half_dst); }

void orc_baseconv(const orc_ctx *c, const int *src, int ns, const int *dst, int nd,
                  const u64 *in, u64 *out, int centered) {
    int N = c->N;
    if (ns == 1 && !centered) {
        u64 qs = c->mod[src[0]];
        for (int d = 0; d < nd; d++) {
            u64 p = c->mod[dst[d]];
            for (int x = 0; x < N; x++) out[(size_t)d * N + x] = (p >= qs) ? in[x] : in[x] % p;
        }
        return;
    }
    bc_tab t; bc_build(c, src, ns, dst, nd, &t);
    #pragma omp parallel for num_threads(g_threads) schedule(static)
    for (int x = 0; x < N; x++) {
        u64 y[64]; double vf = 0.0;
        for (int i = 0; i < ns; i++) {
            u64 qi = c->mod[src[i]], xi = in[(size_t)i * N + x];
            if (centered) { xi += t.half_src[i]; if (xi >= qi) xi -= qi; }
            y[i] = shoup_mul(xi, t.inv[i], t.inv_sh[i], qi);
            vf += (double)y[i] / (double)qi;
        }
        u64 v = (u64)vf;
        for (int d = 0; d < nd; d++) {
            const barrett_t *br = &c->br[dst[d]]; u64 p = br->q;
            u128 acc = 0;
            for (int i = 0; i < ns; i++) acc += (u128)y[i] * t.mat[(size_t)i * nd + d];
            u64 r = barrett_red128(acc, br);
            u64 vq = barrett_mul(v, t.qmod[d], br);
            r = r >= vq ? r - vq : r + p - vq;
            if (centered) r = r >= t.half_dst[d] ? r - t.half_dst[d] : r + p - t.half_dst[d];
            out[(size_t)d * N + x] = r;
        }
    }
    bc_free(&t);
}

/* ------------------------------------------------------------------------------------------
 * rescale: Lattigo ring.DivRoundByLastModulusNTT (recalled), reached from LG/evaluator.go:84-99.
 * HEonGPU analogue for structure: HE/src/heongpu/lib/host/ckks/operator.cu:906-980.
 * ---------------------------------------------------------------------------------------- */
void orc_rescale(const orc_ctx *c, int level, const u64 *in, u64 *out) {
    int N = c->N; u64 ql = c->mod[level], half = (ql - 1) >> 1;
    u64 *last = (u64 *)malloc(N * 8);
    memcpy(last, in + (size_t)level * N, N * 8);
    orc_intt(c, level, last);
    for (int x = 0; x < N; x++) { u64 s = last[x] + half; last[x] = s >= ql ? s - ql : s; }
    #pragma omp parallel for num_threads(g_threads) schedule(dynamic)
    for (int i = 0; i < level; i++) {
        u64 qi = c->mod[i], hq = half % qi;
        u64 qlinv = orc_invmod(ql % qi, qi), qlinv_sh = shoup_pre(qlinv, qi);
        u64 *tmp = (u64 *)malloc(N * 8);
        for (int x = 0; x < N; x++) { u64 u = last[x] % qi; tmp[x] = u >= hq ? u - hq : u + qi - hq; }
        orc_ntt(c, i, tmp);
        for (int x = 0; x < N; x++) {
            u64 a = in[(size_t)i * N + x], d = a >= tmp[x] ? a - tmp[x] : a + qi - tmp[x];
            out[(size_t)i * N + x] = shoup_mul(d, qlinv, qlinv_sh, qi);
        }
        free(tmp);
    }
    free(last);
}

/* ------------------------------------------------------------------------------------------
 * hybrid key switching: Lattigo rlwe.Evaluator.GadgetProduct -> DecomposeNTT / DecomposeSingleNTT
 * -> ring.Decomposer.DecomposeAndSplit, then ModDownQPtoQNTT (all recalled).  Digit j holds the
 * nP consecutive primes Q_{j*nP} .. Q_{(j+1)*nP-1} (clipped to the level).  Reached from
 * LG/evaluator.go:61-81 (Rotate), :297-317 (MulRelin), LG/lineartransform.go:96-113.
 * HEonGPU analogue for structure only: HE/src/heongpu/lib/host/ckks/operator.cu:1308-1429.
 * ---------------------------------------------------------------------------------------- */
int orc_num_digits(const orc_ctx *c, int level) { return (level + 1 + c->nP - 1) / c->nP; }

void orc_decompose(const orc_ctx *c, int level, const u64 *c1_ntt, u64 *digits) {
    int N = c->N, L = level + 1, K = c->nP, W = L + K, beta = orc_num_digits(c, level);
    u64 *coef = (u64 *)malloc((size_t)L * N * 8);
    memcpy(coef, c1_ntt, (size_t)L * N * 8);
    int idxs[64]; for (int i = 0; i < L; i++) idxs[i] = i;
    orc_intt_many(c, idxs, L, coef);
    for (int j = 0; j < beta; j++) {
        int s0 = j * K, s1 = (j + 1) * K < L ? (j + 1) * K : L, ns = s1 - s0;
        u64 *dj = digits + (size_t)j * W * N;
        int src[64], dst[128], nd = 0;
        for (int i = s0; i < s1; i++) src[i - s0] = i;
        for (int t = 0; t < L; t++) if (t < s0 || t >= s1) dst[nd++] = t;
        for (int t = 0; t < K; t++) dst[nd++] = c->nQ + t;
        u64 *conv = (u64 *)malloc((size_t)nd * N * 8);
        orc_baseconv(c, src, ns, dst, nd, coef + (size_t)s0 * N, conv, 0);
        orc_ntt_many(c, dst, nd, conv);
        int k = 0;
        for (int t = 0; t < W; t++) {
            int own = (t >= s0 && t < s1);
            if (own) memcpy(dj + (size_t)t * N, c1_ntt + (size_t)t * N, N * 8);
            else     memcpy(dj + (size_t)t * N, conv + (size_t)(k++) * N, N * 8);
        }
        free(conv);
    }
    free(coef);
}

void orc_inner_product(const orc_ctx *c, int level, const u64 *digits, const u64 *evk, u64 *acc0, u64 *acc1) {
    int N = c->N, L = level + 1, K = c->nP, W = L + K, beta = orc_num_digits(c, level), full = c->nmod;
    #pragma omp parallel for num_threads(g_threads) schedule(dynamic)
    for (int t = 0; t < W; t++) {
        int mi = t < L ? t : c->nQ + (t - L);
        const barrett_t *br = &c->br[mi];
        for (int x = 0; x < N; x++) {
            u64 r0 = 0, r1 = 0;
            for (int j = 0; j < beta; j++) {
                u64 d = digits[((size_t)j * W + t) * N + x];
                const u64 *k0 = evk + (((size_t)j * 2 + 0) * full + mi) * N;
                const u64 *k1 = evk + (((size_t)j * 2 + 1) * full + mi) * N;
                u64 a = barrett_mul(d, k0[x], br), b = barrett_mul(d, k1[x], br);
                r0 += a; if (r0 >= br->q) r0 -= br->q;
                r1 += b; if (r1 >= br->q) r1 -= br->q;
            }
            acc0[(size_t)t * N + x] = r0; acc1[(size_t)t * N + x] = r1;
        }
    }
}

void orc_moddown(const orc_ctx *c, int level, const u64 *accQP, u64 *outQ) {
    int N = c->N, L = level + 1, K = c->nP;
    u64 *p = (u64 *)malloc((size_t)K * N * 8), *conv = (u64 *)malloc((size_t)L * N * 8);
    memcpy(p, accQP + (size_t)L * N, (size_t)K * N * 8);
    int src[64], dst[64];
    for (int i = 0; i < K; i++) src[i] = c->nQ + i;
    for (int i = 0; i < L; i++) dst[i] = i;
    orc_intt_many(c, src, K, p);
    orc_baseconv(c, src, K, dst, L, p, conv, 1);
    orc_ntt_many(c, dst, L, conv);
    #pragma omp parallel for num_threads(g_threads) schedule(dynamic)
    for (int i = 0; i < L; i++) {
        u64 qi = c->mod[i], P = 1;
        for (int k = 0; k < K; k++) P = orc_mulmod(P, c->mod[c->nQ + k] % qi, qi);
        u64 pinv = orc_invmod(P, qi), pinv_sh = shoup_pre(pinv, qi);
        for (int x = 0; x < N; x++) {
            u64 a = accQP[(size_t)i * N + x], b = conv[(size_t)i * N + x];
            u64 d = a >= b ? a - b : a + qi - b;
            outQ[(size_t)i * N + x] = shoup_mul(d, pinv, pinv_sh, qi);
        }
    }
    free(p); free(conv);
}

void orc_keyswitch(const orc_ctx *c, int level, const u64 *c1_ntt, const u64 *evk, u64 *out0, u64 *out1) {
    int N = c->N, L = level + 1, K = c->nP, W = L + K, beta = orc_num_digits(c, level);
    u64 *dig = (u64 *)malloc((size_t)beta * W * N * 8);
    u64 *a0 = (u64 *)malloc((size_t)W * N * 8), *a1 = (u64 *)malloc((size_t)W * N * 8);
    orc_decompose(c, level, c1_ntt, dig);
    orc_inner_product(c, level, dig, evk, a0, a1);
    orc_moddown(c, level, a0, out0);
    orc_moddown(c, level, a1, out1);
    free(dig); free(a0); free(a1);
}

/* ------------------------------------------------------------------------------------------
 * deterministic sampling.  Lattigo seeds a blake2b XOF from the OS and exposes no seed through
 * LG/scheme.go:35-44, so key/encryption randomness can never be bit-matched against it; the
 * oracle and the device share this counter-based generator instead (same function on both
 * sides -> identical keys and ciphertexts for identical seeds).
 * ---------------------------------------------------------------------------------------- */
static inline u64 mix64(u64 z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL; z ^= z >> 27; z *= 0x94D049BB133111EBULL; z ^= z >> 31; return z;
}
u64 orc_rng_u64(u64 seed, u64 stream, u64 index) {
    u64 x = mix64(seed ^ (stream * 0xD6E8FEB86659FD93ULL));
    x = mix64(x ^ (index * 0x9E3779B97F4A7C15ULL));
    return mix64(x + 0x9E3779B97F4A7C15ULL);
}
void orc_sample_uniform(const orc_ctx *c, int mi, u64 seed, u64 stream, u64 *out) {
    u64 q = c->mod[mi]; int bits = 64 - __builtin_clzll(q);
    u64 mask = bits >= 64 ? ~0ULL : (((u64)1 << bits) - 1);
    for (int i = 0; i < c->N; i++) {
        u64 r = 0; int a;
        for (a = 0; a < 64; a++) { r = orc_rng_u64(seed, stream, ((u64)a << 32) | (u64)i) & mask; if (r < q) break; }
        out[i] = a < 64 ? r : r % q;
    }
}
static const u64 GAUSS_CDT[20] = {
    0x1FEA56814B22FE3AULL, 0x5CB44BA97E1C801CULL, 0x9135C4D2AB6097ACULL, 0xBA57841EC437E30AULL,
    0xD790D29879901FE9ULL, 0xEA65B1358D74FB9DULL, 0xF5672D96E886CF31ULL, 0xFB3C9D22CC87D60FULL,
    0xFE0A92B033D640A2ULL, 0xFF439D38007352AEULL, 0xFFBF67E425D189D1ULL, 0xFFEBCDF3A4629F98ULL,
    0xFFFA3F2F78880D45ULL, 0xFFFE81F40111A23BULL, 0xFFFFA5D06BC38AFBULL, 0xFFFFECA4870DFE3CULL,
    0xFFFFFC3B68ED3B4AULL, 0xFFFFFF58143B9611ULL, 0xFFFFFFE850CE70A6ULL, 0xFFFFFFFFFFFFFFFFULL};
int orc_gaussian_cdt(u64 *table, int max_entries) {
    int n = max_entries < 20 ? max_entries : 20; memcpy(table, GAUSS_CDT, n * 8); return 20;
}
/* discrete Gaussian, sigma = 3.2 truncated at 19 (Lattigo's DefaultNoise: sigma 3.2, bound 6 sigma) */
void orc_sample_gaussian_i64(int N, u64 seed, u64 stream, int64_t *out) {
    for (int i = 0; i < N; i++) {
        u64 r = orc_rng_u64(seed, stream, (u64)i), s = orc_rng_u64(seed, stream, ((u64)1 << 32) | (u64)i);
        int k = 0; while (r > GAUSS_CDT[k]) k++;
        out[i] = (s & 1) ? -(int64_t)k : (int64_t)k;
    }
}
/* ternary with exactly h non-zero coefficients (Lattigo ring.Ternary{H: h}; LG/scheme.go:62) */
void orc_sample_ternary_hw(int N, int h, u64 seed, u64 stream, int64_t *out) {
    int *pos = (int *)malloc(N * sizeof(int));
    for (int i = 0; i < N; i++) { pos[i] = i; out[i] = 0; }
    if (h > N) h = N;
    for (int i = 0; i < h; i++) {
        int j = i + (int)(orc_rng_u64(seed, stream, (u64)i) % (u64)(N - i));
        int t = pos[i]; pos[i] = pos[j]; pos[j] = t;
        out[pos[i]] = (orc_rng_u64(seed, stream, ((u64)1 << 32) | (u64)i) & 1) ? -1 : 1;
    }
    free(pos);
}
void orc_small_to_rns(const orc_ctx *c, int mi, const int64_t *in, u64 *out) {
    u64 q = c->mod[mi];
    for (int i = 0; i < c->N; i++) out[i] = in[i] >= 0 ? (u64)in[i] % q : q - ((u64)(-in[i]) % q);
}
