/*
 * ckks_oracle.h -- CPU oracle for the orion-b200 CKKS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of the algorithms on the hot path named by BASELINE.json
 * (NTT/INTT, RNS base conversion, hybrid key switching, automorphisms, rescale, element-wise
 * RNS arithmetic, sampling, encoding).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product library (liborion_b200.so)
 * never links, loads or calls anything in this directory.
 *
 * PARITY STATUS
 *   - NTT/INTT: pinned.  Checked bit-for-bit against the reference's own CPU NTT
 *     (GPU-NTT ntt_cpu.cu, compiled in place into oracle/_ref/libgpuntt_ref.so) on the
 *     in-tree known-answer (prime, psi) pools, and against schoolbook multiplication.
 *   - Everything above the NTT (prime selection order, psi choice, float64 overflow estimate in
 *     the base conversion, digit layout, rounding in rescale / ModDown): "parity unpinned".
 *     The arithmetic lives in github.com/baahl-nyu/lattigo/v6 v6.2.0
 *     (orion/backend/lattigo/go.mod:5), which is not vendored under /root/reference and cannot
 *     be built here (no Go toolchain).  These functions restate the published Lattigo v6
 *     algorithms from memory of the upstream source; each cites the Go call site in the
 *     reference that reaches it.
 */
#ifndef CKKS_ORACLE_H
#define CKKS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef uint64_t u64;
typedef struct orc_ctx orc_ctx;

/* ---- number theory ---- */
int  orc_is_prime(u64 n);
u64  orc_powmod(u64 a, u64 e, u64 q);
u64  orc_invmod(u64 a, u64 q);
u64  orc_mulmod(u64 a, u64 b, u64 q);
/* Lattigo ring.NTTFriendlyPrimesGenerator.NextAlternatingPrimes (recalled): primes = 1 mod nthroot
 * alternating above/below 2^bits.  `skip`/`nskip`: primes already in use, never returned. */
int  orc_gen_ntt_primes(int bits, u64 nthroot, int count, const u64 *skip, int nskip, u64 *out);
/* Lattigo rlwe.GenModuli (recalled), reached from LG/scheme.go:57-65. */
int  orc_gen_moduli(const int *logQ, int nQ, const int *logP, int nP, int log_nthroot, u64 *Q, u64 *P);
u64  orc_primitive_root(u64 q);          /* Lattigo ring.PrimitiveRoot (recalled): smallest g >= 3 */
u64  orc_psi(u64 q, u64 nthroot);        /* g^((q-1)/nthroot) */

/* ---- context: a modulus chain (nQ "Q" primes followed by nP "P" primes) with NTT tables ---- */
orc_ctx *orc_ctx_new(int logN, const u64 *moduli, int nQ, int nP);
/* same, but with caller-chosen 2N-th roots (used to pin against the reference's KAT pools) */
orc_ctx *orc_ctx_new_with_psi(int logN, const u64 *moduli, const u64 *psi, int nQ, int nP);
void     orc_ctx_free(orc_ctx *c);
int      orc_ctx_N(const orc_ctx *c);
u64      orc_ctx_modulus(const orc_ctx *c, int idx);
u64      orc_ctx_psi(const orc_ctx *c, int idx);
void     orc_set_threads(int n);          /* OpenMP threads used by the bulk routines */
int      orc_get_threads(void);

/* ---- negacyclic NTT (natural order in, bit-reversed out) / INTT; in place, one limb ---- */
void orc_ntt(const orc_ctx *c, int mod_idx, u64 *a);
void orc_intt(const orc_ctx *c, int mod_idx, u64 *a);
/* batched over limbs: a is [n][N], mod_idx[n] */
void orc_ntt_many(const orc_ctx *c, const int *mod_idx, int n, u64 *a);
void orc_intt_many(const orc_ctx *c, const int *mod_idx, int n, u64 *a);
void orc_schoolbook_negacyclic(const u64 *a, const u64 *b, u64 *out, int N, u64 q);

/* ---- element-wise, one limb ---- */
void orc_add(const orc_ctx *c, int mod_idx, const u64 *a, const u64 *b, u64 *out);
void orc_sub(const orc_ctx *c, int mod_idx, const u64 *a, const u64 *b, u64 *out);
void orc_neg(const orc_ctx *c, int mod_idx, const u64 *a, u64 *out);
void orc_mul(const orc_ctx *c, int mod_idx, const u64 *a, const u64 *b, u64 *out);
void orc_mul_scalar(const orc_ctx *c, int mod_idx, const u64 *a, u64 s, u64 *out);
void orc_add_scalar(const orc_ctx *c, int mod_idx, const u64 *a, u64 s, u64 *out);

/* ---- automorphism X -> X^galEl ---- */
void orc_automorphism_index_ntt(const orc_ctx *c, u64 galEl, uint32_t *index);   /* out[i]=in[index[i]] */
void orc_automorphism_ntt(const orc_ctx *c, u64 galEl, const u64 *in, u64 *out); /* one limb, NTT domain */
void orc_automorphism_coeff(const orc_ctx *c, int mod_idx, u64 galEl, const u64 *in, u64 *out);
u64  orc_galois_element(const orc_ctx *c, int rot);       /* 5^rot mod 2N  (rot may be negative) */
u64  orc_galois_element_conj(const orc_ctx *c);           /* 2N-1 */

/* ---- RNS base conversion (Lattigo ring.ModUpExact / reconstructRNS, recalled) ----
 * in: [ns][N] coefficient domain on moduli src_idx; out: [nd][N] on moduli dst_idx.
 * centered != 0: add (Q_src/2) before and subtract after (ModUpQtoP / ModUpPtoQ behaviour). */
void orc_baseconv(const orc_ctx *c, const int *src_idx, int ns, const int *dst_idx, int nd,
                  const u64 *in, u64 *out, int centered);

/* ---- rescale (Lattigo ring.DivRoundByLastModulusNTT, recalled; LG/evaluator.go:84-99) ----
 * in: [level+1][N] NTT domain over Q_0..Q_level; out: [level][N]. */
void orc_rescale(const orc_ctx *c, int level, const u64 *in, u64 *out);

/* ---- hybrid key switching (Lattigo rlwe.Evaluator.GadgetProduct, recalled; LG/evaluator.go:61-81) ----
 * Limb order of every "QP at level" buffer: Q_0..Q_level then P_0..P_{nP-1}  -> (level+1+nP) limbs.
 * evk layout: [digit][2][nQ+nP][N] (full chain, NTT domain, standard representation). */
int  orc_num_digits(const orc_ctx *c, int level);
void orc_decompose(const orc_ctx *c, int level, const u64 *c1_ntt, u64 *digits /* [beta][level+1+nP][N] */);
void orc_inner_product(const orc_ctx *c, int level, const u64 *digits, const u64 *evk,
                       u64 *acc0, u64 *acc1 /* each [level+1+nP][N] */);
void orc_moddown(const orc_ctx *c, int level, const u64 *accQP, u64 *outQ /* [level+1][N] */);
void orc_keyswitch(const orc_ctx *c, int level, const u64 *c1_ntt, const u64 *evk, u64 *out0, u64 *out1);

/* ---- deterministic sampling (shared definition with the device samplers) ---- */
u64  orc_rng_u64(u64 seed, u64 stream, u64 index);                     /* counter-based */
void orc_sample_uniform(const orc_ctx *c, int mod_idx, u64 seed, u64 stream, u64 *out);
void orc_sample_gaussian_i64(int N, u64 seed, u64 stream, int64_t *out);      /* sigma 3.2, |x|<=19 */
void orc_sample_ternary_hw(int N, int h, u64 seed, u64 stream, int64_t *out); /* exactly h non-zeros */
void orc_small_to_rns(const orc_ctx *c, int mod_idx, const int64_t *in, u64 *out);
int  orc_gaussian_cdt(u64 *table, int max_entries);                     /* the CDT both sides use */

/* ---- scheme-level pieces (ckks_oracle_ckks.c): canonical-embedding FFT, encode rounding, decode CRT ---- */
void orc_roots(int M, double *re, double *im);
void orc_special_ifft(double *re, double *im, int log_slots, int M);
void orc_special_fft(double *re, double *im, int log_slots, int M);
void orc_encode_round(const orc_ctx *c, const double *re, const double *im, int log_slots, double scale,
                      const int *mods, int nmods, u64 *out);
void orc_scaled_const_rns(const orc_ctx *c, double v, double cscale_hi, double cscale_lo, const int *mods, int n, u64 *out);
void orc_decode_crt(const orc_ctx *c, const u64 *coeffs, int level, double scale, int log_slots, double *re, double *im);

#ifdef __cplusplus
}
#endif
#endif
