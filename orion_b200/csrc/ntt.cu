// ntt.cu -- batched negacyclic NTT / INTT over RNS limbs for sm_100a.
//
// Transform definition (identical to the reference CPU NTT, GPU-NTT src/lib/ntt_merge/ntt_cpu.cu:81-185,
// X^N+1 branch): Cooley-Tukey forward, natural order in / bit-reversed out, twiddle psi^bitrev(m+i) for
// stage m group i; Gentleman-Sande inverse with psi^-1 and a final N^-1.
//
// Decomposition ("4-step" split N = R x M, done as two launches per direction):
//   forward : K_strided  does the first s1 = logN-8 stages on 2^s1 rows that are N/2^s1 apart; one CTA
//             owns a [2^s1 rows] x [C columns] tile (C*8 B contiguous per row -> coalesced),
//             K_contig   does the last 8 stages on contiguous 256-word chunks, 8 chunks per CTA.
//   inverse : the mirror image (contiguous first, strided second, N^-1 folded into the last pass).
// Inside a CTA every thread keeps 8 coefficients in registers and performs up to 3 butterfly stages
// (radix-8) per round; rounds exchange data through XOR-swizzled shared memory.  Butterflies are
// Harvey lazy butterflies on [0,4q) (forward) / [0,2q) (inverse) with Shoup twiddles
// (1 mul.hi.u64 + 2 mul.lo.u64 each); the last pass normalises to [0,q).
// The reference GPU path (NT/src/lib/ntt_merge/ntt.cu:179,709) does one butterfly per thread per
// __syncthreads with Barrett multiplication; this is a different design, not a port.
#include "context.h"
#include "kernels.h"
#include <cooperative_groups.h>
#include <type_traits>

namespace ob {

namespace {

constexpr int LOG_E_STRIDED = 11;   // 2048 words per tile, 256 threads
constexpr int LOG_E_CONTIG = 10;    // 1024 words per tile, 128 threads
constexpr int K2_BITS = 8;          // contiguous pass handles the 8 low index bits
constexpr int IP_STAGES_DEFAULT = 0;  // key ring of ntt2_ip_kernel<ArithF64> (see there)

__device__ __forceinline__ int insert_bits(int x, int pos, int nb, int v) {
    return ((x >> pos) << (pos + nb)) | (v << pos) | (x & ((1 << pos) - 1));
}
// Shared-memory layout: word li of a tile lives at pad(li) = li + (li >> 4) + 2 (li >> 5).  Same bank behaviour as an
// XOR swizzle of the low index bits for every round shape (brute-forced on the host), but ADDITIVE: for the 8 words a
// thread owns in a round, li = base | (k << pos) with the k bits of base zero, so pad(li) = pad(base) + pad(k << pos)
// exactly (no carries) -- one address computation per thread and round, compile-time offsets for the rest.
// Coefficients per thread (LOG_EPT): 16 in the strided pass (radix-16 rounds: ONE exchange per 8 stages, measured 58 -> 48 us
// on 192 limbs), 8 in the contiguous pass (16 there means 128-byte-strided stores and twiddle loads: measured slower).
// The inverse strided pass (final, N^-1 scaling) measured slower with 16 (63 -> 107 us), so it keeps 8 as well.
__host__ __device__ constexpr int log_ept(bool strided, bool fwd) { return (strided && fwd) ? 4 : 3; }
template <int LOG_EPT>
__host__ __device__ constexpr int pad(int li) { return LOG_EPT == 4 ? li + (li >> 4) : li + (li >> 4) + ((li >> 5) << 1); }

// 256-bit global accesses (sm_100: LDG/STG.E.ENL2.256): a lane moves a whole 32-byte sector
__device__ __forceinline__ void ld_global_256(const u64* p, u64& a, u64& b, u64& c, u64& d) {
    asm volatile("ld.global.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}
__device__ __forceinline__ void st_global_256(u64* p, u64 a, u64 b, u64 c, u64 d) {
    asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}

// ---- arithmetic policies ------------------------------------------------------------------------------------
// ArithInt : Harvey lazy butterflies on u64 with Shoup twiddles (any prime < 2^62).  6 IMAD.WIDE + 4 IMAD per
//            butterfly; IMAD.WIDE issues at 32 lanes/clk/SM on B200 (tools/ubench_int.cu), so this path tops out
//            at 2.8 butterflies/clk/SM (tools/ubench_bfly.cu).
// ArithF64 : primes of at most F64_MAX_BITS bits.  Residues are carried as exact integers in doubles, in a signed
//            lazy range; x*w mod q is  h = x*w, l = fma(x,w,-h), t = rint(h/q), r = fma(-t,q,h) + l  (6 DFMA-class
//            ops, all exact -- see the bounds below) and B200 issues DFMA at 64 lanes/clk/SM, which measures
//            6.2 butterflies/clk/SM.  Results are normalised to [0,q) on the way out, so both paths produce the
//            same words.
//   bounds (q < 2^46, logN <= 17): forward values grow by <= 0.7q per stage -> |v| < 13q < 2^49.7 (no reduction
//   needed); rint() via the 1.5*2^52 trick needs |h/q| < 2^51; h - t*q is an integer of magnitude < q + ulp(h)/2
//   < 2^47, so the fma and the final add are exact.  The inverse doubles the sum path each stage, so every value
//   is re-centred at the end of a radix-8 round (|v| <= 8 * 0.7q inside a round).
constexpr int F64_MAX_BITS = 46;
constexpr double F64_MAGIC = 6755399441055744.0;            // 1.5 * 2^52
constexpr unsigned long long F64_EXP52 = 0x4330000000000000ull;   // bit pattern of 2^52

struct ArithInt {
    typedef u64 T;
    typedef ulonglong2 TW;
    typedef ulonglong2 Acc;                              // fused inner product: 128-bit lazy accumulator (x = low, y = high < q)
    u64 q, q2, qinv_neg;
    __device__ __forceinline__ explicit ArithInt(const ModConst& m) : q(m.q), q2(2 * m.q), qinv_neg(m.qinv_neg) {}
    static constexpr int IP_MIN_CTAS = 3;                // fused inner product: resident CTAs per SM the register allocation aims for
    typedef u64 D;                                       // multiply-accumulate operand (a transformed digit word)
    __device__ __forceinline__ D mac_operand(T d) const { return d; }
    __device__ __forceinline__ Acc acc_zero() const { return make_ulonglong2(0, 0); }
    __device__ __forceinline__ T keep(T x) const { return final_fwd(x); }
    __device__ __forceinline__ u64 kept_word(T x) const { return x; }
    __device__ __forceinline__ void mac(Acc& a, T d, u64 k) const { mac128(a.y, a.x, d, k, q); }      // k in Montgomery form
    __device__ __forceinline__ u64 acc_word(Acc a) const { return mont_redc(a.y, a.x, q, qinv_neg); }
    __device__ __forceinline__ T from_word(u64 x) const { return x; }          // normalised input word
    __device__ __forceinline__ T from_raw(u64 x) const { return x; }           // inter-pass / shared-memory word
    __device__ __forceinline__ u64 to_raw(T x) const { return x; }
    __device__ __forceinline__ TW ldtw(const TW* p) const { return __ldg(p); }
    template <int CNT>
    __device__ __forceinline__ void ldtwn(const TW* p, TW* w) const {
#pragma unroll
        for (int m = 0; m < CNT; ++m) w[m] = __ldg(p + m);
    }
    __device__ __forceinline__ void ct(T& a, T& b, TW w) const {
        u64 x = a >= q2 ? a - q2 : a;                 // [0,2q)
        u64 v = shoup_lazy(b, w.x, w.y, q);           // [0,2q)
        a = x + v;                                    // [0,4q)
        b = x - v + q2;                               // (0,4q)
    }
    __device__ __forceinline__ void gs(T& a, T& b, TW w) const {
        u64 s = a + b;                                // [0,4q)
        u64 d = a - b + q2;                           // (0,4q)
        a = s >= q2 ? s - q2 : s;                     // [0,2q)
        b = shoup_lazy(d, w.x, w.y, q);               // [0,2q)
    }
    __device__ __forceinline__ void recentre(T&) const {}
    __device__ __forceinline__ u64 final_fwd(T x) const { x = x >= q2 ? x - q2 : x; return x >= q ? x - q : x; }
    __device__ __forceinline__ u64 final_inv(T x, ulonglong2 ni) const { return shoup_mul(x, ni.x, ni.y, q); }
};

struct ArithF64 {
    typedef double T;
    typedef double TW;
    typedef double Acc;                                  // fused inner product: sum of centred residues (|sum| < beta * 0.7q)
    double q, qinv, rinv;
    __device__ __forceinline__ explicit ArithF64(const ModConst& m) : q(m.qd), qinv(m.qinv_d), rinv(m.rinv_d) {}
    static constexpr int IP_MIN_CTAS = 4;
    typedef double D;
    __device__ __forceinline__ D mac_operand(T d) const { return d; }
    __device__ __forceinline__ Acc acc_zero() const { return 0.0; }
    __device__ __forceinline__ T keep(T x) const { recentre(x); return x; }
    __device__ __forceinline__ u64 kept_word(T x) const { return to_word(x); }
    // k is a Montgomery-form word below q < 2^46: exact as a double, so d*k mod q is the same residue the integer path
    // accumulates; the factor 2^-64 is taken out once at the end (acc_word)
    __device__ __forceinline__ void mac(Acc& a, T d, u64 k) const { a = __dadd_rn(a, mulmod(d, from_word(k))); }
    __device__ __forceinline__ u64 acc_word(Acc a) const { recentre(a); return to_word(mulmod(a, rinv)); }
    __device__ __forceinline__ T from_word(u64 x) const { return __dsub_rn(__longlong_as_double(x | F64_EXP52), 4503599627370496.0); }
    __device__ __forceinline__ T from_raw(u64 x) const { return __longlong_as_double(x); }
    __device__ __forceinline__ u64 to_raw(T x) const { return (u64)__double_as_longlong(x); }
    __device__ __forceinline__ TW ldtw(const TW* p) const { return __ldg(p); }
    // CNT consecutive twiddles starting at a CNT-aligned index: 256-bit / 128-bit loads, so a lane fetches whole sectors
    template <int CNT>
    __device__ __forceinline__ void ldtwn(const TW* p, TW* w) const {
        if (CNT % 4 == 0) {
#pragma unroll
            for (int m = 0; m < CNT; m += 4)
                asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(w[m]), "=d"(w[m + 1]), "=d"(w[m + 2]), "=d"(w[m + 3]) : "l"(p + m));
        } else if (CNT % 2 == 0) {
#pragma unroll
            for (int m = 0; m < CNT; m += 2) { const double2 t = __ldg((const double2*)(p + m)); w[m] = t.x; w[m + 1] = t.y; }
        } else {
#pragma unroll
            for (int m = 0; m < CNT; ++m) w[m] = __ldg(p + m);
        }
    }
    __device__ __forceinline__ T mulmod(T x, T w) const {                     // centred residue of x*w, |r| <= 0.7q
        const double h = __dmul_rn(x, w), l = __fma_rn(x, w, -h);
        const double t = __dsub_rn(__fma_rn(h, qinv, F64_MAGIC), F64_MAGIC);
        return __dadd_rn(__fma_rn(-t, q, h), l);
    }
    __device__ __forceinline__ void ct(T& a, T& b, TW w) const {
        const double r = mulmod(b, w);
        b = __dsub_rn(a, r); a = __dadd_rn(a, r);
    }
    __device__ __forceinline__ void gs(T& a, T& b, TW w) const {
        const double s = __dadd_rn(a, b), d = __dsub_rn(a, b);
        a = s; b = mulmod(d, w);
    }
    __device__ __forceinline__ void recentre(T& x) const {
        const double t = __dsub_rn(__fma_rn(x, qinv, F64_MAGIC), F64_MAGIC);
        x = __fma_rn(-t, q, x);
    }
    __device__ __forceinline__ u64 to_word(T x) const {                        // x integer in (-q, q) -> [0,q)
        x = x < 0.0 ? __dadd_rn(x, q) : x;
        return (u64)__double_as_longlong(__dadd_rn(x, 4503599627370496.0)) & 0x000FFFFFFFFFFFFFull;
    }
    __device__ __forceinline__ u64 final_fwd(T x) const { recentre(x); return to_word(x); }
    __device__ __forceinline__ u64 final_inv(T x, ulonglong2 ni) const { return to_word(mulmod(x, from_word(ni.x))); }
};

// ArithF64 butterflies with the inner product's multiply-accumulate on the INTEGER pipe (ORION_B200_IP_IMAC): the fused kernel is
// FP64-pipe-bound (67 % busy) and a third of its FP64 work is the 8-operation modular MAC; here the transformed digit (an exact
// centred integer, |d| <= q/2 after keep(), [0,q) for own-digit words) is shifted to (0, 2q) once per digit and multiplied
// into a 96-bit lazy accumulator with three IMAD.WIDE (d < 2^47, Montgomery key word < 2^46: 8 digits stay below 2^96), one
// REDC at the end -- the same word the FP64 accumulation produces.
struct ArithF64I : ArithF64 {
    struct Acc { u64 lo; uint32_t hi; };
    typedef u64 D;
    u64 qi, qinv_neg;
    __device__ __forceinline__ explicit ArithF64I(const ModConst& m) : ArithF64(m), qi(m.q), qinv_neg(m.qinv_neg) {}
    __device__ __forceinline__ Acc acc_zero() const { return Acc{0ull, 0u}; }
    __device__ __forceinline__ D mac_operand(T d) const {
        const long long di = __double_as_longlong(__dadd_rn(d, F64_MAGIC)) - __double_as_longlong(F64_MAGIC);   // exact for |d| < 2^51
        return (u64)(di + (long long)qi);
    }
    __device__ __forceinline__ void mac(Acc& a, D du, u64 k) const {
        const uint32_t d0 = (uint32_t)du, d1 = (uint32_t)(du >> 32), k0 = (uint32_t)k, k1 = (uint32_t)(k >> 32);
        const u64 p00 = (u64)d0 * k0;
        const u64 mid = (u64)d0 * k1 + (u64)d1 * k0;                    // < 2^48
        const u64 add = p00 + (mid << 32);
        a.lo += add;
        a.hi += (uint32_t)(mid >> 32) + d1 * k1 + (uint32_t)(add < p00) + (uint32_t)(a.lo < add);
    }
    __device__ __forceinline__ u64 acc_word(Acc a) const { return mont_redc((u64)a.hi, a.lo, qi, qinv_neg); }
};

// One butterfly stage of a radix-2^NB round on EPT>>NB independent groups held in v[] (compile-time stage index I).
// MERGED: the groups' twiddle indices are consecutive (rounds on the lowest index bits), so the stage's twiddles of ALL
// groups are fetched as one aligned run (256-bit loads on the FP64 path).
template <int NB, bool MERGED, int I, int EPT, class A>
__device__ __forceinline__ void fwd_stage(const A& ar, typename A::T (&v)[EPT], const typename A::TW* __restrict__ tw, const int (&twb)[EPT]) {
    if constexpr (I < NB) {
        constexpr int G = EPT >> NB, R = 1 << NB, half = 1 << (NB - 1 - I), cnt = 1 << I;   // cnt twiddles per group
        typename A::TW w[G * cnt];
        if constexpr (MERGED) ar.template ldtwn<G * cnt>(&tw[twb[0] << I], w);
        else {
#pragma unroll
            for (int g = 0; g < G; ++g) ar.template ldtwn<cnt>(&tw[twb[g] << I], w + g * cnt);
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int m = 0; m < cnt; ++m) {
#pragma unroll
                for (int k0 = 0; k0 < half; ++k0) {
                    const int a = g * R + m * 2 * half + k0;
                    ar.ct(v[a], v[a + half], w[g * cnt + m]);
                }
            }
        }
        fwd_stage<NB, MERGED, I + 1>(ar, v, tw, twb);
    }
}
template <int NB, bool MERGED, int EPT, class A>
__device__ __forceinline__ void fwd_radix(const A& ar, typename A::T (&v)[EPT], const typename A::TW* __restrict__ tw, const int (&twb)[EPT]) {
    fwd_stage<NB, MERGED, 0>(ar, v, tw, twb);
}
// inverse (bits ascending); twb[] is the twiddle index at the TOP bit of the round
template <int NB, bool MERGED, int I, int EPT, class A>
__device__ __forceinline__ void inv_stage(const A& ar, typename A::T (&v)[EPT], const typename A::TW* __restrict__ tw, const int (&twb)[EPT]) {
    if constexpr (I < NB) {
        constexpr int G = EPT >> NB, R = 1 << NB, half = 1 << I, cnt = R >> (I + 1), sh = NB - 1 - I;
        typename A::TW w[G * cnt];
        if constexpr (MERGED) ar.template ldtwn<G * cnt>(&tw[twb[0] << sh], w);
        else {
#pragma unroll
            for (int g = 0; g < G; ++g) ar.template ldtwn<cnt>(&tw[twb[g] << sh], w + g * cnt);
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int m = 0; m < cnt; ++m) {
#pragma unroll
                for (int k0 = 0; k0 < half; ++k0) {
                    const int a = g * R + m * 2 * half + k0;
                    ar.gs(v[a], v[a + half], w[g * cnt + m]);
                }
            }
        }
        inv_stage<NB, MERGED, I + 1>(ar, v, tw, twb);
    }
}
template <int NB, bool RECENTRE, bool MERGED, int EPT, class A>
__device__ __forceinline__ void inv_radix(const A& ar, typename A::T (&v)[EPT], const typename A::TW* __restrict__ tw, const int (&twb)[EPT]) {
    constexpr int G = EPT >> NB, R = 1 << NB;
    inv_stage<NB, MERGED, 0>(ar, v, tw, twb);
    if (RECENTRE) {      // only the sum outputs of the round's last stage can have grown (the others are fresh products)
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int k0 = 0; k0 < R / 2; ++k0) ar.recentre(v[g * R + k0]);
    }
}

// One pass.  STRIDED: tile = 2^nbits rows x C cols, transformed bits are local bits [logC, LOG_E).
//            contiguous: tile = 2^LOG_E consecutive words, transformed bits are local bits [0, nbits).
// FWD walks the transformed bits downwards, INV upwards.  Every transform is exactly two passes: the first one
// (!FINAL) reads normalised words and leaves policy-specific lazy values, the second one (FINAL) normalises (and
// scales by N^-1 for INV).
// KEEP (contiguous forward final pass only): the results stay in registers -- thread tid returns words [8 tid, 8 tid + 8)
// of the tile as policy values (A::keep) instead of storing them.
// IO (cluster kernel, ntt_cluster below): 0 = global memory at both ends; 1 = the pass's output goes to the cluster's
// distributed shared-memory staging buffer (`stage` = this CTA's 2048-word slice; word g of the limb lives in the slice of
// the CTA that transforms it next); 2 = the pass's input comes from this CTA's slice.  tile_ / tid_ replace blockIdx.x /
// threadIdx.x when a CTA walks several tiles or runs two tiles side by side.
// LD (first pass of a fused base conversion + NTT, ntt1_bc_kernel): the pass's input values are not loaded but COMPUTED by
// ld(word index in the limb) -- the converted coefficient, already in the policy's lazy form.
// RET (last pass of an inverse transform, intt2_prep_kernel): the normalised output words are not stored but returned in
// retw[] with their word indices in reti[].
struct NoLd {};
template <class A, int LOG_E, int NBITS, bool STRIDED, bool FWD, bool FINAL, bool KEEP = false, int IO = 0, class LD = NoLd, bool RET = false>
__device__ __forceinline__ void ntt_body(const A ar, u64* sm, const u64* __restrict__ src, u64* __restrict__ dst,
                                         const typename A::TW* __restrict__ tw, const ulonglong2 ni, int logN,
                                         typename A::T* kept = nullptr, const bool raw_in = false,
                                         u64* stage = nullptr, const int tile_ = -1, const int tid_ = -1,
                                         const LD* ld = nullptr, u64* retw = nullptr, int* reti = nullptr) {
    typedef typename A::T T;
    constexpr int LOG_EPT = log_ept(STRIDED, FWD), EPT = 1 << LOG_EPT;
    constexpr int nbits = NBITS;          // compile-time: the round schedule below unrolls into straight-line code
    const int tid = tid_ < 0 ? (int)threadIdx.x : tid_;
    const int tile = tile_ < 0 ? (int)blockIdx.x : tile_;
    constexpr int logC = STRIDED ? LOG_E - nbits : 0;
    constexpr int lo_bit = STRIDED ? logC : 0;           // lowest transformed local bit
    const int TOP = STRIDED ? LOG_E - 1 : logN - 1;      // stage = TOP - local bit
    const int tile_base = STRIDED ? 0 : tile << LOG_E;
    auto gidx = [&](int li) -> int {
        if (STRIDED) return ((li >> logC) << (logN - nbits)) + (tile << logC) + (li & ((1 << logC) - 1));
        return tile_base + li;
    };
    // cluster staging (IO != 0): the strided tiles are 2048 words with 2^(19 - logN) columns; word g of the limb is staged
    // for the contiguous pass in slice g >> 11 at g & 2047, for the strided pass in slice (g & 255) >> logCs at
    // ((g >> 8) << logCs) | (g & (Cs - 1))
    auto stage_ptr = [&](int g) -> u64* {
        const int logCs = 19 - logN;
        const int rank = FWD ? (g >> 11) : ((g & 255) >> logCs);
        const int loc = FWD ? (g & 2047) : (((g >> 8) << logCs) | (g & ((1 << logCs) - 1)));
        return cooperative_groups::this_cluster().map_shared_rank(stage, rank) + loc;
    };

    constexpr int NR = (NBITS + LOG_EPT - 1) / LOG_EPT, RBASE = NBITS / NR, REXTRA = NBITS % NR;
    const int row_stride = 1 << (logN - nbits);          // STRIDED: distance between consecutive tile rows in the limb
    int done = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const int nb = RBASE + (r < REXTRA ? 1 : 0);
        const int G = EPT >> nb;
        const int pos = FWD ? lo_bit + nbits - done - nb : lo_bit + done;   // lowest bit of this round
        const int ptop = pos + nb - 1;
        // which register pairs sit in adjacent words (-> 128-bit global accesses): with pos == 0 elements e, e+1;
        // with pos > 0 and G >= 2 the same k of groups g, g+1 (bit 0 of the local index is bit 0 of g)
        const int pair_step = pos == 0 ? 1 : (G >= 2 ? (1 << nb) : 0);
        // per group g: local index with the round's k bits zero, its shared-memory slot and its word index in the limb;
        // element (g, k) adds compile-time offsets to both
        T v[EPT]; int sbase[EPT], gbase[EPT], twb[EPT], lbase[EPT];
#pragma unroll
        for (int g = 0; g < EPT; ++g) {
            if (g < G) {
                const int base = insert_bits(tid * G + g, pos, nb, 0);
                sbase[g] = pad<LOG_EPT>(base);
                gbase[g] = gidx(base);
                lbase[g] = base + (STRIDED ? 0 : ((tile & 1) << LOG_E));      // IO == 2: word index in this CTA's slice
            }
            twb[g] = (1 << (TOP - ptop)) + ((tid * G + g) >> pos) + (tile_base >> (ptop + 1));
        }
        auto soff = [&](int e) { return sbase[e >> nb] + pad<LOG_EPT>((e & ((1 << nb) - 1)) << pos); };
        auto goff = [&](int e) {
            const int k = e & ((1 << nb) - 1);
            return gbase[e >> nb] + (STRIDED ? (k << (pos - logC)) * row_stride : (k << pos));
        };
        const bool quad = pos == 0 && nb >= 2;            // a thread owns runs of >= 4 adjacent words: 256-bit accesses
        if (r == 0 && !std::is_same<LD, NoLd>::value) {
            if constexpr (!std::is_same<LD, NoLd>::value) {
#pragma unroll
                for (int e = 0; e < EPT; ++e) v[e] = (*ld)(goff(e));
            }
        } else if (r == 0) {
            u64 w[EPT];
            if (IO == 2) {
#pragma unroll
                for (int e = 0; e < EPT; ++e) w[e] = stage[lbase[e >> nb] + ((e & ((1 << nb) - 1)) << pos)];
            } else if (quad) {
#pragma unroll
                for (int e = 0; e < EPT; e += 4) ld_global_256(src + goff(e), w[e], w[e + 1], w[e + 2], w[e + 3]);
            } else if (pair_step) {
#pragma unroll
                for (int e = 0; e < EPT; ++e)
                    if (!(e & pair_step)) {
                        const ulonglong2 t = *(const ulonglong2*)(src + goff(e));
                        w[e] = t.x; w[e + pair_step] = t.y;
                    }
            } else {
#pragma unroll
                for (int e = 0; e < EPT; ++e) w[e] = src[goff(e)];
            }
#pragma unroll
            for (int e = 0; e < EPT; ++e) v[e] = (FINAL || raw_in) ? ar.from_raw(w[e]) : ar.from_word(w[e]);
        } else {
#pragma unroll
            for (int e = 0; e < EPT; ++e) v[e] = ar.from_raw(sm[soff(e)]);
        }
        const bool merged = pos == 0;                     // consecutive groups -> consecutive twiddle indices
#define OB_RADIX(FN, ...) do { \
        if (nb == 4) { if (merged) FN<(LOG_EPT >= 4 ? 4 : 3), __VA_ARGS__ true, EPT>(ar, v, tw, twb); else FN<(LOG_EPT >= 4 ? 4 : 3), __VA_ARGS__ false, EPT>(ar, v, tw, twb); } \
        else if (nb == 3) { if (merged) FN<3, __VA_ARGS__ true, EPT>(ar, v, tw, twb); else FN<3, __VA_ARGS__ false, EPT>(ar, v, tw, twb); } \
        else if (nb == 2) { if (merged) FN<2, __VA_ARGS__ true, EPT>(ar, v, tw, twb); else FN<2, __VA_ARGS__ false, EPT>(ar, v, tw, twb); } \
        else { if (merged) FN<1, __VA_ARGS__ true, EPT>(ar, v, tw, twb); else FN<1, __VA_ARGS__ false, EPT>(ar, v, tw, twb); } } while (0)
        if (FWD) OB_RADIX(fwd_radix, );
        else if (FINAL && r == NR - 1) OB_RADIX(inv_radix, false, );
        else OB_RADIX(inv_radix, true, );
#undef OB_RADIX
        done += nb;
        if (KEEP && r == NR - 1) {
            // last round of the contiguous pass: pos = 0, groups of 2^nb adjacent words, group g of thread tid at
            // (tid * G + g) << nb  ->  element e sits at word tid * EPT + e
#pragma unroll
            for (int e = 0; e < EPT; ++e) kept[e] = ar.keep(v[e]);
        } else if (r == NR - 1) {
            u64 w[EPT];
#pragma unroll
            for (int e = 0; e < EPT; ++e) w[e] = !FINAL ? ar.to_raw(v[e]) : (FWD ? ar.final_fwd(v[e]) : ar.final_inv(v[e], ni));
            if (RET) {
#pragma unroll
                for (int e = 0; e < EPT; ++e) { retw[e] = w[e]; reti[e] = goff(e); }
            } else if (IO == 1) {
                if (pair_step) {
#pragma unroll
                    for (int e = 0; e < EPT; ++e)
                        if (!(e & pair_step)) *(ulonglong2*)stage_ptr(goff(e)) = make_ulonglong2(w[e], w[e + pair_step]);
                } else {
#pragma unroll
                    for (int e = 0; e < EPT; ++e) *stage_ptr(goff(e)) = w[e];
                }
            } else if (quad) {
#pragma unroll
                for (int e = 0; e < EPT; e += 4) st_global_256(dst + goff(e), w[e], w[e + 1], w[e + 2], w[e + 3]);
            } else if (pair_step) {
#pragma unroll
                for (int e = 0; e < EPT; ++e)
                    if (!(e & pair_step)) *(ulonglong2*)(dst + goff(e)) = make_ulonglong2(w[e], w[e + pair_step]);
            } else {
#pragma unroll
                for (int e = 0; e < EPT; ++e) dst[goff(e)] = w[e];
            }
        } else {
#pragma unroll
            for (int e = 0; e < EPT; ++e) sm[soff(e)] = ar.to_raw(v[e]);
            __syncthreads();
        }
    }
}

template <int LOG_E, int NBITS, bool STRIDED, bool FWD, bool FINAL>
__global__ void __launch_bounds__(1 << (LOG_E - log_ept(STRIDED, FWD)))
ntt_pass(const u64* in, u64* out, const LimbDesc* __restrict__ limbs,
         const ModConst* __restrict__ mc, const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all,
         const ulonglong2* __restrict__ ninv_all, int logN) {
    OB_PDL_LAUNCH();
    __shared__ __align__(16) u64 sm[(1 << LOG_E) + (1 << (LOG_E - 3))];         // padded layout, see pad()
    const LimbDesc ld = limbs[blockIdx.y];                                       // host-built tables: read before the wait
    const size_t N = (size_t)1 << logN;
    const ModConst& m = mc[ld.mod];
    const u64* src = in + (size_t)ld.in_off * N;
    u64* dst = out + (size_t)ld.out_off * N;
    const ulonglong2 ni = ninv_all[ld.mod];
    // aux bit 0 (forward first pass only): the limb already holds the FP64 policy's lazy values (kernels.h bc_raw_out)
    if (m.bits <= F64_MAX_BITS) {
        const ArithF64 ar(m);
        OB_PDL_WAIT();
        ntt_body<ArithF64, LOG_E, NBITS, STRIDED, FWD, FINAL>(ar, sm, src, dst, twd_all + (size_t)ld.mod * N, ni, logN,
                                                              nullptr, FWD && !FINAL && (ld.aux & 1));
    } else {
        const ArithInt ar(m);
        OB_PDL_WAIT();
        ntt_body<ArithInt, LOG_E, NBITS, STRIDED, FWD, FINAL>(ar, sm, src, dst, tw_all + (size_t)ld.mod * N, ni, logN);
    }
}

// Fused second NTT pass + key-switch inner product (evaluator.cu step 3b + 4).  A CTA owns 1024 consecutive coefficients
// of one QP limb t of ciphertext z and walks the beta digits: the digit's tile (left in lazy form by the strided pass)
// goes through the 8 contiguous stages in shared memory, stays in registers, and is multiplied into the two
// accumulators against the evaluation key.  The transformed digits are never written: 2 x beta*W limb moves per key
// switch less than NTT + inner_product_multi_kernel, and the key stream (the compulsory traffic) overlaps the butterflies.
// Own-digit limbs (t inside digit j) come straight from c1.  Two shared-memory tiles alternate so that a digit's first
// exchange cannot overwrite words a slower warp still reads for the previous digit.
struct LimbTab { unsigned char t[128]; };
// STAGES > 0: the evaluation-key words go through a per-thread shared-memory ring filled with cp.async (LDGSTS.128, L1
// bypass) STAGES-1 digits ahead of their use.  Every thread copies exactly the 2 x 64 bytes it will multiply itself, so the
// ring needs no barrier (cp.async.wait_group orders a thread's own copies), and its 64-byte slots are skewed to an 80-byte
// stride so the four LDS.128 of a quarter-warp fall into distinct bank groups.  The compulsory stream of the kernel (2/3 of
// its bytes) is then in flight in shared memory instead of being tied to registers right before the multiply.
constexpr int IP_SLOT_BYTES = 80;
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }
__device__ __forceinline__ void lds128(const void* p, u64& a, u64& b) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(s));
}
template <class A, int STAGES>
__global__ void __launch_bounds__(1 << (LOG_E_CONTIG - 3), A::IP_MIN_CTAS)
ntt2_ip_kernel(const u64* __restrict__ c1, size_t c1_zstride, const u64* __restrict__ digits, size_t dig_zstride,
               const IpMulti m, u64* __restrict__ out, size_t out_zstride, const __grid_constant__ LimbTab tab,
               int L, int K, int nQ, int nmod, int beta, const ModConst* __restrict__ mc,
               const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all, int logN, int nzx) {
    OB_PDL_SYNC();
    constexpr int EPT = 8, SM_WORDS = (1 << LOG_E_CONTIG) + (1 << (LOG_E_CONTIG - 3)), TPB_ = 1 << (LOG_E_CONTIG - 3);
    // nzx > 1: the ciphertext index is the FASTEST-varying part of blockIdx.x, so the nzx CTAs that work on the same tile of the
    // same limb -- and, in a batch that shares one evaluation key, read the same 2 beta key tiles -- are resident together and the
    // key comes from DRAM once per group instead of once per ciphertext (the rest are L2 hits)
    const int bx = blockIdx.x / nzx;
    extern __shared__ __align__(16) unsigned char ip_smem[];
    u64 (*sm)[SM_WORDS] = (u64 (*)[SM_WORDS])ip_smem;                        // [2][SM_WORDS] exchange tiles
    unsigned char* ring = ip_smem + 2 * SM_WORDS * 8;                        // [STAGES][2][TPB_] slots of IP_SLOT_BYTES
    typedef typename A::T T;
    const int t = tab.t[blockIdx.y], z = blockIdx.z + blockIdx.x % nzx, W = L + K;
    const size_t N = (size_t)1 << logN;
    const int ci = t < L ? t : nQ + (t - L);
    const A ar(mc[ci]);
    const int own = t < L ? t / K : -1;
    const typename A::TW* tw;
    if constexpr (sizeof(typename A::TW) == 8) tw = twd_all + (size_t)ci * N; else tw = tw_all + (size_t)ci * N;
    const size_t off = ((size_t)bx << LOG_E_CONTIG) + threadIdx.x * EPT;      // this thread's 8 words in the limb
    const u64* __restrict__ evk = m.evk[z] + (size_t)ci * N + off;
    const u64* __restrict__ c1t = c1 + (size_t)z * c1_zstride + (size_t)t * N + off;
    const u64* __restrict__ dz = digits + (size_t)z * dig_zstride + (size_t)t * N;
    typename A::Acc a0[EPT], a1[EPT];
#pragma unroll
    for (int e = 0; e < EPT; ++e) { a0[e] = ar.acc_zero(); a1[e] = ar.acc_zero(); }
    int buf = 0;
    // L2 prefetch, one 128-byte line per thread: the key tiles of digit j (2 x 8 KB; direct-load variant only) and the tile
    // of digit j+1 (8 KB), so that the loads that follow the butterflies find their lines in L2
    const size_t line = ((size_t)bx << LOG_E_CONTIG) + (threadIdx.x & 63) * 16;
    const u64* pf_evk = m.evk[z] + ((size_t)(threadIdx.x >> 6) * nmod + ci) * N + line;
    const u64* pf_dig = dz + line;
    auto stage_keys = [&](int j) {               // copies of digit j's key words into ring slot j % STAGES (empty group past the end)
        if (STAGES > 0) {
            if (j < beta) {
                const u64* k0 = evk + (size_t)j * 2 * nmod * N;
                const u64* k1 = k0 + (size_t)nmod * N;
                unsigned char* s0 = ring + ((size_t)((j % STAGES) * 2 + 0) * TPB_ + threadIdx.x) * IP_SLOT_BYTES;
                unsigned char* s1 = s0 + (size_t)TPB_ * IP_SLOT_BYTES;
#pragma unroll
                for (int c = 0; c < 4; ++c) { cp_async16(s0 + 16 * c, k0 + 2 * c); cp_async16(s1 + 16 * c, k1 + 2 * c); }
            }
            cp_async_commit();
        }
    };
    if (STAGES > 0) {
#pragma unroll
        for (int j = 0; j < STAGES - 1; ++j) stage_keys(j);
    }
    for (int j = 0; j < beta; ++j) {
        if (STAGES > 0) stage_keys(j + STAGES - 1);
        else asm volatile("prefetch.global.L2 [%0];" ::"l"(pf_evk + (size_t)j * 2 * nmod * N));
        if (threadIdx.x < 64 && j + 1 < beta && j + 1 != own)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pf_dig + (size_t)(j + 1) * W * N));
        T d[EPT];
        if (j == own) {
            u64 w[EPT];
            ld_global_256(c1t, w[0], w[1], w[2], w[3]); ld_global_256(c1t + 4, w[4], w[5], w[6], w[7]);
#pragma unroll
            for (int e = 0; e < EPT; ++e) d[e] = ar.from_word(w[e]);
        } else {
            ntt_body<A, LOG_E_CONTIG, K2_BITS, false, true, true, true>(ar, sm[buf], dz + (size_t)j * W * N, nullptr, tw,
                                                                         make_ulonglong2(0, 0), logN, d, false, nullptr, bx);
            buf ^= 1;
        }
        typename A::D dd[EPT];
#pragma unroll
        for (int e = 0; e < EPT; ++e) dd[e] = ar.mac_operand(d[e]);
        u64 k[EPT];
        if (STAGES > 0) {
            cp_async_wait<(STAGES > 0 ? STAGES - 1 : 0)>();
            const unsigned char* s0 = ring + ((size_t)((j % STAGES) * 2 + 0) * TPB_ + threadIdx.x) * IP_SLOT_BYTES;
            const unsigned char* s1 = s0 + (size_t)TPB_ * IP_SLOT_BYTES;
#pragma unroll
            for (int c = 0; c < 4; ++c) lds128(s0 + 16 * c, k[2 * c], k[2 * c + 1]);
#pragma unroll
            for (int e = 0; e < EPT; ++e) ar.mac(a0[e], dd[e], k[e]);
#pragma unroll
            for (int c = 0; c < 4; ++c) lds128(s1 + 16 * c, k[2 * c], k[2 * c + 1]);
#pragma unroll
            for (int e = 0; e < EPT; ++e) ar.mac(a1[e], dd[e], k[e]);
        } else {
            const u64* k0 = evk + (size_t)j * 2 * nmod * N;
            const u64* k1 = k0 + (size_t)nmod * N;
            ld_global_256(k0, k[0], k[1], k[2], k[3]); ld_global_256(k0 + 4, k[4], k[5], k[6], k[7]);
#pragma unroll
            for (int e = 0; e < EPT; ++e) ar.mac(a0[e], dd[e], k[e]);
            ld_global_256(k1, k[0], k[1], k[2], k[3]); ld_global_256(k1 + 4, k[4], k[5], k[6], k[7]);
#pragma unroll
            for (int e = 0; e < EPT; ++e) ar.mac(a1[e], dd[e], k[e]);
        }
    }
    u64* o0 = out + (size_t)z * out_zstride + (size_t)t * N + off;
    u64* o1 = o0 + (size_t)W * N;
    u64 r[EPT];
#pragma unroll
    for (int e = 0; e < EPT; ++e) r[e] = ar.acc_word(a0[e]);
    st_global_256(o0, r[0], r[1], r[2], r[3]); st_global_256(o0 + 4, r[4], r[5], r[6], r[7]);
#pragma unroll
    for (int e = 0; e < EPT; ++e) r[e] = ar.acc_word(a1[e]);
    st_global_256(o1, r[0], r[1], r[2], r[3]); st_global_256(o1 + 4, r[4], r[5], r[6], r[7]);
}

// Fused second NTT pass + "subtract and scale" epilogue shared by ModDown and rescale:
//   out[p][i] = (a[p][i] - NTT(b[p][i])) * mulc[i]  (+ addend)      mod q_i
// b holds the output of ntt_forward_first; the transformed limb is never written.
template <class A>
__device__ __forceinline__ void ntt2_submul_body(const A ar, u64* sm, const u64* __restrict__ a, const u64* __restrict__ b,
                                                 const u64* __restrict__ addend, u64* __restrict__ out,
                                                 const typename A::TW* __restrict__ tw, const ulonglong2 w, const u64 q, int logN) {
    constexpr int EPT = 8;
    typename A::T d[EPT];
    ntt_body<A, LOG_E_CONTIG, K2_BITS, false, true, true, true>(ar, sm, b, nullptr, tw, make_ulonglong2(0, 0), logN, d);
    const size_t off = ((size_t)blockIdx.x << LOG_E_CONTIG) + threadIdx.x * EPT;
    u64 x[EPT], r[EPT];
    ld_global_256(a + off, x[0], x[1], x[2], x[3]); ld_global_256(a + off + 4, x[4], x[5], x[6], x[7]);
#pragma unroll
    for (int e = 0; e < EPT; ++e) r[e] = shoup_mul(sub_mod(x[e], ar.kept_word(d[e]), q), w.x, w.y, q);
    if (addend != nullptr) {
        ld_global_256(addend + off, x[0], x[1], x[2], x[3]); ld_global_256(addend + off + 4, x[4], x[5], x[6], x[7]);
#pragma unroll
        for (int e = 0; e < EPT; ++e) r[e] = add_mod(r[e], x[e], q);
    }
    st_global_256(out + off, r[0], r[1], r[2], r[3]); st_global_256(out + off + 4, r[4], r[5], r[6], r[7]);
}

__global__ void __launch_bounds__(1 << (LOG_E_CONTIG - 3))
ntt2_submul_kernel(const u64* __restrict__ a, int a_stride, const u64* __restrict__ b, int nl,
                   const u64* __restrict__ addend, int addend_stride, int addend_polys, int addend_period, int addend_zstride,
                   u64* __restrict__ out, int out_stride, const ulonglong2* __restrict__ mulc, const ModConst* __restrict__ mc,
                   const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all, int logN) {
    OB_PDL_LAUNCH();
    __shared__ __align__(16) u64 sm[(1 << LOG_E_CONTIG) + (1 << (LOG_E_CONTIG - 3))];
    const int i = blockIdx.y, p = blockIdx.z;
    const size_t N = (size_t)1 << logN;
    const ModConst& m = mc[i];
    const ulonglong2 mw = mulc[i];
    const u64 mq = m.q;
    // polynomial p = z*period + c receives addend polynomial c of group z when c < addend_polys
    const int pz = p / addend_period, pc = p - pz * addend_period;
    const u64* ad = (addend != nullptr && pc < addend_polys)
                        ? addend + ((size_t)pz * addend_zstride + (size_t)pc * addend_stride + i) * N : nullptr;
    const u64* ap = a + ((size_t)p * a_stride + i) * N;
    const u64* bp = b + ((size_t)p * nl + i) * N;
    u64* op = out + ((size_t)p * out_stride + i) * N;
    if (m.bits <= F64_MAX_BITS) {
        const ArithF64 ar(m);
        OB_PDL_WAIT();
        ntt2_submul_body<ArithF64>(ar, sm, ap, bp, ad, op, twd_all + (size_t)i * N, mw, mq, logN);
    } else {
        const ArithInt ar(m);
        OB_PDL_WAIT();
        ntt2_submul_body<ArithInt>(ar, sm, ap, bp, ad, op, tw_all + (size_t)i * N, mw, mq, logN);
    }
}

// ---- base conversion fused into the transforms either side of it (key switch: ModUp and ModDown) -------------------------
// Un-fused, a base conversion is a round trip of its own: INTT (2 passes) -> conversion kernel (reads ns limbs, writes nd
// limbs) -> first NTT pass (reads and writes nd limbs).  Fused:
//   intt2_prep_kernel : the LAST pass of the inverse transform, one CTA per (tile, conversion) walking the conversion's ns
//                       source limbs: it leaves y_i = x_i * (Q_S/q_i)^-1 mod q_i in place of the coefficient x_i (as an exact
//                       double when every source prime is an FP64-path prime, else as a word) and the overflow count
//                       v = floor(sum_i y_i / q_i) -- Lattigo's float64 estimate, same operations in the same order as
//                       baseconv_batch_kernel -- as one byte per coefficient.
//   ntt1_bc_kernel    : the FIRST pass of the forward transform of every destination limb computes its inputs from y and v
//                       (ns multiply-accumulates + the -v Q_S correction) instead of loading a converted limb.
// The converted limbs are then written once (lazy inter-pass form) and never read back before the second pass; the y limbs
// (ns per conversion, shared by all its destinations) are served from L2.  Same words as the un-fused path: every value
// is the same exact residue, only the place where it is computed moves.  ns <= BCF_MAX_NS (raw 128-bit accumulation is
// valid for any four primes below 2^62; more sources fall back to the un-fused path).
constexpr int BCF_MAX_NS = 4;
struct BcLimb { uint32_t out_off, mod, conv, d; };      // destination limb: where it is written, its modulus, its conversion and index in it

template <int S1, bool CENTERED>
__global__ void __launch_bounds__(1 << (LOG_E_STRIDED - log_ept(true, false)))
intt2_prep_kernel(u64* buf, unsigned char* __restrict__ vout, const BcDev* __restrict__ descs, const ModConst* __restrict__ mc,
                  const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all, const ulonglong2* __restrict__ ninv_all, int logN) {
    OB_PDL_SYNC();
    constexpr int EPT = 1 << log_ept(true, false);
    __shared__ __align__(16) u64 sm[(1 << LOG_E_STRIDED) + (1 << (LOG_E_STRIDED - 3))];
    const BcDev d = descs[blockIdx.y];
    const size_t N = (size_t)1 << logN;
    double vf[EPT];
    int gi[EPT];
#pragma unroll
    for (int e = 0; e < EPT; ++e) vf[e] = 0.0;
    for (int i = 0; i < d.ns; ++i) {
        const int ci = d.src_mod[i];
        const ModConst& m = mc[ci];
        u64* limb = buf + (size_t)(d.in_off + i) * N;
        const ulonglong2 ni = ninv_all[ci];
        u64 w[EPT];
        if (m.bits <= F64_MAX_BITS)
            ntt_body<ArithF64, LOG_E_STRIDED, S1, true, false, true, false, 0, NoLd, true>(ArithF64(m), sm, limb, nullptr, twd_all + (size_t)ci * N, ni, logN,
                                                                                           nullptr, false, nullptr, -1, -1, nullptr, w, gi);
        else
            ntt_body<ArithInt, LOG_E_STRIDED, S1, true, false, true, false, 0, NoLd, true>(ArithInt(m), sm, limb, nullptr, tw_all + (size_t)ci * N, ni, logN,
                                                                                           nullptr, false, nullptr, -1, -1, nullptr, w, gi);
        const ulonglong2 inv = d.inv[i];
        const u64 q = m.q, half = CENTERED ? d.half_src[i] : 0;
        const double qd = m.qd;
#pragma unroll
        for (int e = 0; e < EPT; ++e) {
            u64 x = w[e];
            if (CENTERED) x = add_mod(x, half, q);
            const u64 y = shoup_mul(x, inv.x, inv.y, q);
            const double yd = __ull2double_rn(y);
            vf[e] = __dadd_rn(vf[e], __ddiv_rn(yd, qd));
            limb[gi[e]] = d.f64_src ? (u64)__double_as_longlong(yd) : y;
        }
        __syncthreads();                                     // the next limb's first exchange reuses sm
    }
    unsigned char* vo = vout + (size_t)blockIdx.y * N;
    const bool single = d.ns == 1 && !CENTERED;              // a single source without centring: v is 0 by definition
#pragma unroll
    for (int e = 0; e < EPT; ++e) vo[gi[e]] = single ? 0 : (unsigned char)(int)vf[e];
}

// input of a destination limb's first NTT pass, computed from the conversion's y limbs and overflow bytes
template <class A, bool CENTERED>
struct BcLoader {
    const u64* y[BCF_MAX_NS];          // source limbs (y form)
    const unsigned char* v;
    const double* s_negd;              // shared memory: [ns+1] doubles, [ns+1] words
    const u64* s_neg;
    double wd[BCF_MAX_NS];             // (Q_S/q_i) mod p_t, standard form, exact double (FP64 contraction)
    u64 wm[BCF_MAX_NS];                // the same in Montgomery form (integer contraction)
    u64 q, qinv_neg, half_dst;
    int ns;
    bool f64_src, f64_mac, raw_ok;
    A ar;
    __device__ __forceinline__ typename A::T operator()(int gi) const {
        const int vb = v[gi];
        if constexpr (std::is_same<A, ArithF64>::value) {
            if (f64_mac) {                                   // sources and destination on the FP64 path
                double a = s_negd[vb];
#pragma unroll
                for (int i = 0; i < BCF_MAX_NS; ++i)
                    if (i < ns) a = __dadd_rn(a, ar.mulmod(__longlong_as_double((long long)y[i][gi]), wd[i]));
                if (!raw_ok) ar.recentre(a);
                return a;
            }
        }
        u64 hi = 0, lo = 0;
#pragma unroll
        for (int i = 0; i < BCF_MAX_NS; ++i)
            if (i < ns) {
                const u64 raw = y[i][gi];
                const u64 yi = f64_src ? (u64)__double2ull_rn(__longlong_as_double((long long)raw)) : raw;
                const u64 pl = yi * wm[i], ph = __umul64hi(yi, wm[i]);
                lo += pl; hi += ph + (lo < pl);
            }
        u64 r = add_mod(mont_redc(hi, lo, q, qinv_neg), s_neg[vb], q);
        if (CENTERED) r = sub_mod(r, half_dst, q);
        return ar.from_word(r);
    }
};

template <int S1, bool CENTERED>
__global__ void __launch_bounds__(1 << (LOG_E_STRIDED - log_ept(true, true)))
ntt1_bc_kernel(const u64* __restrict__ ybuf, const unsigned char* __restrict__ vbuf, u64* out, const BcLimb* __restrict__ limbs,
               const BcDev* __restrict__ descs, const ModConst* __restrict__ mc, const ulonglong2* __restrict__ tw_all,
               const double* __restrict__ twd_all, int logN) {
    OB_PDL_SYNC();
    __shared__ __align__(16) u64 sm[(1 << LOG_E_STRIDED) + (1 << (LOG_E_STRIDED - 3))];
    __shared__ double s_negd[BCF_MAX_NS + 1];
    __shared__ u64 s_neg[BCF_MAX_NS + 1];
    const BcLimb bl = limbs[blockIdx.y];
    const BcDev d = descs[bl.conv];
    const size_t N = (size_t)1 << logN;
    const ModConst& m = mc[bl.mod];
    if (threadIdx.x <= d.ns) {
        s_negd[threadIdx.x] = d.negvq_d[(size_t)bl.d * (d.ns + 1) + threadIdx.x];
        s_neg[threadIdx.x] = d.negvq[(size_t)bl.d * (d.ns + 1) + threadIdx.x];
    }
    __syncthreads();
    u64* dst = out + (size_t)bl.out_off * N;
    const ulonglong2 ni = make_ulonglong2(0, 0);
    auto fill = [&](auto& ld) {
#pragma unroll
        for (int i = 0; i < BCF_MAX_NS; ++i) {
            const bool on = i < d.ns;
            ld.y[i] = on ? ybuf + (size_t)(d.in_off + i) * N : nullptr;
            ld.wd[i] = on ? d.mat_d[(size_t)i * d.nd + bl.d] : 0.0;
            ld.wm[i] = on ? d.mat[(size_t)i * d.nd + bl.d] : 0;
        }
        ld.v = vbuf + (size_t)bl.conv * N;
        ld.s_negd = s_negd; ld.s_neg = s_neg;
        ld.q = m.q; ld.qinv_neg = m.qinv_neg; ld.half_dst = CENTERED ? d.half_dst[bl.d] : 0;
        ld.ns = d.ns; ld.f64_src = d.f64_src != 0;
        ld.f64_mac = !CENTERED && d.f64_src && m.bits <= F64_MAX_BITS;
        ld.raw_ok = d.ns <= 4;                               // growth bound of the forward butterflies (kernels.h bc_raw_out)
    };
    if (m.bits <= F64_MAX_BITS) {
        BcLoader<ArithF64, CENTERED> ld{{}, nullptr, nullptr, nullptr, {}, {}, 0, 0, 0, 0, false, false, false, ArithF64(m)};
        fill(ld);
        ntt_body<ArithF64, LOG_E_STRIDED, S1, true, true, false, false, 0, BcLoader<ArithF64, CENTERED>>(
            ArithF64(m), sm, nullptr, dst, twd_all + (size_t)bl.mod * N, ni, logN, nullptr, false, nullptr, -1, -1, &ld);
    } else {
        BcLoader<ArithInt, CENTERED> ld{{}, nullptr, nullptr, nullptr, {}, {}, 0, 0, 0, 0, false, false, false, ArithInt(m)};
        fill(ld);
        ntt_body<ArithInt, LOG_E_STRIDED, S1, true, true, false, false, 0, BcLoader<ArithInt, CENTERED>>(
            ArithInt(m), sm, nullptr, dst, tw_all + (size_t)bl.mod * N, ni, logN, nullptr, false, nullptr, -1, -1, &ld);
    }
}

// Rescale: the first forward pass of limb i of polynomial p with its input computed on the fly from the coefficient-domain last limb
//   x -> ((last[p][x] + (q_l - 1)/2) mod q_l) mod q_i - ((q_l - 1)/2 mod q_i)
// (what rescale_prep_kernel writes to [polys][level][N] and ntt_forward_first reads back: one launch and one round trip less per
// rescale; the last limb -- N words per polynomial -- is served from L2 to its `level` readers).  Same words.
template <class A>
struct RescaleLoader {
    const u64* last;
    u64 ql, half, q, hq;
    ModConst m;
    A ar;
    __device__ __forceinline__ typename A::T operator()(int gi) const {
        const u64 s = add_mod(last[gi], half, ql);
        return ar.from_word(sub_mod(barrett_red(0, s, m), hq, q));
    }
};
template <int S1>
__global__ void __launch_bounds__(1 << (LOG_E_STRIDED - log_ept(true, true)))
ntt1_rescale_kernel(const u64* __restrict__ last, u64* out, int level, const ModConst* __restrict__ mc,
                    const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all, int logN) {
    OB_PDL_LAUNCH();
    __shared__ __align__(16) u64 sm[(1 << LOG_E_STRIDED) + (1 << (LOG_E_STRIDED - 3))];
    const int i = blockIdx.y % level, p = blockIdx.y / level;
    const size_t N = (size_t)1 << logN;
    const ModConst m = mc[i];
    const u64 ql = mc[level].q, half = (ql - 1) >> 1;
    const u64 hq = barrett_red(0, half, m);
    u64* dst = out + (size_t)blockIdx.y * N;
    const ulonglong2 ni = make_ulonglong2(0, 0);
    if (m.bits <= F64_MAX_BITS) {
        const RescaleLoader<ArithF64> ld{last + (size_t)p * N, ql, half, m.q, hq, m, ArithF64(m)};
        OB_PDL_WAIT();
        ntt_body<ArithF64, LOG_E_STRIDED, S1, true, true, false, false, 0, RescaleLoader<ArithF64>>(
            ArithF64(m), sm, nullptr, dst, twd_all + (size_t)i * N, ni, logN, nullptr, false, nullptr, -1, -1, &ld);
    } else {
        const RescaleLoader<ArithInt> ld{last + (size_t)p * N, ql, half, m.q, hq, m, ArithInt(m)};
        OB_PDL_WAIT();
        ntt_body<ArithInt, LOG_E_STRIDED, S1, true, true, false, false, 0, RescaleLoader<ArithInt>>(
            ArithInt(m), sm, nullptr, dst, tw_all + (size_t)i * N, ni, logN, nullptr, false, nullptr, -1, -1, &ld);
    }
}

// Whole-limb NTT / INTT in ONE launch for small rings (2^12 <= N <= 2^14): a thread-block cluster of N/2048 CTAs owns one
// limb.  Forward: every CTA runs the strided pass on its 2048-word tile and scatters the lazy results through distributed
// shared memory into the slice of the CTA that owns the destination words; after one cluster barrier every CTA runs the
// contiguous pass on the two 1024-word tiles of its slice.  Inverse: the mirror image (two contiguous tiles side by side
// on 256 threads, then the strided final pass).  Same ntt_body, same words as the two-launch path -- the inter-pass limb
// never touches global memory and a transform is one link of a launch-latency-bound chain instead of two.
// MEASURED, NOT ADOPTED (ORION_B200_CLUSTER_NTT=1 enables it for transforms of at most ORION_B200_CLUSTER_NTT_MAX = 48
// limbs): bit-identical words, ResNet-20 at logN=13 57.7 k -> 48.5 k launches per image, but 0.3365 vs 0.3359 s/image and
// LoLA2 at logN=14 2.43 vs 2.31 ms/image -- the two cluster barriers and the scatter through distributed shared memory
// cost what the saved launch gains, and above ~100 limbs the two-launch path is faster (864 limbs: 54 vs 58 us).
template <int NBITS, bool FWD>
__global__ void __launch_bounds__(FWD ? 128 : 256)
ntt_cluster(const u64* in, u64* out, const LimbDesc* __restrict__ limbs,
            const ModConst* __restrict__ mc, const ulonglong2* __restrict__ tw_all, const double* __restrict__ twd_all,
            const ulonglong2* __restrict__ ninv_all, int logN, int raw_flag) {
    OB_PDL_SYNC();
    namespace cg = cooperative_groups;
    __shared__ __align__(16) u64 stage[1 << LOG_E_STRIDED];
    __shared__ __align__(16) u64 sm[(1 << LOG_E_STRIDED) + (1 << (LOG_E_STRIDED - 3))];
    cg::cluster_group cl = cg::this_cluster();
    const int rank = (int)cl.block_rank();
    const LimbDesc ld = limbs[blockIdx.y];
    const size_t N = (size_t)1 << logN;
    const ModConst& m = mc[ld.mod];
    const u64* src = in + (size_t)ld.in_off * N;
    u64* dst = out + (size_t)ld.out_off * N;
    const ulonglong2 ni = ninv_all[ld.mod];
    const bool f64 = m.bits <= F64_MAX_BITS;
    const int tid = threadIdx.x;
    constexpr int SM_C = (1 << LOG_E_CONTIG) + (1 << (LOG_E_CONTIG - 3));
    cl.sync();                                   // every CTA of the cluster is resident before its slice is written
    if (FWD) {
        const bool raw_in = raw_flag && (ld.aux & 1);
        if (f64) ntt_body<ArithF64, LOG_E_STRIDED, NBITS, true, true, false, false, 1>(ArithF64(m), sm, src, nullptr, twd_all + (size_t)ld.mod * N, ni, logN, nullptr, raw_in, stage, rank, tid);
        else     ntt_body<ArithInt, LOG_E_STRIDED, NBITS, true, true, false, false, 1>(ArithInt(m), sm, src, nullptr, tw_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, rank, tid);
        cl.sync();
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            if (f64) ntt_body<ArithF64, LOG_E_CONTIG, K2_BITS, false, true, true, false, 2>(ArithF64(m), sm, nullptr, dst, twd_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, 2 * rank + h, tid);
            else     ntt_body<ArithInt, LOG_E_CONTIG, K2_BITS, false, true, true, false, 2>(ArithInt(m), sm, nullptr, dst, tw_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, 2 * rank + h, tid);
            __syncthreads();
        }
    } else {
        const int h = tid >> 7, t7 = tid & 127;
        if (f64) ntt_body<ArithF64, LOG_E_CONTIG, K2_BITS, false, false, false, false, 1>(ArithF64(m), sm + h * SM_C, src, nullptr, twd_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, 2 * rank + h, t7);
        else     ntt_body<ArithInt, LOG_E_CONTIG, K2_BITS, false, false, false, false, 1>(ArithInt(m), sm + h * SM_C, src, nullptr, tw_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, 2 * rank + h, t7);
        cl.sync();
        if (f64) ntt_body<ArithF64, LOG_E_STRIDED, NBITS, true, false, true, false, 2>(ArithF64(m), sm, nullptr, dst, twd_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, rank, tid);
        else     ntt_body<ArithInt, LOG_E_STRIDED, NBITS, true, false, true, false, 2>(ArithInt(m), sm, nullptr, dst, tw_all + (size_t)ld.mod * N, ni, logN, nullptr, false, stage, rank, tid);
    }
}

}  // namespace

// cluster launch: grid (N/2048, limbs), cluster (N/2048, 1, 1)
bool cluster_ntt_enabled(const Context& c, int limbs) {
    static int on = -1, max_limbs = 0;
    if (on < 0) {
        const char* s = std::getenv("ORION_B200_CLUSTER_NTT"); on = (s && std::atoi(s) != 0) ? 1 : 0;
        const char* m = std::getenv("ORION_B200_CLUSTER_NTT_MAX"); max_limbs = m ? std::atoi(m) : 48;
    }
    return on == 1 && c.logN >= 12 && c.logN <= 14 && limbs <= max_limbs;
}
template <class... KArgs, class... Args>
static void launch_cluster(void (*kernel)(KArgs...), dim3 grid, int threads, int cluster_x, Args... args) {
    g_launches++;
    if (dry_run()) return;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = 0; cfg.stream = stream();
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster_x; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 2 : 1;
    OB_HP("cudaLaunchKernelEx");
    OB_CUDA(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
}
#define OB_NTT_CLUSTER(FWD, ...) \
    switch (c.logN - K2_BITS) { \
        case 4: launch_cluster(ntt_cluster<4, FWD>, dim3(c.N >> LOG_E_STRIDED, p.n), FWD ? 128 : 256, c.N >> LOG_E_STRIDED, __VA_ARGS__); break; \
        case 5: launch_cluster(ntt_cluster<5, FWD>, dim3(c.N >> LOG_E_STRIDED, p.n), FWD ? 128 : 256, c.N >> LOG_E_STRIDED, __VA_ARGS__); break; \
        case 6: launch_cluster(ntt_cluster<6, FWD>, dim3(c.N >> LOG_E_STRIDED, p.n), FWD ? 128 : 256, c.N >> LOG_E_STRIDED, __VA_ARGS__); break; \
        default: OB_FATAL("cluster NTT: unsupported logN"); }

NttPlan make_ntt_plan(Context& c, const std::vector<LimbDesc>& limbs) {
    OB_HP("make_ntt_plan");
    NttPlan p; p.n = (int)limbs.size();
    if (p.n == 0) return p;
    std::vector<LimbDesc> second = limbs;
    for (auto& d : second) d.in_off = d.out_off;          // the second pass works in place on the output
    p.first = c.limb_list(limbs);
    p.second = c.limb_list(second);
    return p;
}

#define OB_NTT_STRIDED(S1, FWD, FINAL, ...) \
    OB_LAUNCH((ntt_pass<LOG_E_STRIDED, S1, true, FWD, FINAL>), g1, 1 << (LOG_E_STRIDED - log_ept(true, FWD)), 0, stream(), __VA_ARGS__)
#define OB_NTT_STRIDED_SWITCH(FWD, FINAL, ...) \
    switch (s1) { \
        case 4: OB_NTT_STRIDED(4, FWD, FINAL, __VA_ARGS__); break; \
        case 5: OB_NTT_STRIDED(5, FWD, FINAL, __VA_ARGS__); break; \
        case 6: OB_NTT_STRIDED(6, FWD, FINAL, __VA_ARGS__); break; \
        case 7: OB_NTT_STRIDED(7, FWD, FINAL, __VA_ARGS__); break; \
        case 8: OB_NTT_STRIDED(8, FWD, FINAL, __VA_ARGS__); break; \
        case 9: OB_NTT_STRIDED(9, FWD, FINAL, __VA_ARGS__); break; \
        default: OB_FATAL("unsupported logN"); }

// Batched forward NTT: limb i is read at in + in_off*N and written at out + out_off*N.
void ntt_forward(Context& c, const u64* in, u64* out, const NttPlan& p) {
    if (p.n == 0) return;
    if (cluster_ntt_enabled(c, p.n)) {
        OB_NTT_CLUSTER(true, in, out, p.first, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.d_ninv.as<ulonglong2>(), c.logN, 1)
        return;
    }
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, p.n), g2(c.N >> LOG_E_CONTIG, p.n);
    OB_NTT_STRIDED_SWITCH(true, false, in, out, p.first, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.d_ninv.as<ulonglong2>(), c.logN)
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, true, true>), g2, 1 << (LOG_E_CONTIG - log_ept(false, true)), 0, stream(),
              (const u64*)out, out, p.second, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.d_ninv.as<ulonglong2>(), c.logN);
}

// Second (contiguous) pass only, in place on limbs that hold the output of a first pass (ntt_forward_first / ntt1_bc).
void ntt_forward_second(Context& c, u64* buf, const NttPlan& p) {
    if (p.n == 0) return;
    dim3 g2(c.N >> LOG_E_CONTIG, p.n);
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, true, true>), g2, 1 << (LOG_E_CONTIG - log_ept(false, true)), 0, stream(),
              (const u64*)buf, buf, p.second, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.d_ninv.as<ulonglong2>(), c.logN);
}

// First (strided) pass only: the limbs are left in the policy's lazy inter-pass form for ks_ntt2_inner_product.
void ntt_forward_first(Context& c, const u64* in, u64* out, const NttPlan& p) {
    if (p.n == 0) return;
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, p.n);
    OB_NTT_STRIDED_SWITCH(true, false, in, out, p.first, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.d_ninv.as<ulonglong2>(), c.logN)
}

// tmp[p][i] (i < level) = first forward pass of the rescale's correction term, computed from last[p] (ntt1_rescale_kernel)
void ntt1_rescale(Context& c, int level, int polys, const u64* last, u64* tmp) {
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, level * polys);
    const int tpb = 1 << (LOG_E_STRIDED - log_ept(true, true));
#define OB_RS(S1V) OB_LAUNCH((ntt1_rescale_kernel<S1V>), g1, tpb, 0, stream(), last, tmp, level, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN)
    switch (s1) { case 4: OB_RS(4); break; case 5: OB_RS(5); break; case 6: OB_RS(6); break; case 7: OB_RS(7); break;
                  case 8: OB_RS(8); break; case 9: OB_RS(9); break; default: OB_FATAL("unsupported logN"); }
#undef OB_RS
}

// acc[z][c][t] = sum_j NTT2(digits[z][j][t]) * evk_z[j][c][chain(t)]  -- see ntt2_ip_kernel.  One launch per arithmetic
// policy (the FP64 instantiation needs half the accumulator registers of the integer one).
void ks_ntt2_inner_product(Context& c, int level, int n, const u64* c1, size_t c1_zstride, const u64* digits, size_t dig_zstride,
                           const IpMulti& m, u64* out, size_t out_zstride) {
    OB_CHECK(n >= 1 && n <= IPM_MAX, "inner product batch size");
    const int L = level + 1, K = c.nP, W = L + K, beta = c.num_digits(level);
    OB_CHECK(W <= 128 && beta <= 64, "fused inner product: chain too long");
    LimbTab tf{}, ti{};
    int nf = 0, ni = 0;
    for (int t = 0; t < W; ++t) {
        const int ci = t < L ? t : c.nQ + (t - L);
        if (c.mc[ci].bits <= F64_MAX_BITS) tf.t[nf++] = (unsigned char)t; else ti.t[ni++] = (unsigned char)t;
    }
    const int tiles = c.N >> LOG_E_CONTIG, tpb = 1 << (LOG_E_CONTIG - 3);
    constexpr size_t SMX = 2 * ((1 << LOG_E_CONTIG) + (1 << (LOG_E_CONTIG - 3))) * 8;      // the two exchange tiles
    // key-ring depth of the FP64 instantiation: ORION_B200_IP_STAGES = 0 (direct loads) / 2 / 3
    static const int stages = [] { const char* e = std::getenv("ORION_B200_IP_STAGES"); const int v = e ? std::atoi(e) : IP_STAGES_DEFAULT;
                                   return (v == 2 || v == 3) ? v : 0; }();
    static bool attr_set = false;
    if (!attr_set) {
        OB_CUDA(cudaFuncSetAttribute(ntt2_ip_kernel<ArithF64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SMX + 2 * 2 * tpb * IP_SLOT_BYTES)));
        OB_CUDA(cudaFuncSetAttribute(ntt2_ip_kernel<ArithF64, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SMX + 3 * 2 * tpb * IP_SLOT_BYTES)));
        attr_set = true;
    }
#define OB_IP_ARGS c1, c1_zstride, digits, dig_zstride, m, out, out_zstride
    // ORION_B200_IP_OVERLAP=1: the integer-path launch (few limbs, multiplier-bound) goes to a side stream so it can share
    // the GPU with the FP64-path launch (HBM-bound) instead of following it
    static const bool overlap = [] { const char* e = std::getenv("ORION_B200_IP_OVERLAP"); return e && std::atoi(e) != 0; }();
    // CTA order: ciphertext index fastest (ORION_B200_IP_ZFAST=0 restores tile-fastest with the ciphertext in blockIdx.z)
    static const bool imac = [] { const char* e = std::getenv("ORION_B200_IP_IMAC"); return e && std::atoi(e) != 0; }();
    static const bool zfast = [] { const char* e = std::getenv("ORION_B200_IP_ZFAST"); return !(e && std::atoi(e) == 0); }();
    const int nzx = zfast ? n : 1;
    const dim3 gf(tiles * nzx, nf, zfast ? 1 : n), gi(tiles * nzx, ni, zfast ? 1 : n);
    StreamFork fork((overlap && nf && ni) ? 2 : 1);
    if (ni && fork.size() > 0) {
        fork.branch(0);
        ProbeScope ps(PROBE_IP_INT);
        OB_LAUNCH((ntt2_ip_kernel<ArithInt, 0>), gi, tpb, SMX, stream(), OB_IP_ARGS, ti, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
        fork.branch(1);
    }
    if (nf) {
        probe_begin(PROBE_IP_F64);
        const size_t ring = (size_t)stages * 2 * tpb * IP_SLOT_BYTES;
        if (stages == 3) OB_LAUNCH((ntt2_ip_kernel<ArithF64, 3>), gf, tpb, SMX + ring, stream(), OB_IP_ARGS, tf, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
        else if (stages == 2) OB_LAUNCH((ntt2_ip_kernel<ArithF64, 2>), gf, tpb, SMX + ring, stream(), OB_IP_ARGS, tf, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
        else if (imac) OB_LAUNCH((ntt2_ip_kernel<ArithF64I, 0>), gf, tpb, SMX, stream(), OB_IP_ARGS, tf, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
        else OB_LAUNCH((ntt2_ip_kernel<ArithF64, 0>), gf, tpb, SMX, stream(), OB_IP_ARGS, tf, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
        probe_end(PROBE_IP_F64);
    }
    if (ni && fork.size() == 0) {
        ProbeScope ps(PROBE_IP_INT);
        OB_LAUNCH((ntt2_ip_kernel<ArithInt, 0>), gi, tpb, SMX, stream(), OB_IP_ARGS, ti, L, K, c.nQ, c.nmod, beta, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN, nzx);
    }
    fork.join();
#undef OB_IP_ARGS
}

// out[p][i] = (a[p*a_stride + i] - NTT(b[p*nl + i])) * mulc[i] (+ addend), limbs i < nl of the Q chain, b from ntt_forward_first
void ntt2_submul(Context& c, int nl, int polys, const u64* a, int a_stride, const u64* b, const ulonglong2* mulc,
                 const u64* addend, int addend_stride, int addend_polys, int addend_period, int addend_zstride,
                 u64* out, int out_stride) {
    if (addend_period <= 0) { addend_period = 1 << 30; addend_zstride = 0; }
    OB_LAUNCH(ntt2_submul_kernel, dim3(c.N >> LOG_E_CONTIG, nl, polys), 1 << (LOG_E_CONTIG - 3), 0, stream(), a, a_stride, b, nl,
              addend, addend_stride, addend_polys, addend_period, addend_zstride, out, out_stride, mulc, c.dmc(),
              c.tw_fwd(), c.twd_fwd(), c.logN);
}

// First (contiguous) pass of the inverse transform only: the limbs are left in the policy's lazy inter-pass form for intt2_prep.
void ntt_inverse_first(Context& c, const u64* in, u64* out, const NttPlan& p) {
    if (p.n == 0) return;
    dim3 g2(c.N >> LOG_E_CONTIG, p.n);
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, false, false>), g2, 1 << (LOG_E_CONTIG - log_ept(false, true)), 0, stream(),
              in, out, p.first, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN);
}

// buf: source limbs after ntt_inverse_first (conversion z's sources at (descs[z].in_off + i) * N); on return y form + vout[z][N]
void intt2_bc_prep(Context& c, u64* buf, unsigned char* vout, const BcDev* descs, int n_conv, bool centered) {
    if (n_conv == 0) return;
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, n_conv);
    const int tpb = 1 << (LOG_E_STRIDED - log_ept(true, false));
#define OB_PREP(S1V) do { if (centered) OB_LAUNCH((intt2_prep_kernel<S1V, true>), g1, tpb, 0, stream(), buf, vout, descs, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN); \
                          else OB_LAUNCH((intt2_prep_kernel<S1V, false>), g1, tpb, 0, stream(), buf, vout, descs, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN); } while (0)
    switch (s1) { case 4: OB_PREP(4); break; case 5: OB_PREP(5); break; case 6: OB_PREP(6); break; case 7: OB_PREP(7); break;
                  case 8: OB_PREP(8); break; case 9: OB_PREP(9); break; default: OB_FATAL("unsupported logN"); }
#undef OB_PREP
}

// first forward NTT pass of every destination limb with its base conversion computed on the fly (see ntt1_bc_kernel)
void ntt1_bc(Context& c, const u64* ybuf, const unsigned char* vbuf, u64* out, const BcLimbDev* limbs, int n_limbs, const BcDev* descs,
             bool centered) {
    if (n_limbs == 0) return;
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, n_limbs);
    const int tpb = 1 << (LOG_E_STRIDED - log_ept(true, true));
    const BcLimb* bl = (const BcLimb*)limbs;
#define OB_N1(S1V) do { if (centered) OB_LAUNCH((ntt1_bc_kernel<S1V, true>), g1, tpb, 0, stream(), ybuf, vbuf, out, bl, descs, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN); \
                        else OB_LAUNCH((ntt1_bc_kernel<S1V, false>), g1, tpb, 0, stream(), ybuf, vbuf, out, bl, descs, c.dmc(), c.tw_fwd(), c.twd_fwd(), c.logN); } while (0)
    switch (s1) { case 4: OB_N1(4); break; case 5: OB_N1(5); break; case 6: OB_N1(6); break; case 7: OB_N1(7); break;
                  case 8: OB_N1(8); break; case 9: OB_N1(9); break; default: OB_FATAL("unsupported logN"); }
#undef OB_N1
}

void ntt_inverse(Context& c, const u64* in, u64* out, const NttPlan& p) {
    if (p.n == 0) return;
    if (cluster_ntt_enabled(c, p.n)) {
        OB_NTT_CLUSTER(false, in, out, p.first, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN, 0)
        return;
    }
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, p.n), g2(c.N >> LOG_E_CONTIG, p.n);
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, false, false>), g2, 1 << (LOG_E_CONTIG - log_ept(false, true)), 0, stream(),
              in, out, p.first, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN);
    OB_NTT_STRIDED_SWITCH(false, true, (const u64*)out, out, p.second, c.dmc(), c.tw_inv(), c.twd_inv(), c.d_ninv.as<ulonglong2>(), c.logN)
}

}  // namespace ob
