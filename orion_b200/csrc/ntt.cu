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

namespace ob {

namespace {

constexpr int LOG_E_STRIDED = 12;   // 4096 words per tile, 512 threads
constexpr int LOG_E_CONTIG = 11;    // 2048 words per tile, 256 threads
constexpr int K2_BITS = 8;          // contiguous pass handles the 8 low index bits

__device__ __forceinline__ int insert_bits(int x, int pos, int nb, int v) {
    return ((x >> pos) << (pos + nb)) | (v << pos) | (x & ((1 << pos) - 1));
}
// bijection on the low 4 bits: keeps 64-bit shared-memory accesses conflict-free for all round shapes
__device__ __forceinline__ int swz(int li) { return li ^ ((li >> 4) & 3) ^ (((li >> 5) & 3) << 2); }

__device__ __forceinline__ void ct_bfly(u64& a, u64& b, ulonglong2 w, u64 q, u64 q2) {
    u64 x = a >= q2 ? a - q2 : a;                 // [0,2q)
    u64 v = shoup_lazy(b, w.x, w.y, q);           // [0,2q)
    a = x + v;                                    // [0,4q)
    b = x - v + q2;                               // (0,4q)
}
__device__ __forceinline__ void gs_bfly(u64& a, u64& b, ulonglong2 w, u64 q, u64 q2) {
    u64 s = a + b;                                // [0,4q)
    u64 d = a - b + q2;                           // (0,4q)
    a = s >= q2 ? s - q2 : s;                     // [0,2q)
    b = shoup_lazy(d, w.x, w.y, q);               // [0,2q)
}

// forward radix-2^NB on 8>>NB independent groups held in v[]
template <int NB>
__device__ __forceinline__ void fwd_radix(u64 (&v)[8], const ulonglong2* __restrict__ tw, const int (&twb)[8], u64 q) {
    constexpr int G = 8 >> NB, R = 1 << NB;
    const u64 q2 = 2 * q;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        const int half = 1 << (NB - 1 - i);
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int m = 0; m < (1 << i); ++m) {
                ulonglong2 w = __ldg(&tw[(twb[g] << i) + m]);
#pragma unroll
                for (int k0 = 0; k0 < half; ++k0) {
                    int a = g * R + m * 2 * half + k0;
                    ct_bfly(v[a], v[a + half], w, q, q2);
                }
            }
        }
    }
}
// inverse radix-2^NB (bits ascending); twb[] is the twiddle index at the TOP bit of the round
template <int NB>
__device__ __forceinline__ void inv_radix(u64 (&v)[8], const ulonglong2* __restrict__ tw, const int (&twb)[8], u64 q) {
    constexpr int G = 8 >> NB, R = 1 << NB;
    const u64 q2 = 2 * q;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        const int half = 1 << i;
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int m = 0; m < (R >> (i + 1)); ++m) {
                ulonglong2 w = __ldg(&tw[(twb[g] << (NB - 1 - i)) + m]);
#pragma unroll
                for (int k0 = 0; k0 < half; ++k0) {
                    int a = g * R + m * 2 * half + k0;
                    gs_bfly(v[a], v[a + half], w, q, q2);
                }
            }
        }
    }
}

struct Rounds { int n; int nb[4]; };
__host__ __device__ inline Rounds split_rounds(int bits) {
    Rounds r; r.n = (bits + 2) / 3;
    int base = bits / r.n, extra = bits % r.n;
    for (int i = 0; i < r.n; ++i) r.nb[i] = base + (i < extra ? 1 : 0);
    return r;
}

// One pass.  STRIDED: tile = 2^nbits rows x C cols, transformed bits are local bits [logC, LOG_E).
//            contiguous: tile = 2^LOG_E consecutive words, transformed bits are local bits [0, nbits).
// FWD walks the transformed bits downwards, INV upwards.  FINAL: normalise (and scale by N^-1 for INV).
template <int LOG_E, int NBITS, bool STRIDED, bool FWD, bool FINAL>
__global__ void __launch_bounds__(1 << (LOG_E - 3))
ntt_pass(const u64* in, u64* out, const LimbDesc* __restrict__ limbs,
         const ModConst* __restrict__ mc, const ulonglong2* __restrict__ tw_all,
         const ulonglong2* __restrict__ ninv_all, int logN) {
    constexpr int nbits = NBITS;          // compile-time: the round schedule below unrolls into straight-line code
    __shared__ u64 sm[1 << LOG_E];
    const int tid = threadIdx.x;
    const LimbDesc ld = limbs[blockIdx.y];
    const int N = 1 << logN;
    const u64 q = mc[ld.mod].q;
    const ulonglong2* tw = tw_all + (size_t)ld.mod * N;
    const u64* src = in + (size_t)ld.in_off * N;
    u64* dst = out + (size_t)ld.out_off * N;

    const int logC = STRIDED ? LOG_E - nbits : 0;
    const int lo_bit = STRIDED ? logC : 0;               // lowest transformed local bit
    const int TOP = STRIDED ? LOG_E - 1 : logN - 1;      // stage = TOP - local bit
    const int tile_base = STRIDED ? 0 : blockIdx.x << LOG_E;
    auto gidx = [&](int li) -> int {
        if (STRIDED) return ((li >> logC) << (logN - nbits)) + (blockIdx.x << logC) + (li & ((1 << logC) - 1));
        return tile_base + li;
    };

    constexpr int NR = (NBITS + 2) / 3, RBASE = NBITS / NR, REXTRA = NBITS % NR;
    int done = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const int nb = RBASE + (r < REXTRA ? 1 : 0);
        const int G = 8 >> nb;
        const int pos = FWD ? lo_bit + nbits - done - nb : lo_bit + done;   // lowest bit of this round
        const int ptop = pos + nb - 1;
        u64 v[8]; int li[8]; int twb[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            int g = e >> nb, k = e & ((1 << nb) - 1);
            li[e] = insert_bits(tid * G + g, pos, nb, k);
        }
#pragma unroll
        for (int g = 0; g < 8; ++g)
            twb[g] = (1 << (TOP - ptop)) + ((tid * G + g) >> pos) + (tile_base >> (ptop + 1));
        if (r == 0) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = src[gidx(li[e])];
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = sm[swz(li[e])];
        }
        if (FWD) {
            if (nb == 3) fwd_radix<3>(v, tw, twb, q); else if (nb == 2) fwd_radix<2>(v, tw, twb, q); else fwd_radix<1>(v, tw, twb, q);
        } else {
            if (nb == 3) inv_radix<3>(v, tw, twb, q); else if (nb == 2) inv_radix<2>(v, tw, twb, q); else inv_radix<1>(v, tw, twb, q);
        }
        done += nb;
        if (r == NR - 1) {
            if (FINAL) {
                if (FWD) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) { u64 x = v[e]; x = x >= 2 * q ? x - 2 * q : x; v[e] = x >= q ? x - q : x; }
                } else {
                    const ulonglong2 ni = ninv_all[ld.mod];
#pragma unroll
                    for (int e = 0; e < 8; ++e) v[e] = shoup_mul(v[e], ni.x, ni.y, q);
                }
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) dst[gidx(li[e])] = v[e];
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) sm[swz(li[e])] = v[e];
            __syncthreads();
        }
    }
}

}  // namespace

NttPlan make_ntt_plan(Context& c, const std::vector<LimbDesc>& limbs) {
    NttPlan p; p.n = (int)limbs.size();
    if (p.n == 0) return p;
    std::vector<LimbDesc> second = limbs;
    for (auto& d : second) d.in_off = d.out_off;          // the second pass works in place on the output
    p.first = c.limb_list(limbs);
    p.second = c.limb_list(second);
    return p;
}

#define OB_NTT_STRIDED(S1, FWD, FINAL, ...) \
    OB_LAUNCH((ntt_pass<LOG_E_STRIDED, S1, true, FWD, FINAL>), g1, 1 << (LOG_E_STRIDED - 3), 0, stream(), __VA_ARGS__)
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
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, p.n), g2(c.N >> LOG_E_CONTIG, p.n);
    OB_NTT_STRIDED_SWITCH(true, false, in, out, p.first, c.dmc(), c.tw_fwd(), c.d_ninv.as<ulonglong2>(), c.logN)
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, true, true>), g2, 1 << (LOG_E_CONTIG - 3), 0, stream(),
              (const u64*)out, out, p.second, c.dmc(), c.tw_fwd(), c.d_ninv.as<ulonglong2>(), c.logN);
}

void ntt_inverse(Context& c, const u64* in, u64* out, const NttPlan& p) {
    if (p.n == 0) return;
    const int s1 = c.logN - K2_BITS;
    dim3 g1(c.N >> LOG_E_STRIDED, p.n), g2(c.N >> LOG_E_CONTIG, p.n);
    OB_LAUNCH((ntt_pass<LOG_E_CONTIG, K2_BITS, false, false, false>), g2, 1 << (LOG_E_CONTIG - 3), 0, stream(),
              in, out, p.first, c.dmc(), c.tw_inv(), c.d_ninv.as<ulonglong2>(), c.logN);
    OB_NTT_STRIDED_SWITCH(false, true, (const u64*)out, out, p.second, c.dmc(), c.tw_inv(), c.d_ninv.as<ulonglong2>(), c.logN)
}

}  // namespace ob
