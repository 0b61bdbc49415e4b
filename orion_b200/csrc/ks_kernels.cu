// ks_kernels.cu -- key-switching kernels for sm_100a: RNS fast base conversion (ModUp per digit, ModDown),
// the evaluation-key inner product, and the ModDown / rescale epilogues.
//
// Base conversion follows the Lattigo definition named in SURVEY.md section 8(a) a10 (float64 overflow
// estimate, accumulated in source-limb order), NOT HEonGPU's float32 estimate
// (HE/src/heongpu/lib/kernel/switchkey.cu:829-845):
//     y_i = x_i * (Q_S/q_i)^-1 mod q_i,  v = (u64) sum_i double(y_i)/double(q_i)
//     out_t = sum_i y_i * (Q_S/q_i mod p_t) - v * Q_S   (mod p_t)
// The contraction over i is done with 128-bit multiply-accumulate against a Montgomery-form matrix and
// one REDC per output, so the per-output cost is ns MACs + 2 wide multiplies.
#include "kernels.h"

namespace ob {

namespace {

constexpr int TPB = 256;
constexpr int MAX_NS = 16;       // source limbs per conversion (digit size or #P)
constexpr int DST_PER_CTA = 4;   // destination limbs handled by one CTA (y_i recomputed per CTA)

template <bool CENTERED>
__global__ void __launch_bounds__(TPB)
baseconv_kernel(const u64* __restrict__ in, u64* __restrict__ out, int N, int ns, int nd,
                const int* __restrict__ src_mod, const int* __restrict__ dst_mod, const LimbDesc* __restrict__ out_off,
                const ulonglong2* __restrict__ inv, const u64* __restrict__ mat, const u64* __restrict__ negvq,
                const u64* __restrict__ half_src, const u64* __restrict__ half_dst,
                const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int x = blockIdx.x * TPB + threadIdx.x;
    u64 y[MAX_NS];
    double vf = 0.0;
#pragma unroll
    for (int i = 0; i < MAX_NS; ++i) {
        if (i < ns) {
            const ModConst& m = mc[src_mod[i]];
            u64 xi = in[(size_t)i * N + x];
            if (CENTERED) xi = add_mod(xi, half_src[i], m.q);
            const ulonglong2 w = inv[i];
            y[i] = shoup_mul(xi, w.x, w.y, m.q);
            vf = __dadd_rn(vf, __ddiv_rn(__ull2double_rn(y[i]), m.qd));
        }
    }
    // a single source modulus without centring is a plain copy + reduction (v is 0 by definition)
    const u64 v = (ns == 1 && !CENTERED) ? 0 : (u64)vf;
    const int d0 = blockIdx.y * DST_PER_CTA;
#pragma unroll
    for (int dd = 0; dd < DST_PER_CTA; ++dd) {
        const int d = d0 + dd;
        if (d >= nd) break;
        const ModConst m = mc[dst_mod[d]];
        u64 hi = 0, lo = 0;
#pragma unroll
        for (int i = 0; i < MAX_NS; ++i)
            if (i < ns) mac128(hi, lo, y[i], __ldg(&mat[(size_t)i * nd + d]), m.q);
        u64 r = mont_redc(hi, lo, m.q, m.qinv_neg);
        r = add_mod(r, __ldg(&negvq[(size_t)d * (ns + 1) + v]), m.q);
        if (CENTERED) r = sub_mod(r, half_dst[d], m.q);
        out[(size_t)out_off[d].in_off * N + x] = r;
    }
}


// Batched base conversion: one CTA column per conversion (blockIdx.y), every thread converts one coefficient to ALL
// destination limbs.  The conversion's constants are staged once per CTA in shared memory, so the inner loop is
// ns MACs + 1 REDC per output with no dependent global loads.
constexpr int BC_TPB = 128;
constexpr int BC_F64_BITS = 46;      // same bound as the NTT's FP64 path (ntt.cu)
// 128-bit multiply-accumulate without the high-word reduction: every product's high word is below q/4 (y < 2^62), so
// up to three terms stay below q * 2^64, which is all REDC needs
__device__ __forceinline__ void mac128_raw(u64& hi, u64& lo, u64 a, u64 b) {
    const u64 pl = a * b, ph = __umul64hi(a, b);
    lo += pl;
    hi += ph + (lo < pl);
}
// 96-bit multiply-accumulate for a 32-bit residue (every source prime below 2^32, e.g. the 30-bit primes of the shipped
// ResNet-20 chain): two IMAD.WIDE instead of four, no high-word reduction
__device__ __forceinline__ void mac96_raw(u64& hi, u64& lo, uint32_t a, u64 b) {
    const u64 p0 = (u64)a * (uint32_t)b, p1 = (u64)a * (uint32_t)(b >> 32);
    const u64 l = p0 + (p1 << 32);
    lo += l;
    hi += (p1 >> 32) + (l < p0) + (lo < l);
}
// MM = highest accumulation mode of the integer path over the conversions of the batch (BcDev::mm is per conversion),
// chosen on the host from the source primes:
//   0: mac128 (conditional high-word reduction after every term)
//   1: mac128_raw -- valid when sum_i q_src_i <= 2^64: sum_i y_i*w_i < q_t * 2^64, which is all REDC needs
//   2: mac96_raw  -- additionally every source prime is below 2^32
// NS = compile-time bound on the source-limb count of every conversion in the batch.  VEC = adjacent coefficients per
// thread (2: 128-bit loads/stores and every per-destination constant fetched once for both -- large rings; 1: twice
// the threads -- small rings).  gridDim.z splits the destination limbs into chunks so that a conversion of a small
// ring (logN = 13, 8 x 61-bit special primes in the bootstrapping chain) still fills the 148 SMs.
template <bool CENTERED, int NS, int VEC, int MM>
__global__ void __launch_bounds__(BC_TPB)
baseconv_batch_kernel(const u64* __restrict__ in, u64* __restrict__ out, int N, const BcDev* __restrict__ descs,
                      const ModConst* __restrict__ mc) {
    OB_PDL_LAUNCH();                                    // constants staged below are host-built: read before the wait
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const BcDev d = descs[blockIdx.y];
    const int ns = d.ns, nd = d.nd;
    const int chunk = (nd + gridDim.z - 1) / gridDim.z;
    const int t0 = blockIdx.z * chunk, ndc = min(nd, t0 + chunk) - t0;          // this CTA's destinations [t0, t0+ndc)
    if (ndc <= 0) { OB_PDL_WAIT(); return; }
    u64* s_mat = (u64*)smem_raw;                       // [ns][chunk]
    u64* s_neg = s_mat + ns * chunk;                   // [chunk][ns+1]
    double* s_matd = (double*)(s_neg + chunk * (ns + 1)); // [ns][chunk]
    double* s_negd = s_matd + ns * chunk;              // [chunk][ns+1]
    u64* s_dq = (u64*)(s_negd + chunk * (ns + 1));     // [chunk]
    u64* s_dqi = s_dq + chunk;                         // [chunk]
    u64* s_dh = s_dqi + chunk;                         // [chunk]
    double* s_dqinv = (double*)(s_dh + chunk);         // [chunk]  1/q as double; 0 marks an integer-path destination
    u64* s_sq = (u64*)(s_dqinv + chunk);               // [ns]
    u64* s_sh = s_sq + ns;                             // [ns]
    double* s_sqd = (double*)(s_sh + ns);              // [ns]
    ulonglong2* s_inv = (ulonglong2*)(s_sqd + ns + (ns & 1));   // [ns], 16-byte aligned
    int* s_off = (int*)(s_inv + ns);                   // [chunk]
    const bool f64 = !CENTERED && d.f64_src;
    const int mm = d.mm & 3;                           // this conversion's accumulation mode (CTA-uniform), at most MM
    const bool w32_ok = (d.mm & 4) != 0;               // 96-bit products for destinations below 2^32 (ORION_B200_BC_W32=0 disables)
    for (int i = threadIdx.x; i < ns * ndc; i += BC_TPB) {
        const int r = i / ndc, t = i - r * ndc;
        s_mat[r * chunk + t] = d.mat[r * nd + t0 + t]; s_matd[r * chunk + t] = d.mat_d[r * nd + t0 + t];
    }
    for (int i = threadIdx.x; i < ndc * (ns + 1); i += BC_TPB) {
        s_neg[i] = d.negvq[t0 * (ns + 1) + i]; s_negd[i] = d.negvq_d[t0 * (ns + 1) + i];
    }
    for (int i = threadIdx.x; i < ndc; i += BC_TPB) {
        const ModConst& m = mc[d.dst_mod[t0 + i]];
        s_dq[i] = m.q; s_dqi[i] = m.qinv_neg; s_dh[i] = CENTERED ? d.half_dst[t0 + i] : 0;
        s_dqinv[i] = (f64 && m.bits <= BC_F64_BITS) ? m.qinv_d : 0.0;
        s_off[i] = d.out_base + (int)d.out_off[t0 + i].in_off;
    }
    for (int i = threadIdx.x; i < ns; i += BC_TPB) {
        const ModConst& m = mc[d.src_mod[i]];
        s_sq[i] = m.q; s_sqd[i] = m.qd; s_inv[i] = d.inv[i]; s_sh[i] = CENTERED ? d.half_src[i] : 0;
    }
    __syncthreads();
    OB_PDL_WAIT();
    const int x = (blockIdx.x * BC_TPB + threadIdx.x) * VEC;
    u64 y[VEC][NS];
    double yd[VEC][NS];
    double vf[VEC];
#pragma unroll
    for (int c = 0; c < VEC; ++c) vf[c] = 0.0;
#pragma unroll
    for (int i = 0; i < NS; ++i) {
        if (i < ns) {
            const u64 q = s_sq[i];
            u64 xi[VEC];
            const u64* src = in + (size_t)(d.in_off + i) * N + x;
            if (VEC == 2) { const ulonglong2 t = *(const ulonglong2*)src; xi[0] = t.x; xi[VEC - 1] = t.y; }
            else xi[0] = *src;
            const ulonglong2 w = s_inv[i];
            const double qd = s_sqd[i];
#pragma unroll
            for (int c = 0; c < VEC; ++c) {
                if (CENTERED) xi[c] = add_mod(xi[c], s_sh[i], q);
                y[c][i] = shoup_mul(xi[c], w.x, w.y, q);
                yd[c][i] = __ull2double_rn(y[c][i]);                      // exact below 2^53
                vf[c] = __dadd_rn(vf[c], __ddiv_rn(yd[c][i], qd));
            }
        }
    }
    // a single source modulus without centring is a plain copy + reduction (v is 0 by definition)
    int v[VEC];
#pragma unroll
    for (int c = 0; c < VEC; ++c) v[c] = (ns == 1 && !CENTERED) ? 0 : (int)vf[c];
#pragma unroll 2
    for (int t = 0; t < ndc; ++t) {
        const u64 q = s_dq[t];
        const double qinv = s_dqinv[t];
        u64 r[VEC];
        if (!CENTERED && qinv != 0.0) {
            // FP64 contraction (sources and destination of at most 46 bits): each product is reduced to a centred
            // residue with 6 exact DFMA-class operations; B200 issues those at twice the rate of IMAD.WIDE
            const double qd = __ull2double_rn(q);
            const double M = 6755399441055744.0;
            double a[VEC];
#pragma unroll
            for (int c = 0; c < VEC; ++c) a[c] = s_negd[t * (ns + 1) + v[c]];
#pragma unroll
            for (int i = 0; i < NS; ++i)
                if (i < ns) {
                    const double w = s_matd[i * chunk + t];
#pragma unroll
                    for (int c = 0; c < VEC; ++c) {
                        const double h = __dmul_rn(yd[c][i], w), l = __fma_rn(yd[c][i], w, -h);
                        const double tq = __dsub_rn(__fma_rn(h, qinv, M), M);
                        a[c] = __dadd_rn(a[c], __dadd_rn(__fma_rn(-tq, qd, h), l));
                    }
                }
            if (ns <= 4) {                       // bc_raw_out(): the NTT that follows takes the unreduced double as it is
#pragma unroll
                for (int c = 0; c < VEC; ++c) r[c] = (u64)__double_as_longlong(a[c]);
            } else {
#pragma unroll
                for (int c = 0; c < VEC; ++c) {
                    const double tq = __dsub_rn(__fma_rn(a[c], qinv, M), M);
                    double ac = __fma_rn(-tq, qd, a[c]);
                    ac = ac < 0.0 ? __dadd_rn(ac, qd) : ac;
                    r[c] = (u64)__double_as_longlong(__dadd_rn(ac, 4503599627370496.0)) & 0x000FFFFFFFFFFFFFull;
                }
            }
        } else {
            u64 hi[VEC], lo[VEC];
#pragma unroll
            for (int c = 0; c < VEC; ++c) hi[c] = lo[c] = 0;
            // destination below 2^32 (and raw accumulation allowed): the Montgomery-form matrix word is the 32-bit operand of
            // a 96-bit product, whatever the size of the sources (ModDown from 61-bit special primes onto 30-bit limbs,
            // ModUp digits that mix a 60-bit and 30-bit primes)
            const bool w32 = MM >= 1 && mm >= 1 && w32_ok && (q >> 32) == 0;
            if (w32 && !(MM == 2 && mm == 2)) {
#pragma unroll
                for (int i = 0; i < NS; ++i)
                    if (i < ns) {
                        const uint32_t w = (uint32_t)s_mat[i * chunk + t];
#pragma unroll
                        for (int c = 0; c < VEC; ++c) mac96_raw(hi[c], lo[c], w, y[c][i]);
                    }
            } else {
#pragma unroll
                for (int i = 0; i < NS; ++i)
                    if (i < ns) {
                        const u64 w = s_mat[i * chunk + t];
#pragma unroll
                        for (int c = 0; c < VEC; ++c) {
                            if (MM == 2 && mm == 2) mac96_raw(hi[c], lo[c], (uint32_t)y[c][i], w);
                            else if ((MM >= 1 && mm >= 1) || NS <= 3) mac128_raw(hi[c], lo[c], y[c][i], w);
                            else mac128(hi[c], lo[c], y[c][i], w, q);
                        }
                    }
            }
            const u64 qi = s_dqi[t];
#pragma unroll
            for (int c = 0; c < VEC; ++c) {
                r[c] = add_mod(mont_redc(hi[c], lo[c], q, qi), s_neg[t * (ns + 1) + v[c]], q);
                if (CENTERED) r[c] = sub_mod(r[c], s_dh[t], q);
            }
        }
        u64* dst = out + (size_t)s_off[t] * N + x;
        if (VEC == 2) *(ulonglong2*)dst = make_ulonglong2(r[0], r[VEC - 1]); else *dst = r[0];
    }
}

// ---- INT8 tensor-core base-conversion contraction (north star: "tried ... kept only if ncu shows a gain" -- it does: DESIGN.md 7) ----
// Same words as baseconv_batch_kernel<false, 8, *, *> on conversions with ns <= 8 sources (used for batches that hold a digit of
// 5..8 sources: the ModUp of the bootstrapping chain; shorter last digits ride along zero-padded).  A warp owns RT tiles of 16 coefficients; lane (g, j) = (lane >> 2, lane & 3) computes y_i for the sources
// i in {j, j + 4} and the rows {g, g + 8} of each tile -- exactly the 32-bit halves its A fragments of mma.m16n8k32.u8 need -- and
// the overflow estimate v from the quad's quotients summed in source order (bit-identical with the scalar kernel).  Per block of
// four destinations: 4 (2 when all four are below 2^32) n8 tiles x 2 K-chunks of MMAs against the precomputed B fragments
// (context.cu base_conv), then lane j recombines the eight exact byte sums of destination 4 blk + j into a 79-bit integer,
// reduces it with one REDC and adds -v Q.  CENTERED: the ModDown variant (+ Q/2 on the way in, - Q/2 on the way out).
// On by default; ORION_B200_BC_TC=0 restores the scalar kernel.
__device__ __forceinline__ void mma_u8(int (&c)[4], const uint32_t (&a)[4], const uint2 b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}
constexpr int TC_TPB = 128;
struct TcDst { u64 q, qi, dh; int off, raw; };          // per destination: modulus, -q^-1, Q/2 mod q (centred), output limb, raw-double flag
// shared memory of a CTA that covers nb destination blocks: B fragments [nb][8][32] uint2, TcDst [nb*4], -vQ table [nb*4][9]
__host__ __device__ inline size_t tc_smem_bytes(int nb) { return (size_t)nb * (8 * 32 * 8 + 4 * sizeof(TcDst) + 4 * 9 * 8); }
template <int RT, bool CENTERED>
__global__ void __launch_bounds__(TC_TPB)
baseconv_tc_kernel(const u64* __restrict__ in, u64* __restrict__ out, int N, const BcDev* __restrict__ descs,
                   const ModConst* __restrict__ mc) {
    OB_PDL_LAUNCH();                                    // the tables staged below are host-built: read before the wait
    extern __shared__ __align__(16) unsigned char tc_smem[];
    const BcDev d = descs[blockIdx.y];
    const int ns = d.ns, nd = d.nd;
    const int nblk = (nd + 3) >> 2, bchunk = (nblk + gridDim.z - 1) / gridDim.z;
    const int blk0 = blockIdx.z * bchunk, blk1 = min(nblk, blk0 + bchunk);
    if (blk0 >= blk1) { OB_PDL_WAIT(); return; }
    const int nb = blk1 - blk0;
    uint2* s_frag = (uint2*)tc_smem;
    TcDst* s_dst = (TcDst*)(s_frag + (size_t)bchunk * 256);
    u64* s_neg = (u64*)(s_dst + (size_t)bchunk * 4);
    // stage this CTA's share of the conversion: one coalesced pass, overlapped with the source loads below
    for (int i = threadIdx.x; i < nb * 256; i += TC_TPB) s_frag[i] = __ldg(d.tc_frag + (size_t)blk0 * 256 + i);
    for (int i = threadIdx.x; i < nb * 4; i += TC_TPB) {
        const int t = 4 * blk0 + i;
        TcDst e{3, 0, 0, -1, 0};
        if (t < nd) {
            const ModConst& m = mc[d.dst_mod[t]];
            e.q = m.q; e.qi = m.qinv_neg; e.off = d.out_base + (int)d.out_off[t].in_off;
            e.raw = (!CENTERED && bc_raw_out(d.f64_src != 0, ns, m.bits)) ? 1 : 0;      // the NTT that follows reads a lazy double (kernels.h)
            if (CENTERED) e.dh = d.half_dst[t];
        }
        s_dst[i] = e;
    }
    for (int i = threadIdx.x; i < nb * 4 * (ns + 1); i += TC_TPB) {
        const int tt = i / (ns + 1), vv = i - tt * (ns + 1), t = 4 * blk0 + tt;
        s_neg[tt * 9 + vv] = t < nd ? d.negvq[(size_t)t * (ns + 1) + vv] : 0;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, j = lane & 3;
    const int xbase = (blockIdx.x * (TC_TPB / 32) + warp) * (16 * RT);
    uint32_t a[RT][2][4];
    int v[RT][2];
    {
        u64 sq[2], sw[2], swp[2], sh[2]; double sqd[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int i = j + 4 * h;
            sq[h] = 1; sw[h] = swp[h] = sh[h] = 0; sqd[h] = 1.0;
            if (i < ns) {
                const ModConst& m = mc[d.src_mod[i]]; sq[h] = m.q; sqd[h] = m.qd; const ulonglong2 w = d.inv[i]; sw[h] = w.x; swp[h] = w.y;
                if (CENTERED) sh[h] = d.half_src[i];
            }
        }
        OB_PDL_WAIT();                                  // `in` is the predecessor's output
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            double quo[2][2];
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int rr = 0; rr < 2; ++rr) {
                    const int i = j + 4 * h, x = xbase + r * 16 + g + 8 * rr;
                    u64 y = 0;
                    if (i < ns) {
                        u64 xi = in[(size_t)(d.in_off + i) * N + x];
                        if (CENTERED) xi = add_mod(xi, sh[h], sq[h]);
                        y = shoup_mul(xi, sw[h], swp[h], sq[h]);
                    }
                    a[r][h][rr] = (uint32_t)y; a[r][h][2 + rr] = (uint32_t)(y >> 32);
                    quo[h][rr] = __ddiv_rn(__ull2double_rn(y), sqd[h]);
                }
#pragma unroll
            for (int rr = 0; rr < 2; ++rr) {
                double vf = 0.0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const double t = __shfl_sync(0xffffffffu, quo[i >> 2][rr], (lane & ~3) | (i & 3));
                    if (i < ns) vf = __dadd_rn(vf, t);
                }
                v[r][rr] = (ns == 1 && !CENTERED) ? 0 : (int)vf;        // a single source is a plain copy + reduction
            }
        }
    }
    __syncthreads();
    for (int b = 0; b < nb; ++b) {
        const TcDst e = s_dst[b * 4 + j];
        const bool valid = e.off >= 0;
        const bool small = __all_sync(0xffffffffu, (e.q >> 32) == 0);
        int c[RT][4][4];
#pragma unroll
        for (int r = 0; r < RT; ++r)
#pragma unroll
            for (int tau = 0; tau < 4; ++tau)
#pragma unroll
                for (int k = 0; k < 4; ++k) c[r][tau][k] = 0;
        const uint2* fr = s_frag + (size_t)b * 256 + lane;
#pragma unroll
        for (int tau = 0; tau < 4; ++tau) {
            if (tau < 2 || !small) {
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    const uint2 bf = fr[(tau * 2 + ch) * 32];
#pragma unroll
                    for (int r = 0; r < RT; ++r) mma_u8(c[r][tau], a[r][ch], bf);
                }
            }
        }
        if (valid) {
            const u64* neg = s_neg + (b * 4 + j) * 9;
#pragma unroll
            for (int r = 0; r < RT; ++r)
#pragma unroll
                for (int rr = 0; rr < 2; ++rr) {
                    // byte sums S_b = c[r][b >> 1][2 rr + (b & 1)], each below 2^22
                    const u64 A = (u64)(uint32_t)c[r][0][2 * rr] + ((u64)(uint32_t)c[r][0][2 * rr + 1] << 8) +
                                  ((u64)(uint32_t)c[r][1][2 * rr] << 16) + ((u64)(uint32_t)c[r][1][2 * rr + 1] << 24);
                    u64 lo = A, hi = 0;
                    if (!small) {
                        const u64 B = (u64)(uint32_t)c[r][2][2 * rr] + ((u64)(uint32_t)c[r][2][2 * rr + 1] << 8) +
                                      ((u64)(uint32_t)c[r][3][2 * rr] << 16) + ((u64)(uint32_t)c[r][3][2 * rr + 1] << 24);
                        lo = A + (B << 32);
                        hi = (B >> 32) + (lo < A);
                    }
                    u64 res = mont_redc(hi, lo, e.q, e.qi);
                    res = add_mod(res, neg[v[r][rr]], e.q);
                    if (CENTERED) res = sub_mod(res, e.dh, e.q);
                    if (e.raw) res = (u64)__double_as_longlong(__ull2double_rn(res));
                    out[(size_t)e.off * N + xbase + r * 16 + g + 8 * rr] = res;
                }
        }
    }
}

// acc[c][t][x] (+)= sum_j dig_j[t][x] * evk[j][c][chain(t)][x]   (evk Montgomery form)
template <bool GATHER, bool ACCUM>
__global__ void __launch_bounds__(TPB)
inner_product_kernel(const u64* __restrict__ c1, const u64* __restrict__ digits, const u64* __restrict__ evk,
                     u64* __restrict__ acc, int N, int L, int K, int nQ, int nmod, int beta,
                     const uint32_t* __restrict__ gather, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, W = L + K;
    const int x = blockIdx.x * TPB + threadIdx.x;
    const int ci = t < L ? t : nQ + (t - L);
    const ModConst m = mc[ci];
    const int own = t < L ? t / K : -1;
    const int xs = GATHER ? (int)__ldg(&gather[x]) : x;
    u64 h0 = 0, l0 = 0, h1 = 0, l1 = 0;
    if ((m.q >> 32) == 0) {                          // limb below 2^32: word-sized products, folded every 2^(64 - 2 bits) terms
        const int fold = 1 << min(30, 64 - 2 * m.bits);
        u64 s0 = 0, s1 = 0;
        int cnt = 0;
        for (int j = 0; j < beta; ++j) {
            const u64 d = (j == own) ? c1[(size_t)t * N + xs] : digits[((size_t)j * W + t) * N + xs];
            s0 += (u64)(uint32_t)d * (uint32_t)evk[(((size_t)j * 2 + 0) * nmod + ci) * N + x];
            s1 += (u64)(uint32_t)d * (uint32_t)evk[(((size_t)j * 2 + 1) * nmod + ci) * N + x];
            if (++cnt == fold) { l0 += s0; h0 += (l0 < s0); l1 += s1; h1 += (l1 < s1); s0 = s1 = 0; cnt = 0; }
        }
        l0 += s0; h0 += (l0 < s0); l1 += s1; h1 += (l1 < s1);
    } else {
        for (int j = 0; j < beta; ++j) {
            const u64 d = (j == own) ? c1[(size_t)t * N + xs] : digits[((size_t)j * W + t) * N + xs];
            const u64 k0 = evk[(((size_t)j * 2 + 0) * nmod + ci) * N + x];
            const u64 k1 = evk[(((size_t)j * 2 + 1) * nmod + ci) * N + x];
            mac128(h0, l0, d, k0, m.q);
            mac128(h1, l1, d, k1, m.q);
        }
    }
    u64 r0 = mont_redc(h0, l0, m.q, m.qinv_neg), r1 = mont_redc(h1, l1, m.q, m.qinv_neg);
    const size_t o0 = (size_t)t * N + x, o1 = ((size_t)W + t) * N + x;
    if (ACCUM) { r0 = add_mod(r0, acc[o0], m.q); r1 = add_mod(r1, acc[o1], m.q); }
    acc[o0] = r0; acc[o1] = r1;
}

// Batched inner product: blockIdx.z selects one of several independent (digits, evaluation key) pairs -- the baby steps
// (shared digits, one key each) or the giant steps (own digits, one key each) of a hoisted BSGS linear transform.
// Fused epilogue: the c0 contribution is added to polynomial 0 and the result is written through the NTT-domain
// automorphism of that step (scatter table), replacing three extra launches per step.
//   ADD_MODE 1: += (P mod q_t) * addend[z][t][x] for t < L   (baby steps: c0 lifted to QP)
//   ADD_MODE 2: += addend[z][t][x] for every t                (giant steps: the c0 half of the inner sum, already in QP)
template <int ADD_MODE>
__global__ void __launch_bounds__(TPB)
inner_product_multi_kernel(const u64* __restrict__ c1, size_t c1_zstride, const u64* __restrict__ digits, size_t dig_zstride,
                           const IpMulti m, u64* __restrict__ out, size_t out_zstride,
                           const u64* __restrict__ addend, size_t addend_zstride,
                           int N, int L, int K, int nQ, int nmod, int beta,
                           const ulonglong2* __restrict__ pmodq, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, z = blockIdx.z, W = L + K;
    const int x = (blockIdx.x * TPB + threadIdx.x) * 2;          // two adjacent coefficients per thread: 128-bit streams
    const int ci = t < L ? t : nQ + (t - L);
    const ModConst mm = mc[ci];
    const int own = t < L ? t / K : -1;
    const u64* __restrict__ evk = m.evk[z];
    const u64* __restrict__ c1z = c1 + (size_t)z * c1_zstride;
    const u64* __restrict__ dz = digits + (size_t)z * dig_zstride;
    u64 h0a = 0, l0a = 0, h1a = 0, l1a = 0, h0b = 0, l0b = 0, h1b = 0, l1b = 0;
    if ((mm.q >> 32) == 0) {
        // limb below 2^32 (30-bit scale primes): each product fits a word and 2^(64 - 2 bits) of them are summed with one
        // IMAD.WIDE each before the partial sum is folded into the 128-bit accumulator (same exact sum, same REDC)
        const int fold = 1 << min(30, 64 - 2 * mm.bits);
        u64 s0a = 0, s1a = 0, s0b = 0, s1b = 0;
        int cnt = 0;
#pragma unroll 2
        for (int j = 0; j < beta; ++j) {
            const ulonglong2 d = *(const ulonglong2*)((j == own) ? c1z + (size_t)t * N + x : dz + ((size_t)j * W + t) * N + x);
            const ulonglong2 k0 = *(const ulonglong2*)(evk + (((size_t)j * 2 + 0) * nmod + ci) * N + x);
            const ulonglong2 k1 = *(const ulonglong2*)(evk + (((size_t)j * 2 + 1) * nmod + ci) * N + x);
            s0a += (u64)(uint32_t)d.x * (uint32_t)k0.x; s0b += (u64)(uint32_t)d.y * (uint32_t)k0.y;
            s1a += (u64)(uint32_t)d.x * (uint32_t)k1.x; s1b += (u64)(uint32_t)d.y * (uint32_t)k1.y;
            if (++cnt == fold) {
                l0a += s0a; h0a += (l0a < s0a); l0b += s0b; h0b += (l0b < s0b);
                l1a += s1a; h1a += (l1a < s1a); l1b += s1b; h1b += (l1b < s1b);
                s0a = s1a = s0b = s1b = 0; cnt = 0;
            }
        }
        l0a += s0a; h0a += (l0a < s0a); l0b += s0b; h0b += (l0b < s0b);
        l1a += s1a; h1a += (l1a < s1a); l1b += s1b; h1b += (l1b < s1b);
    } else {
#pragma unroll 2
        for (int j = 0; j < beta; ++j) {
            const ulonglong2 d = *(const ulonglong2*)((j == own) ? c1z + (size_t)t * N + x : dz + ((size_t)j * W + t) * N + x);
            const ulonglong2 k0 = *(const ulonglong2*)(evk + (((size_t)j * 2 + 0) * nmod + ci) * N + x);
            const ulonglong2 k1 = *(const ulonglong2*)(evk + (((size_t)j * 2 + 1) * nmod + ci) * N + x);
            mac128(h0a, l0a, d.x, k0.x, mm.q); mac128(h0b, l0b, d.y, k0.y, mm.q);
            mac128(h1a, l1a, d.x, k1.x, mm.q); mac128(h1b, l1b, d.y, k1.y, mm.q);
        }
    }
    ulonglong2 r0 = make_ulonglong2(mont_redc(h0a, l0a, mm.q, mm.qinv_neg), mont_redc(h0b, l0b, mm.q, mm.qinv_neg));
    ulonglong2 r1 = make_ulonglong2(mont_redc(h1a, l1a, mm.q, mm.qinv_neg), mont_redc(h1b, l1b, mm.q, mm.qinv_neg));
    if (ADD_MODE == 1) {
        if (t < L) {
            const ulonglong2 w = pmodq[t];
            const ulonglong2 e = *(const ulonglong2*)(addend + (size_t)z * addend_zstride + (size_t)t * N + x);
            r0.x = add_mod(r0.x, shoup_mul(e.x, w.x, w.y, mm.q), mm.q); r0.y = add_mod(r0.y, shoup_mul(e.y, w.x, w.y, mm.q), mm.q);
        }
    } else if (ADD_MODE == 2) {
        const ulonglong2 e = *(const ulonglong2*)(addend + (size_t)z * addend_zstride + (size_t)t * N + x);
        r0.x = add_mod(r0.x, e.x, mm.q); r0.y = add_mod(r0.y, e.y, mm.q);
    }
    const uint32_t* __restrict__ sc = m.scatter[z];
    u64* o = out + (size_t)z * out_zstride;
    if (sc) {
        const uint2 xo = __ldg((const uint2*)(sc + x));
        o[(size_t)t * N + xo.x] = r0.x; o[(size_t)t * N + xo.y] = r0.y;
        o[((size_t)W + t) * N + xo.x] = r1.x; o[((size_t)W + t) * N + xo.y] = r1.y;
    } else {
        *(ulonglong2*)(o + (size_t)t * N + x) = r0; *(ulonglong2*)(o + ((size_t)W + t) * N + x) = r1;
    }
}

// out[p][t][x] = sum_z in[z][p][t][x]  (mod q_t), QP layout with W limbs per polynomial
__global__ void __launch_bounds__(TPB)
sum_qp_kernel(const SumList lst, int n, u64* __restrict__ out, int N, int L, int W, int nQ, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z;
    const int x = (blockIdx.x * TPB + threadIdx.x) * 2;
    const u64 q = mc[t < L ? t : nQ + (t - L)].q;
    const size_t o = ((size_t)p * W + t) * N + x;
    ulonglong2 a = *(const ulonglong2*)(lst.p[0] + o);
    for (int z = 1; z < n; ++z) {
        const ulonglong2 b = *(const ulonglong2*)(lst.p[z] + o);
        a.x = add_mod(a.x, b.x, q); a.y = add_mod(a.y, b.y, q);
    }
    *(ulonglong2*)(out + o) = a;
}

// tmp[p][i][x] = ((last[p][x] + (q_l-1)/2) mod q_l) mod q_i - ((q_l-1)/2 mod q_i)        (coefficient domain)
__global__ void __launch_bounds__(TPB)
rescale_prep_kernel(const u64* __restrict__ last, u64* __restrict__ tmp, int level, int N,
                    const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int i = blockIdx.y, p = blockIdx.z;
    const int x = blockIdx.x * TPB + threadIdx.x;
    const u64 ql = mc[level].q, half = (ql - 1) >> 1;
    const ModConst m = mc[i];
    const u64 s = add_mod(last[(size_t)p * N + x], half, ql);
    const u64 u = barrett_red(0, s, m), hq = barrett_red(0, half, m);
    tmp[((size_t)p * level + i) * N + x] = sub_mod(u, hq, m.q);
}
}  // namespace

BaseConvPlan make_baseconv_plan(Context& c, const std::vector<int>& src, const std::vector<int>& dst,
                                const std::vector<int>& out_off) {
    OB_CHECK((int)src.size() <= MAX_NS, "too many source limbs for one base conversion");
    OB_CHECK(dst.size() == out_off.size(), "out_off size");
    BaseConvPlan p;
    p.bc = &c.base_conv(src, dst);
    std::vector<LimbDesc> tmp(out_off.size());            // destination offsets ride in LimbDesc::in_off
    for (size_t i = 0; i < out_off.size(); ++i) tmp[i] = LimbDesc{(uint32_t)out_off[i], 0u, 0u, 0xB0C0u};
    p.out_off = c.limb_list(tmp);
    return p;
}

void base_convert(Context& c, const BaseConvPlan& p, const u64* in, u64* out, bool centered) {
    const BaseConv& b = *p.bc;
    dim3 grid(c.N / TPB, (b.nd + DST_PER_CTA - 1) / DST_PER_CTA);
    if (centered)
        OB_LAUNCH((baseconv_kernel<true>), grid, TPB, 0, stream(), in, out, c.N, b.ns, b.nd, b.src_mod.as<int>(),
                  b.dst_mod.as<int>(), p.out_off, b.inv.as<ulonglong2>(), b.mat.u(), b.negvq.u(), b.half_src.u(),
                  b.half_dst.u(), c.dmc());
    else
        OB_LAUNCH((baseconv_kernel<false>), grid, TPB, 0, stream(), in, out, c.N, b.ns, b.nd, b.src_mod.as<int>(),
                  b.dst_mod.as<int>(), p.out_off, b.inv.as<ulonglong2>(), b.mat.u(), b.negvq.u(), b.half_src.u(),
                  b.half_dst.u(), c.dmc());
}

BaseConvBatch make_baseconv_batch(Context& c, const std::vector<BaseConvPlan>& plans, const std::vector<int>& in_off,
                                  const std::vector<int>& out_base) {
    BaseConvBatch b; b.n = (int)plans.size();
    std::vector<BcDev> h(plans.size());
    for (size_t z = 0; z < plans.size(); ++z) {
        const BaseConv& bc = *plans[z].bc;
        h[z] = BcDev{bc.ns, bc.nd, in_off[z], out_base[z], bc.src_mod.as<int>(), bc.dst_mod.as<int>(), plans[z].out_off,
                     bc.inv.as<ulonglong2>(), bc.mat.u(), bc.negvq.u(), bc.half_src.u(), bc.half_dst.u(),
                     bc.mat_d.as<double>(), bc.negvq_d.as<double>(), bc.f64_src ? 1 : 0, 0,
                     bc.tc_frag.p ? bc.tc_frag.as<uint2>() : nullptr};
        b.tc_ok = (z == 0 || b.tc_ok) && bc.tc_frag.p != nullptr;
        b.max_nd = std::max(b.max_nd, bc.nd);
        const size_t ns = bc.ns, nd = bc.nd;
        const size_t sm = 8 * (2 * (ns * nd + nd * (ns + 1)) + 4 * nd + 2 * ns + ns + (ns & 1)) + 16 * ns + 8 * nd + 16;
        b.smem = std::max(b.smem, sm);
        b.max_ns = std::max(b.max_ns, bc.ns);
        b.min_nd = std::min(b.min_nd, bc.nd);
        unsigned __int128 qsum = 0;
        bool small = true;
        for (int i : bc.h_src) { qsum += c.mod[i]; small = small && c.mod[i] < (1ull << 32); }
        static const bool w32 = !(std::getenv("ORION_B200_BC_W32") && std::atoi(std::getenv("ORION_B200_BC_W32")) == 0);
        const int mm = qsum <= ((unsigned __int128)1 << 64) ? (small ? 2 : 1) : 0;
        h[z].mm = mm | (w32 ? 4 : 0);
        b.mac_mode = std::max(b.mac_mode, mm);
    }
    capture_abort();
    auto buf = std::make_unique<DevBuf>(h.size() * sizeof(BcDev));
    OB_CUDA(cudaMemcpyAsync(buf->p, h.data(), h.size() * sizeof(BcDev), cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    b.dev = buf->as<BcDev>();
    c.misc.push_back(std::move(buf));
    return b;
}

void base_convert_batch(Context& c, const BaseConvBatch& b, const u64* in, u64* out, bool centered) {
    if (b.n == 0) return;
    if (b.tc_ok && b.max_ns >= 5) {                   // tensor-core path (ORION_B200_BC_TC=0 disables): batches with 5..8-source conversions
        static const int rt = std::getenv("ORION_B200_BC_TC_RT") ? std::atoi(std::getenv("ORION_B200_BC_TC_RT")) : 2;
        static const int zw = std::getenv("ORION_B200_BC_TC_Z") ? std::atoi(std::getenv("ORION_B200_BC_TC_Z")) : 6;
        const int rows = 16 * rt * (TC_TPB / 32);
        const int nblk = (b.max_nd + 3) / 4;
        if (c.N % rows == 0 && tc_smem_bytes(nblk) <= 48 * 1024) {
            const long ctas = (long)(c.N / rows) * b.n;
            const int zsplit = (int)std::min<long>(std::max<long>(1, (148L * zw + ctas - 1) / ctas), std::max(1, nblk / 2));
            dim3 grid(c.N / rows, b.n, zsplit);
            const size_t sm = tc_smem_bytes((nblk + zsplit - 1) / zsplit);
#define OB_TC(RTV) do { if (centered) OB_LAUNCH((baseconv_tc_kernel<RTV, true>), grid, TC_TPB, sm, stream(), in, out, c.N, b.dev, c.dmc()); \
                        else OB_LAUNCH((baseconv_tc_kernel<RTV, false>), grid, TC_TPB, sm, stream(), in, out, c.N, b.dev, c.dmc()); } while (0)
            if (rt == 1) OB_TC(1); else if (rt == 4) OB_TC(4); else OB_TC(2);
#undef OB_TC
            return;
        }
    }
    OB_CHECK(b.smem <= 48 * 1024, "base conversion tables exceed the static shared-memory window");
    // size the launch for >= ~2 CTAs of work per SM: one coefficient per thread and the destination limbs split over
    // gridDim.z when the ring is small, two coefficients per thread and no split when it is large
    const long want = 148L * 2048;
    const int vec = ((long)c.N / 2) * b.n >= want ? 2 : 1;
    const long thr = ((long)c.N / vec) * b.n;
    int nsplit = (int)std::min<long>(std::max<long>(1, (want + thr - 1) / thr), std::max(1, b.min_nd / 4));
    dim3 grid(c.N / (vec * BC_TPB), b.n, nsplit);
#define OB_BC3(C, NSV, MMV) do { if (vec == 2) OB_LAUNCH((baseconv_batch_kernel<C, NSV, 2, MMV>), grid, BC_TPB, b.smem, stream(), in, out, c.N, b.dev, c.dmc()); \
                            else OB_LAUNCH((baseconv_batch_kernel<C, NSV, 1, MMV>), grid, BC_TPB, b.smem, stream(), in, out, c.N, b.dev, c.dmc()); } while (0)
#define OB_BC2(C, NSV) do { if (b.mac_mode == 2) OB_BC3(C, NSV, 2); else if (b.mac_mode == 1 && NSV > 3) OB_BC3(C, NSV, 1); else OB_BC3(C, NSV, 0); } while (0)
#define OB_BC(NSV) do { if (centered) OB_BC2(true, NSV); else OB_BC2(false, NSV); } while (0)
    if (b.max_ns <= 1) OB_BC(1); else if (b.max_ns <= 2) OB_BC(2); else if (b.max_ns <= 3) OB_BC(3); else if (b.max_ns <= 4) OB_BC(4);
    else if (b.max_ns <= 6) OB_BC(6); else if (b.max_ns <= 8) OB_BC(8); else OB_BC(MAX_NS);
#undef OB_BC3
#undef OB_BC2
#undef OB_BC
}

void ks_inner_product(Context& c, int level, const u64* c1_ntt, const u64* digits, const u64* evk,
                      u64* acc, const uint32_t* gather, bool accumulate) {
    const int L = level + 1, K = c.nP, W = L + K, beta = c.num_digits(level);
    dim3 grid(c.N / TPB, W);
#define OB_IP(G, A) OB_LAUNCH((inner_product_kernel<G, A>), grid, TPB, 0, stream(), c1_ntt, digits, evk, acc, c.N, \
                              L, K, c.nQ, c.nmod, beta, gather, c.dmc())
    if (gather) { if (accumulate) OB_IP(true, true); else OB_IP(true, false); }
    else        { if (accumulate) OB_IP(false, true); else OB_IP(false, false); }
#undef OB_IP
}

void ks_inner_product_multi(Context& c, int level, int n, const u64* c1, size_t c1_zstride, const u64* digits, size_t dig_zstride,
                            const IpMulti& m, u64* out, size_t out_zstride, int add_mode, const u64* addend, size_t addend_zstride) {
    OB_CHECK(n >= 1 && n <= IPM_MAX, "inner product batch size");
    const int L = level + 1, K = c.nP, W = L + K, beta = c.num_digits(level);
    dim3 grid(c.N / (2 * TPB), W, n);
#define OB_IPM(MODE) OB_LAUNCH((inner_product_multi_kernel<MODE>), grid, TPB, 0, stream(), c1, c1_zstride, digits, dig_zstride, m, out, \
                               out_zstride, addend, addend_zstride, c.N, L, K, c.nQ, c.nmod, beta, c.d_pmodq.as<ulonglong2>(), c.dmc())
    if (add_mode == 1) OB_IPM(1); else if (add_mode == 2) OB_IPM(2); else OB_IPM(0);
#undef OB_IPM
}

void sum_qp(Context& c, int level, const SumList& lst, int n, u64* out) {
    OB_CHECK(n >= 1 && n <= SUM_MAX, "sum_qp count");
    const int L = level + 1, W = L + c.nP;
    dim3 grid(c.N / (2 * TPB), W, 2);
    OB_LAUNCH(sum_qp_kernel, grid, TPB, 0, stream(), lst, n, out, c.N, L, W, c.nQ, c.dmc());
}

void rescale_prep(Context& c, int level, const u64* last_coeff, u64* tmp, int polys) {
    dim3 grid(c.N / TPB, level, polys);
    OB_LAUNCH(rescale_prep_kernel, grid, TPB, 0, stream(), last_coeff, tmp, level, c.N, c.dmc());
}

}  // namespace ob
