// kernels.h -- launchers for the non-NTT kernels of the CKKS hot path (kernels.cu).
// "QP layout at level l": limbs Q_0..Q_l followed by P_0..P_{nP-1}; W = l+1+nP limbs per polynomial.
#pragma once
#include "context.h"

namespace ob {

enum EwOp { EW_ADD = 0, EW_SUB = 1, EW_NEG = 2, EW_MUL = 3, EW_COPY = 4 };

// out[p][t][x] = a[p][t][x] op b[(p % b_polys)][t][x];  polys x limbs x N words, limb t uses chain index
// (t < L ? t : nQ + t - L).  a/b/out strides are given in limbs so operands at different levels can be mixed.
void ew_binary(Context& c, EwOp op, const u64* a, int a_stride, const u64* b, int b_stride, int b_polys,
               u64* out, int out_stride, int polys, int limbs, int L);
// out = a * s_t  /  a + s_t   with one scalar per limb (host vector, residues < q_t); scalars[t]
void ew_mul_scalar(Context& c, const u64* a, int a_stride, u64* out, int out_stride, int polys, int limbs, int L,
                   const std::vector<u64>& scalars);
void ew_add_scalar(Context& c, const u64* a, int a_stride, u64* out, int out_stride, int polys, int limbs, int L,
                   const std::vector<u64>& scalars, int only_poly);
// degree-2 tensor product: (d0,d1,d2) = (a0*b0, a0*b1 + a1*b0, a1*b1), L limbs each (chain index = limb)
void tensor_product(Context& c, const u64* a, int a_stride, const u64* b, int b_stride, u64* out, int L);
// out[p][t] = sum_{i<n} k[i][t] * src_i[p][t] (+ c0[t] on p = 0) over the first L limbs of the Q chain; src_i has `stride[i]`
// limbs per polynomial (>= L), out has L.  One launch (kernels.cu lincomb_kernel).
constexpr int LINCOMB_MAX = 16, LINCOMB_LIMBS = 96;
void lin_comb(Context& c, int n, const u64* const* src, const int* stride, const std::vector<std::vector<u64>>& k,
              const std::vector<u64>* c0, u64* out, int polys, int L);
// NTT-domain automorphism: out[p][t][i] = in[p][t][idx[i]]
void automorphism_ntt(Context& c, const u64* in, u64* out, int total_limbs, const uint32_t* idx);
// to / from Montgomery form (x * 2^64 mod q), limbs laid out like ew_binary
void to_montgomery(Context& c, const u64* in, u64* out, int polys, int limbs, int L);
void from_montgomery(Context& c, const u64* in, u64* out, int polys, int limbs, int L);

// ---- base conversion: in [ns][N] (coefficient domain, contiguous), out limb d written at out + out_off[d]*N ----
// A non-centred batched conversion (ModUp) whose sources and destination all run on the FP64 path leaves that
// destination limb as the UNREDUCED sum of centred residues in double form (|value| <= (1 + 0.7 ns) q): the forward NTT
// that always follows reads it as its lazy input (LimbDesc::aux bit 0) -- saves the final reduction here and the
// word->double conversion there.  ns <= 4 keeps the NTT's growth bound (ntt.cu) intact.
__host__ __device__ inline bool bc_raw_out(bool f64_src, int ns, int dst_bits) { return f64_src && ns <= 4 && dst_bits <= 46; }
struct BaseConvPlan { const BaseConv* bc = nullptr; const LimbDesc* out_off = nullptr; };
BaseConvPlan make_baseconv_plan(Context& c, const std::vector<int>& src, const std::vector<int>& dst,
                                const std::vector<int>& out_off);
void base_convert(Context& c, const BaseConvPlan& p, const u64* in, u64* out, bool centered);
// several conversions in one launch (all digits of a ModUp, both polynomials of a ModDown): conversion z reads its
// sources at in + (in_off[z] + i)*N and writes destination d at out + (out_base[z] + plan.out_off[d])*N.
struct BcDev {
    int ns, nd, in_off, out_base;
    const int* src_mod; const int* dst_mod; const LimbDesc* out_off;
    const ulonglong2* inv; const u64* mat; const u64* negvq; const u64* half_src; const u64* half_dst;
    const double* mat_d; const double* negvq_d; int f64_src;
    int mm;                       // integer-path accumulation mode of this conversion (ks_kernels.cu: 0 / 1 raw / 2 96-bit)
    const uint2* tc_frag;         // tensor-core trial: B fragments (context.cu base_conv), nullptr when not built
};
struct BaseConvBatch { const BcDev* dev = nullptr; int n = 0; size_t smem = 0; int max_ns = 0; int min_nd = 1 << 30; int mac_mode = 0;
                       bool tc_ok = false; int max_nd = 0; };
BaseConvBatch make_baseconv_batch(Context& c, const std::vector<BaseConvPlan>& plans, const std::vector<int>& in_off,
                                  const std::vector<int>& out_base);
void base_convert_batch(Context& c, const BaseConvBatch& b, const u64* in, u64* out, bool centered);

// ---- base conversion fused into the transforms either side of it (ntt.cu: intt2_prep_kernel, ntt1_bc_kernel) ----
constexpr int BC_FUSED_MAX_NS = 4;
struct BcLimbDev { uint32_t out_off, mod, conv, d; };       // destination limb: output offset (units of N), chain index, conversion, index in it
void intt2_bc_prep(Context& c, u64* buf, unsigned char* vout, const BcDev* descs, int n_conv, bool centered);
void ntt1_bc(Context& c, const u64* ybuf, const unsigned char* vbuf, u64* out, const BcLimbDev* limbs, int n_limbs, const BcDev* descs,
             bool centered);

// ---- key-switch inner product ----
// acc[c][t][x] = sum_j digit_j[t][x] * evk[j][c][chain(t)][x]  (evk in Montgomery form, full-chain layout).
// digit_j limb t is read from `digits + (j*W + t)*N` unless t belongs to digit j itself, in which case it is
// read from c1_ntt + t*N.  gather != nullptr applies an NTT-domain automorphism to the digits on the fly.
void ks_inner_product(Context& c, int level, const u64* c1_ntt, const u64* digits, const u64* evk,
                      u64* acc /* [2][W][N] */, const uint32_t* gather, bool accumulate);

// n independent inner products in one launch (blockIdx.z): product z uses c1 + z*c1_zstride, digits + z*dig_zstride,
// key m.evk[z]; adds the c0 contribution (add_mode 1: P * addend_z over the Q limbs, 2: addend_z over all W limbs, 0: none)
// to polynomial 0 and writes through the scatter table m.scatter[z] (NTT-domain automorphism; nullptr = identity).
constexpr int IPM_MAX = 48;
struct IpMulti { const u64* evk[IPM_MAX]; const uint32_t* scatter[IPM_MAX]; };
void ks_inner_product_multi(Context& c, int level, int n, const u64* c1, size_t c1_zstride, const u64* digits, size_t dig_zstride,
                            const IpMulti& m, u64* out, size_t out_zstride, int add_mode, const u64* addend, size_t addend_zstride);
// Same contraction with the second (contiguous) NTT pass of the digits fused in: `digits` holds the output of
// ntt_forward_first.  Plain accumulate only (no addend, no automorphism).  ntt.cu.
void ks_ntt2_inner_product(Context& c, int level, int n, const u64* c1, size_t c1_zstride, const u64* digits, size_t dig_zstride,
                           const IpMulti& m, u64* out, size_t out_zstride);
// Second NTT pass fused with the ModDown / rescale epilogue (ntt.cu): out = (a - NTT(b)) * mulc (+ addend).
void ntt2_submul(Context& c, int nl, int polys, const u64* a, int a_stride, const u64* b, const ulonglong2* mulc,
                 const u64* addend, int addend_stride, int addend_polys, int addend_period, int addend_zstride,
                 u64* out, int out_stride);
// out[2][W][N] = sum of n QP ciphertexts
constexpr int SUM_MAX = 64;
struct SumList { const u64* p[SUM_MAX]; };
void sum_qp(Context& c, int level, const SumList& lst, int n, u64* out);

// rescale helpers (level -> level-1): see evaluator.cu
void rescale_prep(Context& c, int level, const u64* last_coeff /* [polys][N] */, u64* tmp /* [polys][level][N] */, int polys);
// the same correction term, computed inside the first forward NTT pass of its `level` limbs (ntt.cu ntt1_rescale_kernel): tmp holds
// what rescale_prep + ntt_forward_first leave there
void ntt1_rescale(Context& c, int level, int polys, const u64* last_coeff, u64* tmp);

}  // namespace ob
