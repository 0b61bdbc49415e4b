// kernels.cu -- streaming RNS kernels of the CKKS hot path for sm_100a: element-wise limb arithmetic,
// NTT-domain automorphism, fast base conversion (ModUp / ModDown), the evaluation-key inner product,
// ModDown / rescale epilogues.  All are HBM-bound integer kernels: one or two 64-bit coefficients per
// thread, 16-byte vector accesses where the layout allows, grids of (N/threads-per-block) x limbs.
// Reference kernels these replace (different algorithms, cited for the row they cover):
//   addition.cu:10-36, multiplication.cu:102,313,333, switchkey.cu:179,618,811,1366 under
//   HE/src/heongpu/lib/kernel/.
#include "kernels.h"

namespace ob {

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ int chain_of(int t, int L, int nQ) { return t < L ? t : nQ + (t - L); }

// ------------------------------------------------------------------------------------------------
// element-wise; two coefficients per thread (ulonglong2)
// ------------------------------------------------------------------------------------------------
template <int OP>
__global__ void __launch_bounds__(TPB)
ew_binary_kernel(const u64* __restrict__ a, int a_stride, const u64* __restrict__ b, int b_stride, int b_polys,
                 u64* __restrict__ out, int out_stride, int limbs, int L, int nQ, int N,
                 const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z;
    const int x = (blockIdx.x * TPB + threadIdx.x) * 2;
    if (x >= N) return;
    const ModConst m = mc[chain_of(t, L, nQ)];
    const ulonglong2 va = *(const ulonglong2*)(a + ((size_t)p * a_stride + t) * N + x);
    ulonglong2 vb = make_ulonglong2(0, 0);
    if (OP == EW_ADD || OP == EW_SUB || OP == EW_MUL)
        vb = *(const ulonglong2*)(b + ((size_t)(p % b_polys) * b_stride + t) * N + x);
    ulonglong2 r;
    if (OP == EW_ADD) { r.x = add_mod(va.x, vb.x, m.q); r.y = add_mod(va.y, vb.y, m.q); }
    else if (OP == EW_SUB) { r.x = sub_mod(va.x, vb.x, m.q); r.y = sub_mod(va.y, vb.y, m.q); }
    else if (OP == EW_NEG) { r.x = va.x ? m.q - va.x : 0; r.y = va.y ? m.q - va.y : 0; }
    else if (OP == EW_MUL) { r.x = barrett_mul(va.x, vb.x, m); r.y = barrett_mul(va.y, vb.y, m); }
    else r = va;
    *(ulonglong2*)(out + ((size_t)p * out_stride + t) * N + x) = r;
}

// MODE 0: out = a * s (Shoup);  MODE 1: out = a + s (only on poly `only_poly`, copy elsewhere).
// The per-limb scalars travel BY VALUE in the kernel parameters (up to 96 limbs = 1.5 KB of the 4 KB parameter space):
// no staging copy and no allocation per call -- at logN = 13 a polynomial evaluation issues thousands of these and the
// pageable copy was a visible share of the host-side enqueue time.
constexpr int SCAL_MAX = 96;
struct ScalarTab { ulonglong2 v[SCAL_MAX]; };
template <int MODE>
__global__ void __launch_bounds__(TPB)
ew_scalar_kernel(const u64* __restrict__ a, int a_stride, u64* __restrict__ out, int out_stride, int L, int nQ, int N,
                 const ModConst* __restrict__ mc, const __grid_constant__ ScalarTab tab, const ulonglong2* __restrict__ scal,
                 int only_poly) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z;
    const int x = (blockIdx.x * TPB + threadIdx.x) * 2;
    if (x >= N) return;
    const u64 q = mc[chain_of(t, L, nQ)].q;
    const ulonglong2 s = scal ? scal[t] : tab.v[t];
    ulonglong2 v = *(const ulonglong2*)(a + ((size_t)p * a_stride + t) * N + x);
    if (MODE == 0) { v.x = shoup_mul(v.x, s.x, s.y, q); v.y = shoup_mul(v.y, s.x, s.y, q); }
    else if (p == only_poly) { v.x = add_mod(v.x, s.x, q); v.y = add_mod(v.y, s.x, q); }
    *(ulonglong2*)(out + ((size_t)p * out_stride + t) * N + x) = v;
}

// out[p][t][x] = sum_i k_i[t] * src_i[p][t][x]  (+ c0[t] on polynomial 0): the leaves of the polynomial evaluator
// (sum of constant multiples of power-basis ciphertexts) in ONE launch instead of a multiply, an add and, for
// operands that sit at a higher level, a copy per term.  Operands are read in place with their own limb stride, so a
// higher-level operand is "dropped" by simply not reading its upper limbs.  Residues travel in the kernel parameters.
struct LinCombArgs {
    const u64* src[LINCOMB_MAX];
    int stride[LINCOMB_MAX];
    u64 k[(LINCOMB_MAX + 1) * LINCOMB_LIMBS];             // [term][limb]; row n = additive constant
};
__global__ void __launch_bounds__(TPB)
lincomb_kernel(const __grid_constant__ LinCombArgs a, int n, int has_c0, u64* __restrict__ out, int L, int N,
               const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z;
    const int x = (blockIdx.x * TPB + threadIdx.x) * 2;
    if (x >= N) return;
    const ModConst m = mc[t];
    ulonglong2 acc = make_ulonglong2(0, 0);
    for (int i = 0; i < n; ++i) {
        const ulonglong2 v = *(const ulonglong2*)(a.src[i] + ((size_t)p * a.stride[i] + t) * N + x);
        const u64 k = a.k[i * LINCOMB_LIMBS + t];
        acc.x = add_mod(acc.x, barrett_mul(v.x, k, m), m.q);
        acc.y = add_mod(acc.y, barrett_mul(v.y, k, m), m.q);
    }
    if (has_c0 && p == 0) {
        const u64 c = a.k[n * LINCOMB_LIMBS + t];
        acc.x = add_mod(acc.x, c, m.q); acc.y = add_mod(acc.y, c, m.q);
    }
    *(ulonglong2*)(out + ((size_t)p * L + t) * N + x) = acc;
}

__global__ void __launch_bounds__(TPB)
tensor_kernel(const u64* __restrict__ a, int a_stride, const u64* __restrict__ b, int b_stride,
              u64* __restrict__ out, int L, int N, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y;
    const int x = blockIdx.x * TPB + threadIdx.x;
    const ModConst m = mc[t];
    const u64 a0 = a[(size_t)t * N + x], a1 = a[((size_t)a_stride + t) * N + x];
    const u64 b0 = b[(size_t)t * N + x], b1 = b[((size_t)b_stride + t) * N + x];
    u64 hi = 0, lo = 0;
    out[(size_t)t * N + x] = barrett_mul(a0, b0, m);
    // a0*b1 + a1*b0 < 2q^2 < 2^(2*bits+1) <= 2^(bits+62): one Barrett reduction for the sum
    u64 p0l = a0 * b1, p0h = mulhi64(a0, b1), p1l = a1 * b0, p1h = mulhi64(a1, b0);
    lo = p0l + p1l; hi = p0h + p1h + (lo < p0l);
    out[((size_t)L + t) * N + x] = barrett_red(hi, lo, m);
    out[((size_t)2 * L + t) * N + x] = barrett_mul(a1, b1, m);
}

__global__ void __launch_bounds__(TPB)
automorphism_kernel(const u64* __restrict__ in, u64* __restrict__ out, int N, const uint32_t* __restrict__ idx) {
    OB_PDL_SYNC();
    const int x = blockIdx.x * TPB + threadIdx.x;
    const size_t base = (size_t)blockIdx.y * N;
    out[base + x] = in[base + __ldg(&idx[x])];
}

template <bool TO>
__global__ void __launch_bounds__(TPB)
mont_kernel(const u64* __restrict__ in, u64* __restrict__ out, int limbs, int L, int nQ, int N,
            const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z;
    const int x = blockIdx.x * TPB + threadIdx.x;
    const ModConst m = mc[chain_of(t, L, nQ)];
    const size_t o = ((size_t)p * limbs + t) * N + x;
    out[o] = TO ? shoup_mul(in[o], m.r1, m.r1_sh, m.q) : mont_redc(0, in[o], m.q, m.qinv_neg);
}

}  // namespace

void ew_binary(Context& c, EwOp op, const u64* a, int a_stride, const u64* b, int b_stride, int b_polys,
               u64* out, int out_stride, int polys, int limbs, int L) {
    dim3 grid(c.N / (2 * TPB), limbs, polys);
    if (op != EW_COPY) alg_bytes(8.0 * c.N * limbs * polys * (op == EW_NEG ? 2.0 : 3.0));       // operands in, result out
#define OB_EW(OPV) OB_LAUNCH((ew_binary_kernel<OPV>), grid, TPB, 0, stream(), a, a_stride, b, b_stride, \
                             b_polys < 1 ? 1 : b_polys, out, out_stride, limbs, L, c.nQ, c.N, c.dmc())
    switch (op) {
        case EW_ADD: OB_EW(EW_ADD); break;
        case EW_SUB: OB_EW(EW_SUB); break;
        case EW_NEG: OB_EW(EW_NEG); break;
        case EW_MUL: OB_EW(EW_MUL); break;
        default: OB_EW(EW_COPY); break;
    }
#undef OB_EW
}

// scalars -> (value, Shoup companion) per limb, in the parameter table when they fit, else in a device buffer
static const ulonglong2* stage_scalars(Context& c, int limbs, int L, const std::vector<u64>& scalars, ScalarTab& tab, DevBuf& buf) {
    std::vector<ulonglong2> big;
    ulonglong2* h = tab.v;
    if (limbs > SCAL_MAX) { big.resize(limbs); h = big.data(); }
    for (int t = 0; t < limbs; ++t) {
        u64 q = c.mod[t < L ? t : c.nQ + (t - L)];
        u64 s = scalars[t] % q;
        h[t] = make_ulonglong2(s, shoup_companion(s, q));
    }
    if (limbs <= SCAL_MAX) return nullptr;
    capture_abort();                                          // a synchronising upload: not capturable
    buf.alloc(limbs * sizeof(ulonglong2));
    OB_CUDA(cudaMemcpyAsync(buf.p, h, limbs * sizeof(ulonglong2), cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));               // `big` dies at return
    return buf.as<ulonglong2>();
}

void ew_mul_scalar(Context& c, const u64* a, int a_stride, u64* out, int out_stride, int polys, int limbs, int L,
                   const std::vector<u64>& scalars) {
    ScalarTab tab; DevBuf buf; const ulonglong2* s = stage_scalars(c, limbs, L, scalars, tab, buf);
    dim3 grid(c.N / (2 * TPB), limbs, polys);
    alg_bytes(16.0 * c.N * limbs * polys);
    OB_LAUNCH((ew_scalar_kernel<0>), grid, TPB, 0, stream(), a, a_stride, out, out_stride, L, c.nQ, c.N, c.dmc(), tab, s, 0);
}
void ew_add_scalar(Context& c, const u64* a, int a_stride, u64* out, int out_stride, int polys, int limbs, int L,
                   const std::vector<u64>& scalars, int only_poly) {
    ScalarTab tab; DevBuf buf; const ulonglong2* s = stage_scalars(c, limbs, L, scalars, tab, buf);
    dim3 grid(c.N / (2 * TPB), limbs, polys);
    alg_bytes(16.0 * c.N * limbs * (only_poly >= 0 ? 1 : polys));
    OB_LAUNCH((ew_scalar_kernel<1>), grid, TPB, 0, stream(), a, a_stride, out, out_stride, L, c.nQ, c.N, c.dmc(), tab, s, only_poly);
}

void lin_comb(Context& c, int n, const u64* const* src, const int* stride, const std::vector<std::vector<u64>>& k,
              const std::vector<u64>* c0, u64* out, int polys, int L) {
    OB_CHECK(n >= 1 && n <= LINCOMB_MAX && L <= LINCOMB_LIMBS, "lin_comb: too many terms or limbs");
    LinCombArgs a;
    for (int i = 0; i < n; ++i) {
        a.src[i] = src[i]; a.stride[i] = stride[i];
        for (int t = 0; t < L; ++t) a.k[i * LINCOMB_LIMBS + t] = k[i][t];
    }
    if (c0) for (int t = 0; t < L; ++t) a.k[n * LINCOMB_LIMBS + t] = (*c0)[t];
    dim3 grid(c.N / (2 * TPB), L, polys);
    alg_bytes(8.0 * c.N * L * polys * (n + 1.0));
    OB_LAUNCH(lincomb_kernel, grid, TPB, 0, stream(), a, n, c0 ? 1 : 0, out, L, c.N, c.dmc());
}

void tensor_product(Context& c, const u64* a, int a_stride, const u64* b, int b_stride, u64* out, int L) {
    dim3 grid(c.N / TPB, L);
    alg_bytes(8.0 * c.N * L * 7.0);                       // two degree-1 operands in, one degree-2 result out
    OB_LAUNCH(tensor_kernel, grid, TPB, 0, stream(), a, a_stride, b, b_stride, out, L, c.N, c.dmc());
}

void automorphism_ntt(Context& c, const u64* in, u64* out, int total_limbs, const uint32_t* idx) {
    OB_CHECK(in != out, "automorphism cannot run in place");
    dim3 grid(c.N / TPB, total_limbs);
    alg_bytes(16.0 * c.N * total_limbs);
    OB_LAUNCH(automorphism_kernel, grid, TPB, 0, stream(), in, out, c.N, idx);
}

void to_montgomery(Context& c, const u64* in, u64* out, int polys, int limbs, int L) {
    dim3 grid(c.N / TPB, limbs, polys);
    OB_LAUNCH((mont_kernel<true>), grid, TPB, 0, stream(), in, out, limbs, L, c.nQ, c.N, c.dmc());
}
void from_montgomery(Context& c, const u64* in, u64* out, int polys, int limbs, int L) {
    dim3 grid(c.N / TPB, limbs, polys);
    OB_LAUNCH((mont_kernel<false>), grid, TPB, 0, stream(), in, out, limbs, L, c.nQ, c.N, c.dmc());
}

}  // namespace ob
