// context.h -- CKKS parameter set resident in HBM: modulus chain, NTT twiddles (Shoup pairs, bit-reversed),
// base-conversion matrices, Galois index tables.  Replaces the state behind LG/scheme.go:16-86
// (ckks.NewParametersFromLiteral) and HEonGPU's HEContext (HE/src/heongpu/lib/host/ckks/context.cu).
#pragma once
#include "common.h"

namespace ob {

// host number theory (context.cu)
bool is_prime_u64(u64 n);
u64 powmod_u64(u64 a, u64 e, u64 q);
u64 invmod_u64(u64 a, u64 q);
u64 mulmod_u64(u64 a, u64 b, u64 q);
u64 shoup_companion(u64 w, u64 q);
std::vector<u64> gen_ntt_primes(int bits, u64 nthroot, int count, const std::vector<u64>& skip);
void gen_moduli(const std::vector<int>& logQ, const std::vector<int>& logP, int log_nthroot,
                std::vector<u64>& Q, std::vector<u64>& P);
u64 primitive_root(u64 q);
u64 find_psi(u64 q, u64 nthroot);
ModConst make_modconst(u64 q);

// RNS fast base conversion tables for one (source set -> destination set) pair, in HBM
struct BaseConv {
    int ns = 0, nd = 0;
    DevBuf src_mod, dst_mod;      // int[ns], int[nd]    chain indices
    DevBuf inv;                   // u64[ns][2]          (Q_S/q_i)^-1 mod q_i with Shoup companion
    DevBuf mat;                   // u64[ns][nd]         (Q_S/q_i) * 2^64 mod p_t   (Montgomery form)
    DevBuf negvq;                 // u64[nd][ns+1]       (-v * Q_S) mod p_t for the overflow count v
    DevBuf half_src, half_dst;    // u64[ns], u64[nd]    floor(Q_S/2) mod q_i / mod p_t
    DevBuf mat_d, negvq_d;        // double[ns][nd], double[nd][ns+1]: the same constants in standard form, as exact
                                  // doubles (FP64 contraction for destinations of at most 46 bits)
    bool f64_src = false;         // every source prime has at most 46 bits
    DevBuf tc_frag;               // uint2[ceil(nd/4)][4][2][32]: INT8 tensor-core contraction (ks_kernels.cu baseconv_tc_kernel), the
                                  // matrix as per-lane B fragments of mma.m16n8k32.u8; for ns <= 8 unless ORION_B200_BC_TC=0
    std::vector<int> h_src, h_dst;
};

struct Context {
    int logN = 0, N = 0, nQ = 0, nP = 0, nmod = 0;
    std::vector<u64> mod, psi;
    std::vector<ModConst> mc;

    DevBuf d_mc;        // ModConst[nmod]
    DevBuf d_tw_fwd;    // ulonglong2[nmod][N]  (psi^bitrev(i), shoup)
    DevBuf d_tw_inv;    // ulonglong2[nmod][N]  (psi^-bitrev(i), shoup)
    DevBuf d_twd_fwd;   // double[nmod][N]      the same twiddles as exact doubles (FP64 butterfly path, primes <= 46 bits)
    DevBuf d_twd_inv;   // double[nmod][N]
    DevBuf d_ninv;      // ulonglong2[nmod]     (N^-1, shoup)
    DevBuf d_pinv;      // ulonglong2[nQ]       (P^-1 mod q_i, shoup)          ModDown
    DevBuf d_qlinv;     // ulonglong2[nQ][nQ]   [l][i] = q_l^-1 mod q_i, shoup  rescale
    DevBuf d_pmodq;     // ulonglong2[nQ]       (P mod q_i, shoup)              key generation

    Context(int logN, const std::vector<u64>& Q, const std::vector<u64>& P);
    ~Context();

    int num_digits(int level) const { return nP ? (level + 1 + nP - 1) / nP : 0; }
    int chain_index_qp(int level, int t) const { return t <= level ? t : nQ + (t - level - 1); }
    const ModConst* dmc() const { return d_mc.as<ModConst>(); }
    const ulonglong2* tw_fwd() const { return d_tw_fwd.as<ulonglong2>(); }
    const ulonglong2* tw_inv() const { return d_tw_inv.as<ulonglong2>(); }
    const double* twd_fwd() const { return d_twd_fwd.as<double>(); }
    const double* twd_inv() const { return d_twd_inv.as<double>(); }

    // cached device-side descriptors
    const LimbDesc* limb_list(const std::vector<LimbDesc>& v);
    const BaseConv& base_conv(const std::vector<int>& src, const std::vector<int>& dst);
    const uint32_t* galois_index(u64 galEl);          // NTT-domain gather table, out[i] = in[idx[i]]
    const uint32_t* galois_scatter(u64 galEl);        // inverse permutation: out[sc[i]] = in[i]
    u64 galois_element(int rot) const;                // 5^rot mod 2N
    u64 galois_conj() const { return 2ull * N - 1; }

    std::vector<std::unique_ptr<DevBuf>> misc;        // small device tables owned by the context (launch descriptors)

  private:
    std::map<std::vector<uint32_t>, std::unique_ptr<DevBuf>> limb_cache_;
    std::map<std::vector<int>, std::unique_ptr<BaseConv>> bc_cache_;
    std::map<u64, std::unique_ptr<DevBuf>> gal_cache_;
};


// ---- batched NTT (ntt.cu) ----
struct NttPlan { const LimbDesc* first = nullptr; const LimbDesc* second = nullptr; int n = 0; };
NttPlan make_ntt_plan(Context& c, const std::vector<LimbDesc>& limbs);
void ntt_forward(Context& c, const u64* in, u64* out, const NttPlan& p);
void ntt_forward_first(Context& c, const u64* in, u64* out, const NttPlan& p);   // strided pass only (lazy form), see ks_ntt2_inner_product
void ntt_forward_second(Context& c, u64* buf, const NttPlan& p);                  // contiguous pass only, in place
void ntt_inverse(Context& c, const u64* in, u64* out, const NttPlan& p);
void ntt_inverse_first(Context& c, const u64* in, u64* out, const NttPlan& p);   // contiguous pass only (lazy form), see intt2_bc_prep

}  // namespace ob
