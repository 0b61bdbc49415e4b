// evaluator.h -- device-resident polynomial-level operations built from the kernels: hybrid key switching
// (decompose / inner product / ModDown), rescale.  These are the building blocks behind
// LG/evaluator.go:61-99, :297-317 and LG/lineartransform.go:96-113.
#pragma once
#include "kernels.h"

namespace ob {

struct KsPlan {                       // everything a key switch at one level needs, built once
    int level = -1, L = 0, K = 0, W = 0, beta = 0;
    NttPlan intt_c1;                  // c1 (L limbs) -> coefficient scratch
    std::vector<BaseConvPlan> modup;  // per digit
    std::vector<int> digit_src_off;   // first source limb of each digit
    NttPlan ntt_digits;               // all converted (non-own) limbs of the digit buffer, in place
    NttPlan intt_p[3];                // [polys]: P limbs of acc -> scratch [polys][K][N]
    BaseConvPlan moddown;             // P -> Q_0..Q_level (centered), one polynomial
    BaseConvBatch modup_all;          // every digit in one launch
    BaseConvBatch moddown_b[3];       // [polys]
    NttPlan ntt_conv[3];              // [polys]: conv [polys][L][N] in place
};

class PolyEvaluator {
  public:
    explicit PolyEvaluator(Context& c) : ctx(c) {}
    Context& ctx;

    const KsPlan& ks_plan(int level);

    // digits: [beta][W][N]; own-digit limbs are left untouched (they are read from c1 by the inner product)
    void decompose(int level, const u64* c1_ntt, u64* digits);
    // acc [polys][W][N] (QP, NTT) -> out [polys][L][N] = round(acc / P) (+ addend)
    void moddown(int level, const u64* acc, int polys, const u64* addend, int addend_stride, int addend_polys,
                 u64* out, int out_stride);
    // full key switch of c1 with evk: out = (out0, out1) [2][L][N], optionally + addend (c0 on poly 0)
    void keyswitch(int level, const u64* c1_ntt, const u64* evk, const u64* addend, int addend_stride,
                   int addend_polys, u64* out, int out_stride);
    // in [polys][level+1][N] (stride in_stride) -> out [polys][level][N]
    void rescale(int level, const u64* in, int in_stride, u64* out, int out_stride, int polys);

    // ---- batched variants: n independent polynomials in one set of launches (hoisted BSGS giant steps) ----
    // c1s: polynomial z at c1s + z*c1_stride*N ([L][N]);  digits: [n][beta][W][N]
    void decompose_n(int level, const u64* c1s, int c1_stride, int n, u64* digits, bool second_pass = true);
    // acc: polynomial z at acc + z*acc_stride*N ([W][N] QP);  out: polynomial z at out + z*out_stride*N ([L][N])
    void moddown_n(int level, const u64* acc, int acc_stride, int n, const u64* addend, int addend_stride, int addend_polys,
                   u64* out, int out_stride, int addend_period = 0, int addend_zstride = 0);
    // n independent key switches in one launch sequence: c1 of ciphertext z at c1s + z*c1_zstride*N, key evks[z],
    // optional addend (c0) at addend + z*addend_zstride*N; out: [n][2][L][N] contiguous
    void keyswitch_n(int level, int n, const u64* c1s, int c1_zstride, const u64* const* evks, const u64* addend,
                     int addend_zstride, u64* out);

    size_t digits_words(int level) const { return (size_t)ctx.num_digits(level) * (level + 1 + ctx.nP) * ctx.N; }

  private:
    std::map<int, std::unique_ptr<KsPlan>> plans_;
    std::map<int, NttPlan> rescale_intt_[3], rescale_ntt_[3];
    // fl / n_fl: destination-limb list of the fused conversion + first NTT pass (ntt1_bc); fusable: every conversion has at
    // most BC_FUSED_MAX_NS source limbs
    struct DecompNPlan { NttPlan intt, ntt; BaseConvBatch bc; const BcLimbDev* fl = nullptr; int n_fl = 0; bool fusable = false; };
    struct ModDownNPlan { NttPlan intt, ntt; BaseConvBatch bc; const BcLimbDev* fl = nullptr; int n_fl = 0; bool fusable = false; };
    const BcLimbDev* upload_bc_limbs(const std::vector<BcLimbDev>& v);
    std::map<std::vector<int>, DecompNPlan> decomp_n_;
    std::map<std::vector<int>, ModDownNPlan> moddown_n_;
};

}  // namespace ob
