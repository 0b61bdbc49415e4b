// evaluator.cu -- hybrid key switching and rescale as fixed sequences of kernels on one stream.
//
// Key switch at level l (L = l+1 limbs, K special primes, beta = ceil(L/K) digits, W = L+K):
//   1. INTT c1 (L limbs)                                        -> coef  [L][N]
//   2. per digit j: base-convert Q_digit -> (Q \ digit) u P      -> digits[j][t] for t not in digit j
//   3. NTT of the beta*W - L converted limbs (one batched launch pair)
//   4. inner product with the evaluation key (own-digit limbs read straight from c1)   -> acc [2][W][N]
//   5. INTT of the 2K P-limbs, centred base conversion P -> Q, NTT of 2L limbs, (acc_Q - conv) * P^-1 (+ c0)
// Limb-NTT count beta*W + L + 2K  (= SURVEY.md 8(d): beta(L+k) + 2L + 2k with the L own limbs never transformed).
// Data flow of the reference GPU path for the same op: HE/src/heongpu/lib/host/ckks/operator.cu:1308-1429.
#include "evaluator.h"

namespace ob {

const KsPlan& PolyEvaluator::ks_plan(int level) {
    auto it = plans_.find(level);
    if (it != plans_.end()) return *it->second;
    Context& c = ctx;
    OB_CHECK(c.nP > 0, "key switching needs at least one special prime");
    OB_CHECK(level >= 0 && level < c.nQ, "level out of range");
    auto p = std::make_unique<KsPlan>();
    p->level = level; p->L = level + 1; p->K = c.nP; p->W = p->L + p->K; p->beta = c.num_digits(level);
    const int L = p->L, K = p->K, W = p->W, beta = p->beta;
    std::vector<LimbDesc> l;
    for (int i = 0; i < L; ++i) l.push_back(LimbDesc{(uint32_t)i, (uint32_t)i, (uint32_t)i, 0});
    p->intt_c1 = make_ntt_plan(c, l);
    std::vector<LimbDesc> dl;
    for (int j = 0; j < beta; ++j) {
        int s0 = j * K, s1 = std::min((j + 1) * K, L);
        std::vector<int> src, dst, off;
        for (int i = s0; i < s1; ++i) src.push_back(i);
        for (int t = 0; t < W; ++t) {
            if (t >= s0 && t < s1) continue;
            int ci = t < L ? t : c.nQ + (t - L);
            dst.push_back(ci); off.push_back(j * W + t);
            dl.push_back(LimbDesc{(uint32_t)(j * W + t), (uint32_t)(j * W + t), (uint32_t)ci, 0});
        }
        p->modup.push_back(make_baseconv_plan(c, src, dst, off));
        p->digit_src_off.push_back(s0);
    }
    p->ntt_digits = make_ntt_plan(c, dl);
    p->modup_all = make_baseconv_batch(c, p->modup, p->digit_src_off, std::vector<int>(beta, 0));
    std::vector<int> psrc, qdst, qoff;
    for (int k = 0; k < K; ++k) psrc.push_back(c.nQ + k);
    for (int i = 0; i < L; ++i) { qdst.push_back(i); qoff.push_back(i); }
    p->moddown = make_baseconv_plan(c, psrc, qdst, qoff);
    for (int polys = 1; polys <= 2; ++polys) {
        std::vector<LimbDesc> ip, nc;
        for (int pp = 0; pp < polys; ++pp) {
            for (int k = 0; k < K; ++k)
                ip.push_back(LimbDesc{(uint32_t)(pp * W + L + k), (uint32_t)(pp * K + k), (uint32_t)(c.nQ + k), 0});
            for (int i = 0; i < L; ++i)
                nc.push_back(LimbDesc{(uint32_t)(pp * L + i), (uint32_t)(pp * L + i), (uint32_t)i, 0});
        }
        p->intt_p[polys] = make_ntt_plan(c, ip);
        {
            std::vector<BaseConvPlan> pl(polys, p->moddown);
            std::vector<int> io, ob;
            for (int pp = 0; pp < polys; ++pp) { io.push_back(pp * K); ob.push_back(pp * L); }
            p->moddown_b[polys] = make_baseconv_batch(c, pl, io, ob);
        }
        p->ntt_conv[polys] = make_ntt_plan(c, nc);
    }
    return *plans_.emplace(level, std::move(p)).first->second;
}

void PolyEvaluator::decompose(int level, const u64* c1_ntt, u64* digits) {
    const KsPlan& p = ks_plan(level);
    DevBuf coef((size_t)p.L * ctx.N * 8);
    ntt_inverse(ctx, c1_ntt, coef.u(), p.intt_c1);
    base_convert_batch(ctx, p.modup_all, coef.u(), digits, false);
    ntt_forward(ctx, digits, digits, p.ntt_digits);
}

void PolyEvaluator::moddown(int level, const u64* acc, int polys, const u64* addend, int addend_stride,
                            int addend_polys, u64* out, int out_stride) {
    OB_CHECK(polys == 1 || polys == 2, "moddown handles 1 or 2 polynomials");
    const KsPlan& p = ks_plan(level);
    const size_t N = ctx.N;
    DevBuf pc((size_t)polys * p.K * N * 8), conv((size_t)polys * p.L * N * 8);
    ntt_inverse(ctx, acc, pc.u(), p.intt_p[polys]);
    base_convert_batch(ctx, p.moddown_b[polys], pc.u(), conv.u(), true);
    ntt_forward(ctx, conv.u(), conv.u(), p.ntt_conv[polys]);
    moddown_finish(ctx, level, acc, p.W, conv.u(), addend, addend_stride, addend_polys, out, out_stride, polys);
}

void PolyEvaluator::keyswitch(int level, const u64* c1_ntt, const u64* evk, const u64* addend, int addend_stride,
                              int addend_polys, u64* out, int out_stride) {
    const KsPlan& p = ks_plan(level);
    DevBuf digits(digits_words(level) * 8), acc((size_t)2 * p.W * ctx.N * 8);
    decompose(level, c1_ntt, digits.u());
    ks_inner_product(ctx, level, c1_ntt, digits.u(), evk, acc.u(), nullptr, false);
    moddown(level, acc.u(), 2, addend, addend_stride, addend_polys, out, out_stride);
}

void PolyEvaluator::rescale(int level, const u64* in, int in_stride, u64* out, int out_stride, int polys) {
    OB_CHECK(level >= 1, "cannot rescale at level 0");
    OB_CHECK(polys >= 1 && polys <= 2, "rescale handles 1 or 2 polynomials");
    const size_t N = ctx.N;
    auto& ci = rescale_intt_[polys]; auto& cf = rescale_ntt_[polys];
    int key = level * 4096 + in_stride;
    if (!ci.count(key)) {
        std::vector<LimbDesc> a, b;
        for (int pp = 0; pp < polys; ++pp) {
            a.push_back(LimbDesc{(uint32_t)(pp * in_stride + level), (uint32_t)pp, (uint32_t)level, 0});
            for (int i = 0; i < level; ++i)
                b.push_back(LimbDesc{(uint32_t)(pp * level + i), (uint32_t)(pp * level + i), (uint32_t)i, 0});
        }
        ci[key] = make_ntt_plan(ctx, a); cf[key] = make_ntt_plan(ctx, b);
    }
    DevBuf last((size_t)polys * N * 8), tmp((size_t)polys * level * N * 8);
    ntt_inverse(ctx, in, last.u(), ci[key]);
    rescale_prep(ctx, level, last.u(), tmp.u(), polys);
    ntt_forward(ctx, tmp.u(), tmp.u(), cf[key]);
    rescale_finish(ctx, level, in, in_stride, tmp.u(), out, out_stride, polys);
}

}  // namespace ob
