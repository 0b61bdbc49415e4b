// lintrans.cu -- hoisted BSGS diagonal matrix-vector product on the device.
//
//   out = sum_j rot_j( sum_i diag'_{j+i} (.) rot_i(ct) ),   diag'_{j+i} = diag_{j+i} rotated right by j
//
// Data flow (everything stays in the extended basis QP until the final ModDown, as Lattigo's
// MultiplyByDiagMatrixBSGS does; reference call site LG/lineartransform.go:96-113):
//   1. decompose c1 once (hoisting)                                       [beta][W][N]
//   2. per baby step i != 0: inner product with evk_i, + P*c0, automorphism  -> rot_i  [2][W][N]  (lazy, not / P)
//      rot_0 = (P*c0, P*c1) with zero P-limbs
//   3. per giant step j: 128-bit multiply-accumulate of the Montgomery-form diagonals against rot_i,
//      one REDC per output; j != 0: ModDown the c1 half, key switch it with evk_j (lazy), add the c0 half,
//      automorphism, accumulate
//   4. ModDown the accumulator: level and scale bookkeeping only, no rescale (PY/lt_evaluator.py:192 rescales).
// The reference GPU path instead does one full rotation + plaintext encode per diagonal (HG/bindings.py:1745-1826).
#include "lintrans.h"
#include <algorithm>
#include <cmath>

namespace ob {

namespace {
constexpr int TPB = 256;

// acc[z][c][t][x] = REDC( sum_k diag[pairs_z[k].x][td][x] * rot[pairs_z[k].y][c][t][x] ),  td = diag limb of t;
// blockIdx.z = giant step z (all giant steps of a transform in one launch)
constexpr int LTM_MAX = 64;
struct LtMacMulti { const int2* pairs[LTM_MAX]; int np[LTM_MAX]; };
// Two adjacent coefficients per thread (128-bit streams).  Limbs below 2^32 (the 30-bit scale primes of the shipped
// ResNet-20 / LoLA chains: 33 of 35 limbs): every product fits a word, 2^(64 - 2 bits) of them are summed with one IMAD.WIDE
// each before the partial sum is folded into the 128-bit accumulator; larger limbs keep the 128-bit multiply-accumulate
// with its conditional high-word reduction.  Same words either way (one REDC of the exact sum).
__global__ void __launch_bounds__(TPB)
lt_mac_kernel(const u64* __restrict__ diags, const u64* __restrict__ rots, const LtMacMulti mm, u64* __restrict__ acc_all,
              int N, int L, int W, int L_lt, int W_lt, int nQ, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = (blockIdx.x * TPB + threadIdx.x) * 2, z = blockIdx.z;
    const int ci = t < L ? t : nQ + (t - L);
    const int td = t < L ? t : L_lt + (t - L);
    const ModConst m = mc[ci];
    const int2* __restrict__ pairs = mm.pairs[z];
    const int npairs = mm.np[z];
    u64* acc = acc_all + (size_t)z * 2 * W * N;
    u64 h0a = 0, l0a = 0, h1a = 0, l1a = 0, h0b = 0, l0b = 0, h1b = 0, l1b = 0;
    if ((m.q >> 32) == 0) {
        const int fold = 1 << min(30, 64 - 2 * m.bits);
        u64 s0a = 0, s1a = 0, s0b = 0, s1b = 0;
        int cnt = 0;
        for (int k = 0; k < npairs; ++k) {
            const int2 pr = __ldg(&pairs[k]);
            const ulonglong2 d = *(const ulonglong2*)(diags + ((size_t)pr.x * W_lt + td) * N + x);
            const u64* r = rots + (size_t)pr.y * 2 * W * N;
            const ulonglong2 r0 = *(const ulonglong2*)(r + (size_t)t * N + x), r1 = *(const ulonglong2*)(r + ((size_t)W + t) * N + x);
            s0a += (u64)(uint32_t)d.x * (uint32_t)r0.x; s0b += (u64)(uint32_t)d.y * (uint32_t)r0.y;
            s1a += (u64)(uint32_t)d.x * (uint32_t)r1.x; s1b += (u64)(uint32_t)d.y * (uint32_t)r1.y;
            if (++cnt == fold) {
                l0a += s0a; h0a += (l0a < s0a); l0b += s0b; h0b += (l0b < s0b);
                l1a += s1a; h1a += (l1a < s1a); l1b += s1b; h1b += (l1b < s1b);
                s0a = s1a = s0b = s1b = 0; cnt = 0;
            }
        }
        l0a += s0a; h0a += (l0a < s0a); l0b += s0b; h0b += (l0b < s0b);
        l1a += s1a; h1a += (l1a < s1a); l1b += s1b; h1b += (l1b < s1b);
    } else {
        for (int k = 0; k < npairs; ++k) {
            const int2 pr = __ldg(&pairs[k]);
            const ulonglong2 d = *(const ulonglong2*)(diags + ((size_t)pr.x * W_lt + td) * N + x);
            const u64* r = rots + (size_t)pr.y * 2 * W * N;
            const ulonglong2 r0 = *(const ulonglong2*)(r + (size_t)t * N + x), r1 = *(const ulonglong2*)(r + ((size_t)W + t) * N + x);
            mac128(h0a, l0a, d.x, r0.x, m.q); mac128(h0b, l0b, d.y, r0.y, m.q);
            mac128(h1a, l1a, d.x, r1.x, m.q); mac128(h1b, l1b, d.y, r1.y, m.q);
        }
    }
    *(ulonglong2*)(acc + (size_t)t * N + x) = make_ulonglong2(mont_redc(h0a, l0a, m.q, m.qinv_neg), mont_redc(h0b, l0b, m.q, m.qinv_neg));
    *(ulonglong2*)(acc + ((size_t)W + t) * N + x) = make_ulonglong2(mont_redc(h1a, l1a, m.q, m.qinv_neg), mont_redc(h1b, l1b, m.q, m.qinv_neg));
}

// out[p][t][x] = t < L ? (P mod q_t) * c[p][t][x] : 0        (lift of a Q-basis ciphertext to "P * ct" in QP)
__global__ void __launch_bounds__(TPB)
lift_times_p_kernel(const u64* __restrict__ c, int c_stride, u64* __restrict__ out, int N, int L, int W,
                    const ulonglong2* __restrict__ pmodq, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, p = blockIdx.z, x = blockIdx.x * TPB + threadIdx.x;
    u64 v = 0;
    if (t < L) { const ulonglong2 w = pmodq[t]; v = shoup_mul(c[((size_t)p * c_stride + t) * N + x], w.x, w.y, mc[t].q); }
    out[((size_t)p * W + t) * N + x] = v;
}
// acc[t][x] += (P mod q_t) * c[t][x]   for t < L
__global__ void __launch_bounds__(TPB)
add_times_p_kernel(u64* __restrict__ acc, const u64* __restrict__ c, int N,
                   const ulonglong2* __restrict__ pmodq, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = blockIdx.x * TPB + threadIdx.x;
    const u64 q = mc[t].q; const ulonglong2 w = pmodq[t];
    const size_t o = (size_t)t * N + x;
    acc[o] = add_mod(acc[o], shoup_mul(c[o], w.x, w.y, q), q);
}

HandleTable<LinearTransform>& g_lts = *new HandleTable<LinearTransform>();   // heap-allocated and never destroyed: no static-destruction-order hazards (and no CUDA calls) at process exit
}  // namespace

int lt_push(std::unique_ptr<LinearTransform> lt) { return g_lts.add(std::move(lt)); }
LinearTransform& lt_get(int id) { return g_lts.get(id, "linear transform"); }
void lt_delete(int id) { g_lts.del(id); }
void lt_reset() { if (!g_lts.live.empty()) cudaStreamSynchronize(stream()); g_lts.reset(); }

void bsgs_index(const std::vector<int>& diags, int slots, int N1, std::map<int, std::vector<int>>& index,
                std::vector<int>& giants, std::vector<int>& babies) {
    index.clear();
    std::set<int> g, b;
    for (int rot : diags) {
        rot &= (slots - 1);
        const int j = ((rot / N1) * N1) & (slots - 1), i = rot & (N1 - 1);
        index[j].push_back(i);
        g.insert(j); b.insert(i);
    }
    for (auto& kv : index) std::sort(kv.second.begin(), kv.second.end());
    giants.assign(g.begin(), g.end()); babies.assign(b.begin(), b.end());
}

int find_best_bsgs_ratio(const std::vector<int>& diags, int slots, int log_max_ratio) {
    const double max_ratio = (double)(1 << log_max_ratio);
    for (int N1 = 1; N1 < slots; N1 <<= 1) {
        std::map<int, std::vector<int>> idx; std::vector<int> g, b;
        bsgs_index(diags, slots, N1, idx, g, b);
        const double r = (double)((int)g.size() - 1) / (double)((int)b.size() - 1);   // IEEE: x/0 = inf, 0/0 = nan
        if (r == max_ratio) return N1;
        if (r < max_ratio) return N1 / 2 > 0 ? N1 / 2 : 1;
    }
    return 1;
}

std::unique_ptr<LinearTransform> lt_build(Scheme& s, const std::map<int, std::vector<double2>>& diagonals, int level,
                                          scale_t scale, int log_slots, int log_bsgs_ratio, bool encode) {
    capture_abort();                                          // encoded diagonals are persistent state
    OB_CHECK(level >= 0 && level <= s.max_level, "linear transform level out of range");
    OB_CHECK(s.nP > 0, "linear transforms need special primes (LogP)");
    auto lt = std::make_unique<LinearTransform>();
    const int slots = 1 << log_slots, N = s.N;
    lt->level = level; lt->log_slots = log_slots; lt->scale = scale; lt->W = level + 1 + s.nP;
    for (auto& kv : diagonals) {
        OB_CHECK(!encode || (int)kv.second.size() == slots, "diagonal length must equal the slot count");
        lt->diag_idx.push_back(kv.first & (slots - 1));
    }
    std::sort(lt->diag_idx.begin(), lt->diag_idx.end());
    lt->diag_idx.erase(std::unique(lt->diag_idx.begin(), lt->diag_idx.end()), lt->diag_idx.end());
    OB_CHECK(lt->diag_idx.size() == diagonals.size(), "duplicate diagonal indices");
    lt->N1 = log_bsgs_ratio < 0 ? slots : find_best_bsgs_ratio(lt->diag_idx, slots, log_bsgs_ratio);
    bsgs_index(lt->diag_idx, slots, lt->N1, lt->index, lt->giants, lt->babies);
    const int nd = (int)lt->diag_idx.size();
    for (int k = 0; k < nd; ++k) lt->slot_of[lt->diag_idx[k]] = k;
    std::map<int, int> baby_slot;
    for (size_t k = 0; k < lt->babies.size(); ++k) baby_slot[lt->babies[k]] = (int)k;

    // encode every diagonal, pre-rotated right by its giant step, at `scale` into QP (Montgomery form)
    if (encode) lt->diags.alloc((size_t)nd * lt->W * N * 8);
    const int chunk = std::max(1, std::min(nd, (int)((256u << 20) / ((size_t)lt->W * N * 8))));
    std::vector<double2> host((size_t)chunk * slots);
    DevBuf dv((size_t)chunk * slots * sizeof(double2));
    std::map<int, const std::vector<double2>*> by_idx;
    for (auto& kv : diagonals) by_idx[kv.first & (slots - 1)] = &kv.second;
    for (int k0 = 0; encode && k0 < nd; k0 += chunk) {
        const int nb = std::min(chunk, nd - k0);
        for (int k = 0; k < nb; ++k) {
            const int d = lt->diag_idx[k0 + k], j = ((d / lt->N1) * lt->N1) & (slots - 1);
            const std::vector<double2>& v = *by_idx[d];
            double2* dst = host.data() + (size_t)k * slots;
            for (int x = 0; x < slots; ++x) dst[x] = v[(x - j + slots) & (slots - 1)];
        }
        OB_CUDA(cudaMemcpyAsync(dv.p, host.data(), (size_t)nb * slots * sizeof(double2), cudaMemcpyHostToDevice, stream()));
        OB_CUDA(cudaStreamSynchronize(stream()));
        s.encode_device_batch(dv.as<double2>(), log_slots, nb, scale, level, true, lt->diags.u() + (size_t)k0 * lt->W * N, true);
    }
    for (auto& kv : lt->index) {
        std::vector<int2> pr;
        for (int i : kv.second) pr.push_back(make_int2(lt->slot_of[(kv.first + i) & (slots - 1)], baby_slot[i]));
        DevBuf& b = lt->pairs[kv.first];
        b.alloc(pr.size() * sizeof(int2));
        OB_CUDA(cudaMemcpyAsync(b.p, pr.data(), pr.size() * sizeof(int2), cudaMemcpyHostToDevice, stream()));
        OB_CUDA(cudaStreamSynchronize(stream()));
    }
    return lt;
}

std::vector<u64> lt_galois_elements(Scheme& s, const LinearTransform& lt) {
    std::set<u64> g;
    for (int j : lt.giants) if (j) g.insert(s.ctx->galois_element(j));
    for (int i : lt.babies) if (i) g.insert(s.ctx->galois_element(i));
    return std::vector<u64>(g.begin(), g.end());
}

std::unique_ptr<Ciphertext> lt_apply(Scheme& s, const LinearTransform& lt, const Ciphertext& in) {
    OB_HP("lt_apply (total)");
    OpPin pin;                                                  // the Galois keys collected below stay resident until the launches are queued
    OB_CHECK(lt.has_data(), "linear transform has no encoded diagonals (removed?)");
    OB_CHECK(in.degree == 1, "EvaluateLinearTransform expects a degree-1 ciphertext");
    Context& c = *s.ctx;
    const int level = std::min(in.level, lt.level), L = level + 1, K = c.nP, W = L + K, N = c.N;
    const int in_stride = in.level + 1;
    const size_t PW = (size_t)2 * W * N;                       // one QP ciphertext
    const int nb = (int)lt.babies.size();
    const ulonglong2* pmodq = c.d_pmodq.as<ulonglong2>();

    // compact c0, c1 at the working level
    DevBuf ct((size_t)2 * L * N * 8);
    ew_binary(c, EW_COPY, in.data.u(), in_stride, nullptr, 0, 1, ct.u(), L, 2, L, L);
    const u64* c0 = ct.u(); const u64* c1 = ct.u() + (size_t)L * N;

    // ---- baby steps: one decomposition (hoisting), then ALL rotations in one launch: inner product with evk_i,
    //      + P*c0, written through the automorphism's scatter table ----
    const size_t dig_words = s.pev->digits_words(level);
    DevBuf digits(dig_words * 8), rots((size_t)nb * PW * 8);
    int k_first = 0;
    if (nb > 0 && lt.babies[0] == 0) {
        OB_LAUNCH(lift_times_p_kernel, dim3(N / TPB, W, 2), TPB, 0, stream(), ct.u(), L, rots.u(), N, L, W, pmodq, c.dmc());
        k_first = 1;
    }
    if (nb > k_first) {
        s.pev->decompose(level, c1, digits.u());
        for (int k0 = k_first; k0 < nb; k0 += IPM_MAX) {
            const int n = std::min(IPM_MAX, nb - k0);
            IpMulti m{};
            for (int k = 0; k < n; ++k) {
                const u64 g = c.galois_element(lt.babies[k0 + k]);
                m.evk[k] = s.galois_key(g).data.u(); m.scatter[k] = c.galois_scatter(g);
            }
            ks_inner_product_multi(c, level, n, c1, 0, digits.u(), 0, m, rots.u() + (size_t)k0 * PW, PW, 1, c0, 0);
        }
    }

    // ---- giant steps, batched: all inner sums in one launch; then for the rotated ones ModDown(c1 half) -> decompose ->
    //      inner product with evk_j (+ c0 half, through the automorphism) as one launch sequence over all of them ----
    std::vector<int> gj; std::vector<const DevBuf*> gp; std::vector<int> gn;
    for (auto& kv : lt.index) { gj.push_back(kv.first); gp.push_back(&lt.pairs.at(kv.first)); gn.push_back((int)kv.second.size()); }
    const int ng = (int)gj.size();
    OB_CHECK(ng >= 1, "linear transform without diagonals");
    const int z_first = gj[0] == 0 ? 1 : 0;                    // giant step 0 needs no rotation (std::map: it comes first)
    // chunk so that the per-chunk digit buffers stay below ~6 GB
    const int chunk = std::max(1, (int)std::min<size_t>(std::min(IPM_MAX, LTM_MAX), ((size_t)6 << 30) / (dig_words * 8 + 3 * PW * 8)));
    DevBuf out(PW * 8);
    bool have_out = false;
    for (int zc = 0; zc < ng; zc += chunk) {
        const int n = std::min(chunk, ng - zc);
        DevBuf accs((size_t)n * PW * 8);
        LtMacMulti mm{};
        for (int z = 0; z < n; ++z) { mm.pairs[z] = gp[zc + z]->as<int2>(); mm.np[z] = gn[zc + z]; }
        OB_LAUNCH(lt_mac_kernel, dim3(N / (2 * TPB), W, n), TPB, 0, stream(), lt.diags.u(), rots.u(), mm, accs.u(), N, L, W,
                  lt.level + 1, lt.W, c.nQ, c.dmc());
        const int zr = (zc == 0) ? z_first : 0;                 // first rotated step of this chunk
        const int nr = n - zr;
        SumList lst{}; int ns = 0;
        if (have_out) lst.p[ns++] = out.u();
        if (zr == 1) lst.p[ns++] = accs.u();
        DevBuf rot_out((size_t)std::max(nr, 1) * PW * 8);
        if (nr > 0) {
            DevBuf u1((size_t)nr * L * N * 8), dig2((size_t)nr * dig_words * 8);
            const u64* acc_r = accs.u() + (size_t)zr * PW;
            s.pev->moddown_n(level, acc_r + (size_t)W * N, 2 * W, nr, nullptr, 0, 0, u1.u(), L);
            s.pev->decompose_n(level, u1.u(), L, nr, dig2.u());
            IpMulti m{};
            for (int z = 0; z < nr; ++z) {
                const u64 g = c.galois_element(gj[zc + zr + z]);
                m.evk[z] = s.galois_key(g).data.u(); m.scatter[z] = c.galois_scatter(g);
                lst.p[ns++] = rot_out.u() + (size_t)z * PW;
            }
            ks_inner_product_multi(c, level, nr, u1.u(), (size_t)L * N, dig2.u(), dig_words, m, rot_out.u(), PW, 2, acc_r, PW);
        }
        if (ns == 1 && lst.p[0] != out.u())
            OB_CUDA(cudaMemcpyAsync(out.p, lst.p[0], PW * 8, cudaMemcpyDeviceToDevice, stream()));
        else if (ns > 1)
            sum_qp(c, level, lst, ns, out.u());
        have_out = true;
    }
    auto res = s.new_ct(level, 1, in.scale * lt.scale, in.log_slots);
    s.pev->moddown(level, out.u(), 2, nullptr, 0, 0, res->data.u(), L);
    // ct in, every QP diagonal once, one evaluation key per rotated baby / giant step, ct out
    alg_bytes(8.0 * N * (4.0 * L + (double)lt.diag_idx.size() * W + (double)((nb - k_first) + (ng - z_first)) * 2.0 * c.num_digits(level) * W));
    return res;
}

}  // namespace ob
