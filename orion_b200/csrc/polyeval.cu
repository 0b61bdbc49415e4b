// polyeval.cu -- depth-optimal polynomial evaluation (monomial or Chebyshev basis on [-1,1]) with an exact target
// output scale.  Replaces LG/polyeval.go:33-84 (polynomial.Evaluator.Evaluate in Lattigo); SURVEY.md 8a row a13.
//
// Contract (as the reference's): input at level l, any scale; output at level l - ceil(log2(d+1)), scale == outScale.
// Algorithm (Paterson-Stockmeyer / baby-step giant-step with backward scale propagation):
//   power basis  P_n for n < B = 2^ceil(bitlen(d)/2) and n = 2^k:  P_{a+b} = P_a P_b  (Chebyshev: 2 P_a P_b - P_{a-b}),
//                a = 2^(ceil(log2 n)-1): depth ceil(log2 n); the subtraction happens BEFORE the rescale with P_{a-b}
//                multiplied by the integer round(scale(P_a P_b)/scale(P_{a-b})) so that scales match to 2^-30.
//   E(p, L, S)   a ciphertext at level L and scale exactly S:
//                  every needed power is in the basis at level >= L ->  sum_i round(c_i * S/s_i) * P_i  + round(c_0 * S)
//                  otherwise p = q * P_g + r (g = largest power of two <= deg p; Chebyshev division) and
//                  E = relin( rescale(E(q, L+1, (S/s_g) * q_{L+1})) * P_g ) + E(r, L, S)
//   result       rescale( E(p, l_out + 1, outScale * q_{l_out+1}) )
// All constants are integers times 1, so every scale is hit exactly; the only roundings are the rescales.
#include "polyeval.h"
#include <cmath>
#include <set>

namespace ob {

namespace {
HandleTable<Polynomial>& g_polys = *new HandleTable<Polynomial>();   // heap-allocated and never destroyed: no static-destruction-order hazards (and no CUDA calls) at process exit

int bitlen(int x) { int b = 0; while (x) { ++b; x >>= 1; } return b; }

struct PowerBasis {
    Scheme& s;
    bool cheb;
    std::map<int, std::unique_ptr<Ciphertext>> P;
    PowerBasis(Scheme& s_, bool c) : s(s_), cheb(c) {}

    const Ciphertext& get(int n) {
        auto it = P.find(n);
        if (it != P.end()) return *it->second;
        OB_CHECK(n >= 2, "power basis: bad index");
        const int a = 1 << (bitlen(n - 1) - 1), b = n - a;       // a = 2^(ceil(log2 n) - 1)
        const Ciphertext& A = get(a);
        const Ciphertext& Bc = get(b);
        auto prod = std::make_unique<Ciphertext>();
        s.mul_relin(A, Bc, *prod);
        if (cheb) {
            // T_{a+b} = 2 T_a T_b - T_{a-b}: the doubling and the subtraction in one fused launch.  T_{a-b} is multiplied by
            // the integer k = round(scale(T_a T_b) / scale(T_{a-b})) so the scales match up to 1/k; T_0 = 1 is the constant
            // round(scale).  Same residues as mul_int(2), mul_const(k) and a subtraction.
            const int c = a - b;
            if (c == 0) {
                s.lin_comb_int({prod.get()}, {2.0L}, -roundl(prod->scale), prod->level, prod->scale, *prod);
            } else {
                const Ciphertext& C = get(c);
                const scale_t k = roundl(prod->scale / C.scale);
                s.lin_comb_int({prod.get(), &C}, {2.0L, -k}, 0.0L, prod->level, prod->scale, *prod);
            }
        }
        s.rescale(*prod, *prod);
        return *P.emplace(n, std::move(prod)).first->second;
    }
    int level_of(int n, int in_level) const { return in_level - (n <= 1 ? 0 : bitlen(n - 1)); }

    // closure of the powers a set of requests will touch (same recursion as get())
    void require(int n, std::set<int>& need) const {
        if (n <= 1 || need.count(n)) return;
        need.insert(n);
        const int a = 1 << (bitlen(n - 1) - 1), b = n - a;
        require(a, need); require(b, need);
        if (cheb && a != b) require(a - b, need);
    }
    // compute every power of `need` depth by depth; the products of one depth only read shallower powers, so they are
    // issued on side streams and overlap on the GPU
    void precompute(const std::set<int>& need) {
        std::map<int, std::vector<int>> by_depth;
        for (int n : need) if (!P.count(n)) by_depth[bitlen(n - 1)].push_back(n);
        for (auto& kv : by_depth) {
            StreamFork f((int)std::min<size_t>(kv.second.size(), 8));
            int i = 0;
            for (int n : kv.second) { f.branch(i++); get(n); }
            f.join();
        }
    }
};

struct Evaluator {
    Scheme& s;
    PowerBasis& pb;
    int B, in_level;

    bool in_basis(int n) const { return n < B || (n & (n - 1)) == 0; }

    // dry run of E(): which powers will be read (same control flow, no ciphertexts)
    void collect(const std::vector<double>& p, int L, std::set<int>& need) const {
        int deg = (int)p.size() - 1;
        while (deg > 0 && p[deg] == 0.0) --deg;
        bool direct = true;
        for (int i = 1; i <= deg; ++i)
            if (p[i] != 0.0 && !(in_basis(i) && pb.level_of(i, in_level) >= L)) direct = false;
        if (direct) {
            for (int i = 1; i <= deg; ++i) if (p[i] != 0.0) pb.require(i, need);
            return;
        }
        if (deg < 2) return;                                       // E() reports the error
        const int g = 1 << (bitlen(deg) - 1);
        pb.require(g, need);
        std::vector<double> q(deg - g + 1, 0.0), r(p.begin(), p.begin() + g);
        if (pb.cheb) {
            q[0] = p[g];
            for (int k = 1; k <= deg - g; ++k) { q[k] = 2.0 * p[g + k]; r[g - k] -= p[g + k]; }
        } else {
            for (int k = 0; k <= deg - g; ++k) q[k] = p[g + k];
        }
        if (deg != g) collect(q, L + 1, need);
        bool r_zero = true;
        for (double v : r) if (v != 0.0) r_zero = false;
        if (!r_zero) collect(r, L, need);
    }

    // ct at level L, scale exactly S
    std::unique_ptr<Ciphertext> E(const std::vector<double>& p, int L, scale_t S) {
        int deg = (int)p.size() - 1;
        while (deg > 0 && p[deg] == 0.0) --deg;
        bool direct = true;
        for (int i = 1; i <= deg; ++i)
            if (p[i] != 0.0 && !(in_basis(i) && pb.level_of(i, in_level) >= L)) direct = false;
        if (direct) {
            // leaf: sum_i round(c_i * S/s_i) * P_i + round(c_0 * S) -- one fused launch (Scheme::lin_comb_scaled); powers that
            // sit above level L are read in place with their own limb stride instead of being copied down
            std::vector<const Ciphertext*> ops; std::vector<double> cs;
            for (int i = 1; i <= deg; ++i)
                if (p[i] != 0.0) { ops.push_back(&pb.get(i)); cs.push_back(p[i]); }
            auto acc = std::make_unique<Ciphertext>();
            if (!ops.empty()) {
                s.lin_comb_scaled(ops, cs, p[0], L, S, *acc);
                return acc;
            }
            // constant polynomial: 0 * P_1 + c_0
            auto c = s.clone(pb.get(1)); s.drop_level(*c, L);
            s.mul_int(*c, 0, *acc);
            acc->scale = S;
            if (p[0] != 0.0) s.add_const(*acc, p[0], *acc);
            return acc;
        }
        OB_CHECK(deg >= 2, "polynomial evaluation: level budget exhausted");
        const int g = 1 << (bitlen(deg) - 1);
        OB_CHECK(pb.level_of(g, in_level) >= L, "polynomial evaluation: not enough levels for this degree");
        std::vector<double> q(deg - g + 1, 0.0), r(p.begin(), p.begin() + g);
        if (pb.cheb) {
            q[0] = p[g];
            for (int k = 1; k <= deg - g; ++k) { q[k] = 2.0 * p[g + k]; r[g - k] -= p[g + k]; }
        } else {
            for (int k = 0; k <= deg - g; ++k) q[k] = p[g + k];
        }
        const Ciphertext& Pg = pb.get(g);
        std::unique_ptr<Ciphertext> pg;
        const Ciphertext* usePg = &Pg;
        if (Pg.level > L) { pg = s.clone(Pg); s.drop_level(*pg, L); usePg = pg.get(); }
        auto res = std::make_unique<Ciphertext>();
        bool r_zero = true;
        for (double v : r) if (v != 0.0) r_zero = false;
        std::unique_ptr<Ciphertext> R;
        {
            // quotient branch and remainder branch are independent: two side streams
            StreamFork f((deg != g && !r_zero) ? 2 : 1);
            f.branch(0);
            if (deg == g) {                                       // q is the constant c_g: no ciphertext product needed
                s.mul_const_scaled(*usePg, q[0], S / Pg.scale, *res);
            } else {
                OB_CHECK(L + 1 <= in_level, "polynomial evaluation: not enough levels for this degree");
                const scale_t qS = (S / Pg.scale) * s.q_at(L + 1);
                auto A = E(q, L + 1, qS);
                s.rescale(*A, *A);
                A->scale = S / Pg.scale;
                s.mul_relin(*A, *usePg, *res);
            }
            res->scale = S;
            if (!r_zero) { f.branch(1); R = E(r, L, S); }
            f.join();
        }
        if (!r_zero) s.add_ct(*res, *R, *res, false);
        return res;
    }
};
}  // namespace

int poly_push(std::unique_ptr<Polynomial> p) { return g_polys.add(std::move(p)); }
Polynomial& poly_get(int id) { return g_polys.get(id, "polynomial"); }
void poly_reset() { g_polys.reset(); }
int poly_depth(const Polynomial& p) { return bitlen(p.degree()); }

std::unique_ptr<Ciphertext> poly_evaluate(Scheme& s, const Ciphertext& in, const Polynomial& poly, scale_t out_scale) {
    OB_HP("poly_evaluate (total)");
    OB_CHECK(in.degree == 1, "EvaluatePolynomial expects a degree-1 ciphertext");
    const int d = poly.degree();
    OB_CHECK(d >= 1, "EvaluatePolynomial: degree must be at least 1");
    const int depth = bitlen(d), out_level = in.level - depth;
    OB_CHECK(out_level >= 0, "EvaluatePolynomial: not enough levels (need " + std::to_string(depth) + ")");
    PowerBasis pb(s, poly.chebyshev);
    pb.P[1] = s.clone(in);
    Evaluator ev{s, pb, 1 << ((depth + 1) / 2), in.level};
    std::vector<double> c(poly.coeffs.begin(), poly.coeffs.begin() + d + 1);
    {
        std::set<int> need;
        ev.collect(c, out_level + 1, need);
        pb.precompute(need);
    }
    auto res = ev.E(c, out_level + 1, out_scale * s.q_at(out_level + 1));
    s.rescale(*res, *res);
    res->scale = out_scale;
    return res;
}

}  // namespace ob
