// polyeval.h -- polynomial objects and depth-optimal evaluation (polyeval.cu); LG/polyeval.go:33-84.
#pragma once
#include "scheme.h"

namespace ob {

struct Polynomial {
    bool chebyshev = false;            // basis on [-1, 1]
    std::vector<double> coeffs;        // ascending: c_0 first
    int degree() const { int d = (int)coeffs.size() - 1; while (d > 0 && coeffs[d] == 0.0) --d; return d; }
};

int poly_push(std::unique_ptr<Polynomial> p);
Polynomial& poly_get(int id);
void poly_reset();
int poly_depth(const Polynomial& p);    // ceil(log2(degree + 1)), what orion/nn/activation.py:22-23 budgets
std::unique_ptr<Ciphertext> poly_evaluate(Scheme& s, const Ciphertext& in, const Polynomial& poly, scale_t out_scale);

}  // namespace ob
