"""Composite minimax approximation of sign(x) -- host-only replacement for GenerateMinimaxSignCoeffs
(orion/backend/lattigo/polyeval.go:91-167, which calls Lattigo's minimax.GenMinimaxCompositePolynomial).

Same arguments and output layout as the Go export: for degrees [d_0, ..., d_{k-1}] it returns the flattened Chebyshev
coefficients (basis on [-1, 1], d_i + 1 values each, ascending) of odd polynomials p_0, ..., p_{k-1} such that
p_{k-1}(...p_0(x)) ~ sign(x) for 2^-logalpha <= |x| <= 1; the last polynomial is then divided by 2 and shifted by 0.5
so that the composition outputs [0, 1] (polyeval.go:136-143).  Each stage is the minimax (Remez) odd polynomial
approximating 1 on the interval the previous stage maps [2^-logalpha, 1] into; intermediate stages are rescaled so
their range stays inside [-1, 1].  Target errors are ~2^-logerr >> 2^-52, so the exchange runs in float64
(`prec` is accepted for signature compatibility).
"""
import functools

import numpy as np
from numpy.polynomial import chebyshev as C


def _remez_odd_one(a, b, deg, iters=60):
    """Minimax odd polynomial of degree `deg` (Chebyshev basis) approximating 1 on [a, b], 0 < a < b <= 1."""
    ks = np.arange(1, deg + 1, 2)
    m = len(ks)
    t = np.cos(np.pi * (np.arange(m + 1) + 0.5) / (m + 1))[::-1]
    x = np.exp(np.log(a) + (np.log(b) - np.log(a)) * (t + 1) / 2) if a / b < 0.25 else a + (b - a) * (t + 1) / 2
    grid = np.unique(np.concatenate([np.exp(np.linspace(np.log(a), np.log(b), 8192)), np.linspace(a, b, 8192)]))

    def basis(pts):
        out = np.empty((len(pts), m))
        for j, k in enumerate(ks):
            c = np.zeros(k + 1); c[k] = 1.0
            out[:, j] = C.chebval(pts, c)
        return out

    G = basis(grid)
    coef, E = None, 0.0
    for _ in range(iters):
        A = np.hstack([basis(x), ((-1.0) ** np.arange(m + 1))[:, None]])
        sol = np.linalg.solve(A, np.ones(m + 1))
        coef, E = sol[:m], sol[m]
        err = G @ coef - 1.0
        # local extrema of the error (including the end points), then an alternating subset of m+1 of them
        idx = [0] + [i for i in range(1, len(grid) - 1) if (err[i] - err[i - 1]) * (err[i + 1] - err[i]) <= 0] + [len(grid) - 1]
        pts = []
        for i in idx:
            if pts and np.sign(err[i]) == np.sign(err[pts[-1]]):
                if abs(err[i]) > abs(err[pts[-1]]):
                    pts[-1] = i
            else:
                pts.append(i)
        while len(pts) > m + 1:                      # drop the weaker end until m+1 alternating points remain
            if abs(err[pts[0]]) < abs(err[pts[-1]]):
                pts.pop(0)
            else:
                pts.pop()
        if len(pts) < m + 1:
            break
        newx = grid[pts]
        emax = np.max(np.abs(err))
        if emax - abs(E) <= 1e-6 * abs(E) or np.allclose(newx, x, rtol=0, atol=1e-15):
            x = newx
            break
        x = newx
    full = np.zeros(deg + 1)
    full[ks] = coef
    return full, float(np.max(np.abs(G @ coef - 1.0)))


@functools.lru_cache(maxsize=None)
def _composite(degrees, logalpha):
    a, b = 2.0 ** (-logalpha), 1.0
    polys = []
    for i, d in enumerate(degrees):
        d = d if d % 2 == 1 else d - 1               # odd polynomials; an even degree keeps a zero leading coefficient
        c, err = _remez_odd_one(a, b, d)
        c = np.concatenate([c, np.zeros(degrees[i] - d)])
        if i < len(degrees) - 1:
            top = (1.0 + err) * (1.0 + 1e-6)         # the error is measured on a grid: a 1e-6 margin keeps the true
            c = c / top                              # range inside [-1, 1], the next stage's Chebyshev domain
            a, b = (1.0 - err) / top, 1.0
        polys.append(c)
    return polys


def generate_minimax_sign_coeffs(degrees, prec=128, logalpha=12, logerr=12, debug=False):
    polys = [p.copy() for p in _composite(tuple(int(d) for d in degrees), int(logalpha))]
    polys[-1] = polys[-1] / 2.0
    polys[-1][0] += 0.5
    return [float(v) for p in polys for v in p]


def evaluate_composite(coeff_flat, degrees, x):
    """Reference evaluation of the composition (used by the tests)."""
    out, pos = np.asarray(x, dtype=np.float64), 0
    for d in degrees:
        out = C.chebval(out, np.asarray(coeff_flat[pos:pos + d + 1]))
        pos += d + 1
    return out
