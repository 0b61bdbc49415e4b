"""Composite minimax approximation of sign(x) -- host-only replacement for GenerateMinimaxSignCoeffs
(orion/backend/lattigo/polyeval.go:91-167, which calls Lattigo's minimax.GenMinimaxCompositePolynomial).

Same arguments and output layout as the Go export: for degrees [d_0, ..., d_{k-1}] it returns the flattened Chebyshev
coefficients (basis on [-1, 1], d_i + 1 values each, ascending) of odd polynomials p_0, ..., p_{k-1} such that
p_{k-1}(...p_0(x)) ~ sign(x) for 2^-logalpha <= |x| <= 1; the last polynomial is then divided by 2 and shifted by 0.5
so that the composition outputs [0, 1] (polyeval.go:136-143).  Each stage is the minimax (Remez) odd polynomial
approximating 1 on the interval the previous stage maps [2^-logalpha, 1] into, widened by 2^-logerr on both sides (the
room the caller leaves for CKKS noise between stages); intermediate stages are rescaled so their range stays inside
[-1, 1], the next stage's Chebyshev domain.

Multi-precision is not optional here (the Go export takes `prec` for the same reason): the later stages approximate 1 on
a narrow interval next to 1 -- [0.95, 1] for the degree-27 stage of the default [15, 15, 27] -- where the minimax error is
~1e-20 and the exchange system is singular to float64.  A float64 exchange stops at SOME polynomial that meets the error
bound, but not the minimax one, and such a polynomial can oscillate between -2 and 4 on [0, 0.95] where the true minimax
polynomial rises monotonically from 0 to 1: inputs that fall into the gap (-alpha, alpha) -- a few per cent of the
activations of a ReLU layer -- then come out with magnitudes > 1 and their noise amplified a thousandfold, which a
bootstrapped ResNet-20 does not survive.  The exchange therefore runs in mpmath at max(prec, 128) + 128 bits.
"""
import functools

import mpmath as mp
import numpy as np
from numpy.polynomial import chebyshev as C


def _clenshaw_odd(coef, ks_max, x):
    """sum_j coef[j] * T_{2j+1}(x) by the three-term recurrence (mpf arithmetic)."""
    t0, t1 = mp.mpf(1), x
    acc = coef[0] * t1
    two_x = 2 * x
    j = 1
    for k in range(2, ks_max + 1):
        t0, t1 = t1, two_x * t1 - t0
        if k & 1:
            acc += coef[j] * t1
            j += 1
    return acc


def _remez_odd_one(a, b, deg, prec_bits, iters=80):
    """Minimax odd polynomial of degree `deg` (Chebyshev basis on [-1, 1]) approximating 1 on [a, b], 0 < a < b <= 1.
    Returns (coefficients as float64 array of length deg + 1, max error as float)."""
    with mp.workprec(prec_bits):
        a, b = mp.mpf(a), mp.mpf(b)
        ks = list(range(1, deg + 1, 2))
        m = len(ks)
        # initial reference: Chebyshev points of [a, b] (log-spaced when the interval spans orders of magnitude)
        ts = [mp.cos(mp.pi * (mp.mpf(i) + mp.mpf(1) / 2) / (m + 1)) for i in range(m + 1)][::-1]
        if a / b < mp.mpf(1) / 4:
            la, lb = mp.log(a), mp.log(b)
            xs = [mp.exp(la + (lb - la) * (t + 1) / 2) for t in ts]
        else:
            xs = [a + (b - a) * (t + 1) / 2 for t in ts]

        def basis_row(x):
            row, t0, t1 = [], mp.mpf(1), x
            two_x = 2 * x
            row.append(t1)
            for k in range(2, deg + 1):
                t0, t1 = t1, two_x * t1 - t0
                if k & 1:
                    row.append(t1)
            return row

        # scan grid: dense in both linear and logarithmic spacing (the error equioscillates on a log scale near a << b)
        G = 48 * (m + 1)
        grid = sorted(set([a + (b - a) * mp.mpf(i) / G for i in range(G + 1)] +
                          ([mp.exp(mp.log(a) + (mp.log(b) - mp.log(a)) * mp.mpf(i) / G) for i in range(G + 1)] if a / b < mp.mpf(1) / 4 else [])))
        coef, E = None, mp.mpf(0)
        for _ in range(iters):
            A = mp.matrix(m + 1, m + 1)
            rhs = mp.matrix(m + 1, 1)
            for i, x in enumerate(xs):
                row = basis_row(x)
                for j in range(m):
                    A[i, j] = row[j]
                A[i, m] = mp.mpf(-1) ** i
                rhs[i] = 1
            sol = mp.lu_solve(A, rhs)
            coef, E = [sol[j] for j in range(m)], sol[m]
            err = lambda x: _clenshaw_odd(coef, deg, x) - 1
            ev = [err(x) for x in grid]
            # local extrema of the error on the grid (end points included), refined by a golden-section search
            cand = [0] + [i for i in range(1, len(grid) - 1) if (ev[i] - ev[i - 1]) * (ev[i + 1] - ev[i]) <= 0] + [len(grid) - 1]
            pts = []
            for i in cand:
                x, e = grid[i], ev[i]
                if 0 < i < len(grid) - 1:
                    lo, hi = grid[i - 1], grid[i + 1]
                    sgn = 1 if e > 0 else -1
                    gr = (mp.sqrt(5) - 1) / 2
                    c, d = hi - gr * (hi - lo), lo + gr * (hi - lo)
                    fc, fd = sgn * err(c), sgn * err(d)
                    for _g in range(40):
                        if fc > fd:
                            hi, d, fd = d, c, fc
                            c = hi - gr * (hi - lo); fc = sgn * err(c)
                        else:
                            lo, c, fc = c, d, fd
                            d = lo + gr * (hi - lo); fd = sgn * err(d)
                    x = (lo + hi) / 2
                    e = err(x)
                if pts and (e > 0) == (pts[-1][1] > 0):
                    if abs(e) > abs(pts[-1][1]):
                        pts[-1] = (x, e)
                else:
                    pts.append((x, e))
            while len(pts) > m + 1:                      # drop the weaker end until m + 1 alternating points remain
                if abs(pts[0][1]) < abs(pts[-1][1]):
                    pts.pop(0)
                else:
                    pts.pop()
            if len(pts) < m + 1:
                raise RuntimeError("minimax exchange lost alternation points (raise prec): degree %d on [%s, %s]" %
                                   (deg, mp.nstr(a, 8), mp.nstr(b, 8)))
            emax = max(abs(e) for _, e in pts)
            xs = [x for x, _ in pts]
            if emax - abs(E) <= abs(E) * mp.mpf(10) ** -6:
                break
        else:
            raise RuntimeError("minimax exchange did not converge: degree %d on [%s, %s]" % (deg, mp.nstr(a, 8), mp.nstr(b, 8)))
        full = np.zeros(deg + 1)
        for j, k in enumerate(ks):
            full[k] = float(coef[j])
        return full, coef, emax


@functools.lru_cache(maxsize=None)
def _composite(degrees, prec, logalpha, logerr):
    prec_bits = max(int(prec), 128) + 128
    with mp.workprec(prec_bits):
        a, b = mp.mpf(2) ** (-logalpha), mp.mpf(1)
        e = mp.mpf(2) ** (-logerr)
        polys = []
        for i, d in enumerate(degrees):
            dd = d if d % 2 == 1 else d - 1              # odd polynomials; an even degree keeps a zero leading coefficient
            _, coef, err = _remez_odd_one(a, b, dd, prec_bits)
            if i < len(degrees) - 1:
                top = 1 + err                            # the stage maps [a, b] into [1 - err, 1 + err]: normalise to [., 1]
                coef = [c / top for c in coef]
                # next stage's interval, with 2^-logerr of room below for the noise the evaluation adds in between
                a, b = (1 - err) / top - e, mp.mpf(1)
                if a <= 0:
                    raise ValueError("degrees %r cannot separate +-2^-%d" % (list(degrees), logalpha))
            c = np.zeros(d + 1)
            for j, k in enumerate(range(1, dd + 1, 2)):
                c[k] = float(coef[j])
            if i < len(degrees) - 1:
                # float64 rounding of the coefficients must not push the range above 1
                x = np.cos(np.linspace(0, np.pi / 2, 20001))
                peak = float(np.max(np.abs(C.chebval(x, c))))
                if peak > 1.0:
                    c = c / (peak * (1 + 1e-15))
            polys.append(c)
    return polys


def generate_minimax_sign_coeffs(degrees, prec=128, logalpha=12, logerr=12, debug=False):
    polys = [p.copy() for p in _composite(tuple(int(d) for d in degrees), int(prec), int(logalpha), int(logerr))]
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
