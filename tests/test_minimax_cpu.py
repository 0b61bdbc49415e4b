"""Host-side composite minimax sign generator (orion_b200/minimax.py) -- SURVEY.md rows a14 / 8(f)-4.

Replaces the Go export GenerateMinimaxSignCoeffs (orion/backend/lattigo/polyeval.go:91-167, Lattigo's
minimax.GenMinimaxCompositePolynomial).  Lattigo is not buildable here, so the acceptance test is the contract the
caller relies on (orion/nn/activation.py:201-299): for the defaults degrees=[15,15,27], logalpha=6, logerr=12 the
composition stays within 2^-logerr of (sign(x)+1)/2 on [-1,-2^-logalpha] u [2^-logalpha,1], and the output has the layout
the Go shim produces (polyeval.go:136-143: flattened Chebyshev coefficients, d_i+1 per stage, last stage halved and
shifted by 0.5)."""
import numpy as np
import pytest
from numpy.polynomial import chebyshev as C

from orion_b200.minimax import evaluate_composite, generate_minimax_sign_coeffs


def _domain(logalpha, n=200001):
    a = 2.0 ** (-logalpha)
    pos = np.unique(np.concatenate([np.linspace(a, 1.0, n), np.exp(np.linspace(np.log(a), 0.0, n))]))
    return pos


@pytest.mark.parametrize("degrees,logalpha,logerr", [([15, 15, 27], 6, 12),      # orion/nn/activation.py:264-269 defaults (ReLU)
                                                       ([15, 15, 27], 5, 12),
                                                       ([15, 15, 15, 27], 8, 12),
                                                       ([15, 27], 4, 10)])
def test_composite_sign_meets_the_error_bound(degrees, logalpha, logerr):
    flat = generate_minimax_sign_coeffs(degrees, 128, logalpha, logerr)
    # layout: sum(d_i + 1) float64 values, stage after stage, Chebyshev basis on [-1, 1], ascending order
    assert len(flat) == sum(d + 1 for d in degrees)
    assert all(isinstance(v, float) for v in flat)
    pos = _domain(logalpha)
    hi = evaluate_composite(flat, degrees, pos)
    lo = evaluate_composite(flat, degrees, -pos)
    bound = 2.0 ** (-logerr)
    assert np.max(np.abs(hi - 1.0)) <= bound, np.max(np.abs(hi - 1.0))
    assert np.max(np.abs(lo - 0.0)) <= bound, np.max(np.abs(lo))
    # inside the gap (-alpha, alpha) the TRUE minimax composition rises monotonically from 0 to 1.  A float64 exchange
    # converges to a different polynomial on the narrow last-stage interval (error bound met, but values between -2 and 4
    # in the gap -- see the module docstring): this assertion is what tells the two apart.
    mid = evaluate_composite(flat, degrees, np.linspace(-2.0 ** -logalpha, 2.0 ** -logalpha, 20001))
    assert np.min(mid) >= -bound and np.max(mid) <= 1.0 + bound, (np.min(mid), np.max(mid))
    assert np.all(np.diff(mid) >= -1e-9)
    assert abs(evaluate_composite(flat, degrees, np.array([0.0]))[0] - 0.5) < 1e-12          # odd composition: p(0) = 1/2


def test_stage_structure_matches_the_shims_post_processing():
    """polyeval.go:136-143: every stage but the last is an odd polynomial with range inside [-1, 1] (the next stage's Chebyshev
    domain); the last one is (odd polynomial)/2 + 0.5: its even coefficients vanish except c_0 = 0.5."""
    degrees = [15, 15, 27]
    flat = np.array(generate_minimax_sign_coeffs(degrees, 128, 6, 12))
    pos, stages = 0, []
    for d in degrees:
        stages.append(flat[pos:pos + d + 1]); pos += d + 1
    x = np.linspace(-1, 1, 100001)
    for c in stages[:-1]:
        assert np.all(c[0::2] == 0.0)                               # odd
        assert np.max(np.abs(C.chebval(x, c))) <= 1.0 + 1e-12       # stays inside the next stage's domain
    last = stages[-1]
    assert last[0] == 0.5 and np.all(last[2::2] == 0.0)
    y = C.chebval(x, last)
    assert np.allclose(y + y[::-1], 1.0, atol=1e-12)                # p(x) + p(-x) = 1


def test_generator_is_memoised_and_deterministic():
    """polyeval.go:116-125 memoises by (degrees, prec, logalpha, logerr); the callers get equal lists on every call."""
    a = generate_minimax_sign_coeffs([15, 15, 27], 128, 6, 12)
    b = generate_minimax_sign_coeffs([15, 15, 27], 128, 6, 12)
    assert a == b and a is not b


def test_binding_layer_returns_one_list_per_stage():
    """The Python-visible call (LG/bindings.py GenerateMinimaxSignCoeffs wrapper -> orion/backend/python/poly_evaluator.py)
    goes through B200Library without touching the GPU."""
    from orion_b200.bindings import B200Library
    B = B200Library()
    flat = B.GenerateMinimaxSignCoeffs([15, 15, 27], 128, 6, 12, 0)
    assert len(flat) == 16 + 16 + 28
    relu_in = np.array([-1.0, -0.5, -2.0 ** -6, 2.0 ** -6, 0.5, 1.0])
    s = evaluate_composite(flat, [15, 15, 27], relu_in)
    assert np.max(np.abs(relu_in * s - np.maximum(relu_in, 0.0))) <= 2.0 ** -12
