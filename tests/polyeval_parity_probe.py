"""Probe (not a test): EvaluatePolynomial on the GPU against oracle/scheme.py poly_evaluate, word for word.
Run on a B200:  python tests/polyeval_parity_probe.py   -- prints, per polynomial, whether the output ciphertexts are equal.
Seen green on a B200 at the end of round 1 (all five polynomials equal); the same comparison is now
tests/test_gpu_scheme.py::test_polynomial_evaluation_words_equal_the_oracles.  Kept as a stand-alone probe."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.scheme import OScheme
from orion_b200.bindings import B200Library

LOGN, LOGQ, LOGP, LOGSCALE, H, SEED = 12, [50] + [40] * 7, [56, 56], 40, 64, 7
B = B200Library(); B.B200_SetSeed(SEED)
B.NewScheme(LOGN, LOGQ, LOGP, LOGSCALE, H, "standard", "", "none")
B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewPolynomialEvaluator()
O = OScheme(LOGN, LOGQ, LOGP, LOGSCALE, H, SEED)
O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
rng = np.random.default_rng(1)
x = rng.uniform(-1, 1, 1 << (LOGN - 1)).astype(np.float32)
ct = B.Encrypt(B.Encode(x.tolist(), 7, 1 << LOGSCALE))
oct_ = O.encrypt(O.encode(x.astype(np.float64), 7, 2.0 ** LOGSCALE))


def words(c):
    level, deg = B.GetCiphertextLevel(c), B.GetCiphertextDegree(c)
    out = np.empty((deg + 1, level + 1, 1 << LOGN), dtype=np.uint64)
    B.B200_ExportCiphertext(c, out)
    return out


print("input ciphertext equal:", np.array_equal(words(ct), oct_.data))
for deg, cheb in ((7, True), (15, True), (27, True), (31, True), (5, False)):
    c = (rng.uniform(-1, 1, deg + 1) / (1 if cheb else deg)).astype(np.float32)
    poly = B.GenerateChebyshev(c.tolist()) if cheb else B.GenerateMonomial(c.tolist())
    g = words(B.EvaluatePolynomial(ct, poly, 1 << LOGSCALE))
    o = O.poly_evaluate(oct_, c, cheb, 2.0 ** LOGSCALE).data
    print("degree %2d %s: output ciphertext equal: %s" % (deg, "chebyshev" if cheb else "monomial", g.shape == o.shape and np.array_equal(g, o)))
