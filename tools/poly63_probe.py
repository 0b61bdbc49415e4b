"""Debug helper (not a test): scheme-level parity probes on a chain of 60-bit primes (scales near 2^60)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.scheme import OScheme, LD
from orion_b200.bindings import B200Library
logn, h, seed = 12, 64, 7
logq, logp, logscale, level = [55] + [60] * 8, [61, 61, 61, 61], 40, 8
B = B200Library(); B.B200_SetSeed(seed)
B.NewScheme(logn, logq, logp, logscale, h, "standard", "", "none")
B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewPolynomialEvaluator(); B.NewEvaluator()
O = OScheme(logn, logq, logp, logscale, h, seed)
O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
rng = np.random.default_rng(1)
x = rng.uniform(-1, 1, 1 << (logn - 1)).astype(np.float32)
sc = O.Q[level]
ct = B.Encrypt(B.Encode(x.tolist(), level, sc))
oct_ = O.encrypt(O.encode(x.astype(np.float64), level, LD(sc)))
def ex(c):
    lv, dg = B.GetCiphertextLevel(c), B.GetCiphertextDegree(c)
    out = np.empty((dg + 1, lv + 1, O.N), dtype=np.uint64); B.B200_ExportCiphertext(c, out); return out
print('input equal', np.array_equal(ex(ct), oct_.data))
m = B.MulRelinCiphertextNew(ct, ct); om = O.mul_relin(oct_, oct_)
print('mul_relin equal', np.array_equal(ex(m), om.data), 'scale', B.GetCiphertextScale(m) == int(om.scale))
r = B.Rescale(m); orr = O.rescale(om)
print('rescale equal', np.array_equal(ex(r), orr.data))
a = B.AddScalarNew(ct, 0.37); oa = O.add_const(oct_, float(np.float32(0.37)))
print('add_const equal', np.array_equal(ex(a), oa.data))
k = B.MulScalarIntNew(ct, 3); ok = O.mul_int(oct_, 3)
print('mul_int equal', np.array_equal(ex(k), ok.data))
f = B.MulScalarFloatNew(ct, 0.37); of = O.mul_float(oct_, float(np.float32(0.37)))
print('mul_float equal', np.array_equal(ex(f), of.data))
for deg in (3, 7, 15):
    c = (rng.uniform(-1, 1, deg + 1)).astype(np.float32)
    poly = B.GenerateChebyshev(c.tolist())
    osc = O.Q[level - int(deg).bit_length()]
    got = ex(B.EvaluatePolynomial(ct, poly, osc))
    want = O.poly_evaluate(oct_, c, True, LD(osc)).data
    print('cheb deg', deg, 'equal', np.array_equal(got, want))
B.DeleteScheme()
