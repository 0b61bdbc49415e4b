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
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewPolynomialEvaluator()
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
def probe(c, tag):
    c = np.asarray(c, dtype=np.float32)
    deg = len(c) - 1
    poly = B.GenerateChebyshev(c.tolist())
    osc = O.Q[level - int(deg).bit_length()]
    got = ex(B.EvaluatePolynomial(ct, poly, osc))
    want = O.poly_evaluate(oct_, c, True, LD(osc)).data
    bad = np.argwhere(got != want)
    from numpy.polynomial import chebyshev as Cc
    truth = Cc.chebval(x.astype(np.float64), c.astype(np.float64))
    g_ct = B.EvaluatePolynomial(ct, poly, osc)
    re, im = np.empty(len(x)), np.empty(len(x)); B.B200_DecodeF64(B.Decrypt(g_ct), re, im)
    o_val = np.real(O.decode(O.decrypt(O.poly_evaluate(oct_, c, True, LD(osc)))))
    print('   max err GPU %.3e  oracle %.3e' % (np.max(np.abs(re - truth)), np.max(np.abs(o_val - truth))))
    print('%-28s equal %s' % (tag, len(bad) == 0), '' if len(bad) == 0 else 'differing words %d, limbs %s polys %s' % (len(bad), sorted(set(bad[:, 1])), sorted(set(bad[:, 0]))))
for k in range(2, 8):
    c = [0.0] * 8
    c[k] = 0.5; c[7] = c[7] or 0.25
    probe(c, 'T%d + T7' % k)
probe([0, 0, 0, 0.5, 0.25], 'T3 + T4')
probe([0.1, 0, 0, 0.5, 0.25], 'c0 + T3 + T4')
probe([0, 0.3, 0, 0, 0.25], 'T1 + T4')
probe([0, 0, 0.3, 0, 0.25], 'T2 + T4')
probe([0, 0, 0, 0, 0, 0.25], 'T5')
probe([0, 0, 0, 0, 0, 0, 0.25], 'T6')
B.DeleteScheme()
