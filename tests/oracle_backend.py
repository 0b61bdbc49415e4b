"""The CPU oracle behind the reference's backend-call names -- TEST INFRASTRUCTURE ONLY (lives in tests/, imports oracle/).

`OracleBackend` exposes the subset of the Go shim's surface (orion/backend/lattigo/*.go) that the
recorded MLP / LoLA traces use, implemented with `oracle.scheme.OScheme`, so that `orion_b200.replay.Replayer` can drive
the ORACLE with the exact call sequence the reference's Python layer issues:
  * BASELINE.json configs[0]: "MLP (configs/mlp.yml) ... runs without a GPU" -- the whole encrypted inference on host cores,
    checked with the reference's own assertion (tests/models/test_mlp.py:45-48, MAE < 0.005 against the cleartext output);
  * the output CIPHERTEXT of the GPU replay must equal the oracle's, word for word.
Handles are smallest-free integers, in-place ops return the same id, `RescaleNew` rescales its input and returns a copy
(LG/evaluator.go:91-99) -- the conventions of the boundary, restated once more here so the replay is the same on both sides.
"""
import math

import numpy as np

from oracle.scheme import OScheme


class _Table:
    def __init__(self):
        self.d = {}

    def add(self, obj):
        i = 0
        while i in self.d:
            i += 1
        self.d[i] = obj
        return i


class OracleBackend:
    def __init__(self, seed):
        self.seed = seed
        self.O = None
        self.ct, self.pt, self.lt = _Table(), _Table(), _Table()

    # ---- scheme / keys (LG/scheme.go:35-86, LG/keygenerator.go:13-35) ----
    def NewScheme(self, logN, logQ, logP, logScale, h, ringType, keysPath, ioMode):
        # RingType ConjugateInvariant is served the way the product serves it (csrc/scheme.cu constructor): the standard
        # ring of degree 2N over the same prime chain (primes = 1 mod 4N), N real slots
        ci = ringType.lower() != "standard"
        self.O = OScheme(logN + (1 if ci else 0), list(logQ), list(logP), logScale, h, self.seed)

    def GenerateSecretKey(self): self.O.gen_secret_key()
    def GeneratePublicKey(self): self.O.gen_public_key()
    def GenerateRelinearizationKey(self): self.O.gen_relin_key()

    def _noop(self, *a): return None
    NewKeyGenerator = GenerateEvaluationKeys = NewEncoder = NewEncryptor = NewDecryptor = _noop
    NewPolynomialEvaluator = NewLinearTransformEvaluator = _noop

    def NewEvaluator(self):
        # LG/evaluator.go:25-31: the evaluator is created with the Galois keys of every power-of-two rotation
        i = 1
        while i < (1 << self.O.log_max_slots):
            self.O.galois_key(self.O.ctx.galois_element(i))
            i *= 2

    def GenerateConsolidatedRotationKeys(self, gal_els):
        for g in gal_els:
            self.O.galois_key(int(g))

    # ---- encoder / encryptor ----
    def Encode(self, values, level, scale):
        return self.pt.add(self.O.encode(np.asarray(values, dtype=np.float32).astype(np.float64), level, scale))

    def Decode(self, pt):
        return np.real(self.O.decode(self.pt.d[pt])).astype(np.float32).tolist()

    def Encrypt(self, pt): return self.ct.add(self.O.encrypt(self.pt.d[pt]))
    def Decrypt(self, ct): return self.pt.add(self.O.decrypt(self.ct.d[ct]))
    def DeletePlaintext(self, i): self.pt.d.pop(i)
    def DeleteCiphertext(self, i): self.ct.d.pop(i)
    def DeleteLinearTransform(self, i): self.lt.d.pop(i)
    def SetCiphertextScale(self, ct, scale): self.ct.d[ct].scale = np.longdouble(scale)
    def GetCiphertextLevel(self, ct): return self.ct.d[ct].level

    # ---- evaluator (LG/evaluator.go:48-317) ----
    def _inplace(self, ct, res):
        self.ct.d[ct] = res
        return ct

    def Rescale(self, ct): return self._inplace(ct, self.O.rescale(self.ct.d[ct]))

    def RescaleNew(self, ct):
        self.Rescale(ct)
        return self.ct.add(self.ct.d[ct].copy())

    def RotateNew(self, ct, rot): return self.ct.add(self.O.rotate(self.ct.d[ct], rot))
    def Rotate(self, ct, rot): return self._inplace(ct, self.O.rotate(self.ct.d[ct], rot))
    def AddCiphertext(self, a, b): return self._inplace(a, self.O.add(self.ct.d[a], self.ct.d[b]))
    def AddCiphertextNew(self, a, b): return self.ct.add(self.O.add(self.ct.d[a], self.ct.d[b]))
    def SubCiphertext(self, a, b): return self._inplace(a, self.O.add(self.ct.d[a], self.ct.d[b], sub=True))
    def AddPlaintext(self, a, p): return self._inplace(a, self.O.add(self.ct.d[a], self.pt.d[p]))
    def AddPlaintextNew(self, a, p): return self.ct.add(self.O.add(self.ct.d[a], self.pt.d[p]))
    def MulPlaintextNew(self, a, p): return self.ct.add(self.O.mul_pt(self.ct.d[a], self.pt.d[p]))
    def MulRelinCiphertextNew(self, a, b): return self.ct.add(self.O.mul_relin(self.ct.d[a], self.ct.d[b]))
    def MulRelinCiphertext(self, a, b): return self._inplace(a, self.O.mul_relin(self.ct.d[a], self.ct.d[b]))
    def MulScalarIntNew(self, a, k): return self.ct.add(self.O.mul_int(self.ct.d[a], k))
    def MulScalarFloatNew(self, a, c): return self.ct.add(self.O.mul_float(self.ct.d[a], float(np.float32(c))))
    def AddScalarNew(self, a, c): return self.ct.add(self.O.add_const(self.ct.d[a], float(np.float32(c))))

    # ---- linear transforms (LG/lineartransform.go:37-113) ----
    def GenerateLinearTransform(self, diag_idx, flat, level, bsgs_ratio, io_mode):
        O = self.O
        slots = 1 << O.log_max_slots
        flat = np.asarray(flat, dtype=np.float32).astype(np.float64)
        assert flat.size == len(diag_idx) * slots
        diags = {int(d): flat[i * slots:(i + 1) * slots] for i, d in enumerate(diag_idx)}
        log_ratio = int(math.log(bsgs_ratio))                       # natural log, as LG/lineartransform.go:67
        return self.lt.add(O.lt_build(diags, level, O.Q[level], log_bsgs_ratio=log_ratio))

    def GetLinearTransformRotationKeys(self, lt): return self.O.lt_galois_elements(self.lt.d[lt])
    def EvaluateLinearTransform(self, lt, ct): return self.ct.add(self.O.lt_apply(self.lt.d[lt], self.ct.d[ct]))

    # what the parity tests read
    def words(self, ct): return self.ct.d[ct].data
