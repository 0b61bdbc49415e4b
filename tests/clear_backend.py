"""A cleartext stand-in for the backend object -- TEST INFRASTRUCTURE ONLY.

`ClearBackend` implements the reference's backend-call names (orion/backend/lattigo/*.go) on plain
float64 slot vectors: a "ciphertext" is the vector it would decrypt to.  Replaying a recorded trace through it with the same
seeded payloads computes, in the clear, exactly the function the encrypted replay computes -- the cleartext side of the
reference's own end-to-end assertion (tests/models/test_mlp.py:45-48).  (The ResNet-20 trace carries payload SIZES only; with
random weights its activations leave the [-1, 1] domain of the sign polynomials in the clear as well, so that trace is a
structural and timing workload; its numerics are covered per operation: test_gpu_scheme.py::test_bootstrap, test_polynomial_evaluation.)
Semantics per call: linear transform = sum_d diag_d * roll(x, -d) (LG/lineartransform.go:96-113), rotation by r =
roll(x, -r), polynomial = Chebyshev series on [-1, 1] / monomial (LG/polyeval.go:38-84), rescale / bootstrap / level and
scale bookkeeping = identity."""
import numpy as np
from numpy.polynomial import chebyshev as C


class ClearBackend:
    """Handles are integers into one object table (the replayer compares handles)."""

    def __init__(self):
        self.slots, self.obj, self.next_id = None, {}, 0

    def NewScheme(self, logN, logQ, logP, logScale, h, ringType, keysPath, ioMode):
        self.slots = 1 << (logN - 1 if ringType.lower() == "standard" else logN)

    def __getattr__(self, name):                     # New*, Generate*Key*, Set*Scale, AddRotationKey, ...: nothing to do
        if name.startswith(("New", "Generate", "Set", "AddRotationKey", "Remove", "Load")):
            return lambda *a: None
        raise AttributeError(name)

    def _new(self, x):
        self.next_id += 1
        self.obj[self.next_id] = x
        return self.next_id

    def _set(self, h, x):
        self.obj[h] = x
        return h

    def DeletePlaintext(self, h): self.obj.pop(h, None)
    DeleteCiphertext = DeleteLinearTransform = DeletePlaintext

    def Encode(self, values, level, scale):
        v = np.zeros(self.slots)
        a = np.asarray(values, dtype=np.float32).astype(np.float64)
        v[: a.size] = a
        return self._new(v)

    def Encrypt(self, pt): return self._new(self.obj[pt].copy())
    def Decrypt(self, ct): return self._new(self.obj[ct].copy())
    def Decode(self, pt): return self.obj[pt].astype(np.float32).tolist()

    def GenerateLinearTransform(self, diag_idx, flat, level, ratio, io_mode):
        flat = np.asarray(flat, dtype=np.float32).astype(np.float64).reshape(len(diag_idx), self.slots)
        return self._new({int(d) & (self.slots - 1): flat[i] for i, d in enumerate(diag_idx)})

    def EvaluateLinearTransform(self, lt, ct):
        x, out = self.obj[ct], np.zeros(self.slots)
        for d, v in self.obj[lt].items():
            out += v * np.roll(x, -d)
        return self._new(out)

    def GenerateChebyshev(self, coeffs, *_): return self._new(("cheb", np.asarray(coeffs, dtype=np.float32).astype(np.float64)))
    def GenerateMonomial(self, coeffs, *_): return self._new(("mono", np.asarray(coeffs, dtype=np.float32).astype(np.float64)))

    def EvaluatePolynomial(self, ct, poly, out_scale):
        kind, c = self.obj[poly]
        x = self.obj[ct]
        return self._new(C.chebval(x, c) if kind == "cheb" else np.polynomial.polynomial.polyval(x, c))

    def Bootstrap(self, ct, slots): return self._new(self.obj[ct].copy())
    def Rescale(self, ct): return ct
    def RescaleNew(self, ct): return self._new(self.obj[ct].copy())
    def CloneCiphertext(self, ct): return self._new(self.obj[ct].copy())

    # (name, in place?, function of the operand vectors / scalar)
    def _unary(fn, inplace):
        def f(self, ct, *a):
            r = fn(self.obj[ct], *a)
            return self._set(ct, r) if inplace else self._new(r)
        return f

    def _binary(fn, inplace):
        def f(self, a, b):
            r = fn(self.obj[a], self.obj[b])
            return self._set(a, r) if inplace else self._new(r)
        return f

    _f32 = lambda c: float(np.float32(c))
    Negate = _unary(lambda x: -x, False)
    Rotate, RotateNew = _unary(lambda x, r: np.roll(x, -r), True), _unary(lambda x, r: np.roll(x, -r), False)
    AddScalar, AddScalarNew = _unary(lambda x, c: x + float(np.float32(c)), True), _unary(lambda x, c: x + float(np.float32(c)), False)
    SubScalar, SubScalarNew = _unary(lambda x, c: x - float(np.float32(c)), True), _unary(lambda x, c: x - float(np.float32(c)), False)
    MulScalarInt, MulScalarIntNew = _unary(lambda x, k: x * int(k), True), _unary(lambda x, k: x * int(k), False)
    MulScalarFloat, MulScalarFloatNew = _unary(lambda x, c: x * float(np.float32(c)), True), _unary(lambda x, c: x * float(np.float32(c)), False)
    AddPlaintext = AddCiphertext = _binary(lambda x, y: x + y, True)
    AddPlaintextNew = AddCiphertextNew = _binary(lambda x, y: x + y, False)
    SubPlaintext = SubCiphertext = _binary(lambda x, y: x - y, True)
    SubPlaintextNew = SubCiphertextNew = _binary(lambda x, y: x - y, False)
    MulPlaintext = MulRelinCiphertext = _binary(lambda x, y: x * y, True)
    MulPlaintextNew = MulRelinCiphertextNew = _binary(lambda x, y: x * y, False)
