"""Python binding of the Orion-facing C ABI (include/orion_b200_scheme.h): the drop-in `backend` object.

`B200Library` plays the role of `LattigoLibrary` (orion/backend/lattigo/bindings.py:91-746): it is what
`orion.core.orion.Scheme.setup_backend` (orion/core/orion.py:90-106) would instantiate for `backend: b200`, and it
exposes the same attribute names as the Go exports (NewScheme, Encode, Encrypt, Rotate, EvaluateLinearTransform,
EvaluatePolynomial, Bootstrap, ...) with the same Python-side conventions: lists in, ints / lists out, arrays
returned by the library freed with FreeCArray, serialized blobs returned as (numpy uint8 array, pointer).
Plus the extras the drifted Python layer needs from the heongpu adapter (CloneCiphertext, _ModDropCiphertext,
GenerateConsolidatedRotationKeys, GetPolyDepth, arithmeticoperator_handle).

There is no CPU path: every compute call aborts loudly in the library when no CUDA device is present.
"""
import ctypes

import numpy as np

from . import raw

c_int, c_long, c_ulong, c_float, c_double = ctypes.c_int, ctypes.c_long, ctypes.c_ulong, ctypes.c_float, ctypes.c_double
c_char_p, c_vp, c_u64 = ctypes.c_char_p, ctypes.c_void_p, ctypes.c_uint64
intp, floatp, ubytep = ctypes.POINTER(c_int), ctypes.POINTER(c_float), ctypes.POINTER(ctypes.c_ubyte)
u64p, i64p, f64p = ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(c_double)


class ArrayResultFloat(ctypes.Structure):
    _fields_ = [("Data", floatp), ("Length", c_ulong)]


class ArrayResultDouble(ctypes.Structure):
    _fields_ = [("Data", f64p), ("Length", c_ulong)]


class ArrayResultInt(ctypes.Structure):
    _fields_ = [("Data", intp), ("Length", c_ulong)]


class ArrayResultUInt64(ctypes.Structure):
    _fields_ = [("Data", u64p), ("Length", c_ulong)]


class ArrayResultByte(ctypes.Structure):
    _fields_ = [("Data", ubytep), ("Length", c_ulong)]


# name -> (argtypes, restype).  A pointer argtype consumes a Python list and is followed by its length (as in
# orion/backend/lattigo/bindings.py:54-62); the explicit length parameter is NOT listed here.
LIST = "list"
SCHEME_SIGNATURES = {
    "NewScheme": ([c_int, (intp, c_int), (intp, c_int), c_int, c_int, c_char_p, c_char_p, c_char_p], None),
    "DeleteScheme": ([], None),
    "FreeCArray": ([c_vp], None),
    "B200_SetSeed": ([c_u64], None),
    "DeletePlaintext": ([c_int], None),
    "DeleteCiphertext": ([c_int], None),
    "GetPlaintextScale": ([c_int], c_ulong),
    "GetCiphertextScale": ([c_int], c_ulong),
    "SetPlaintextScale": ([c_int, c_ulong], None),
    "SetCiphertextScale": ([c_int, c_ulong], None),
    "GetPlaintextLevel": ([c_int], c_int),
    "GetCiphertextLevel": ([c_long], c_int),
    "GetPlaintextSlots": ([c_long], c_int),
    "GetCiphertextSlots": ([c_long], c_int),
    "GetCiphertextDegree": ([c_long], c_int),
    "GetModuliChain": ([], ArrayResultUInt64),
    "GetLivePlaintexts": ([], ArrayResultInt),
    "GetLiveCiphertexts": ([], ArrayResultInt),
    "NewKeyGenerator": ([], None),
    "GenerateSecretKey": ([], None),
    "GeneratePublicKey": ([], None),
    "GenerateRelinearizationKey": ([], None),
    "GenerateEvaluationKeys": ([], None),
    "SerializeSecretKey": ([], ArrayResultByte),
    "LoadSecretKey": ([(ubytep, c_ulong)], None),
    "NewEncoder": ([], None),
    "Encode": ([(floatp, c_int), c_int, c_ulong], c_int),
    "Decode": ([c_int], ArrayResultFloat),
    "NewEncryptor": ([], None),
    "NewDecryptor": ([], None),
    "Encrypt": ([c_int], c_int),
    "Decrypt": ([c_int], c_int),
    "NewEvaluator": ([], None),
    "AddRotationKey": ([c_int], None),
    "Negate": ([c_int], c_int),
    "Rotate": ([c_int, c_int], c_int),
    "RotateNew": ([c_int, c_int], c_int),
    "Rescale": ([c_int], c_int),
    "RescaleNew": ([c_int], c_int),
    "AddScalar": ([c_int, c_float], c_int),
    "AddScalarNew": ([c_int, c_float], c_int),
    "SubScalar": ([c_int, c_float], c_int),
    "SubScalarNew": ([c_int, c_float], c_int),
    "MulScalarInt": ([c_int, c_int], c_int),
    "MulScalarIntNew": ([c_int, c_int], c_int),
    "MulScalarFloat": ([c_int, c_float], c_int),
    "MulScalarFloatNew": ([c_int, c_float], c_int),
    "AddPlaintext": ([c_int, c_int], c_int),
    "AddPlaintextNew": ([c_int, c_int], c_int),
    "SubPlaintext": ([c_int, c_int], c_int),
    "SubPlaintextNew": ([c_int, c_int], c_int),
    "MulPlaintext": ([c_int, c_int], c_int),
    "MulPlaintextNew": ([c_int, c_int], c_int),
    "AddCiphertext": ([c_int, c_int], c_int),
    "AddCiphertextNew": ([c_int, c_int], c_int),
    "SubCiphertext": ([c_int, c_int], c_int),
    "SubCiphertextNew": ([c_int, c_int], c_int),
    "MulRelinCiphertext": ([c_int, c_int], c_int),
    "MulRelinCiphertextNew": ([c_int, c_int], c_int),
    "CloneCiphertext": ([c_int], c_int),
    "ModDropCiphertext": ([c_int], c_int),
    "NewLinearTransformEvaluator": ([], None),
    "GenerateLinearTransform": ([(intp, c_int), (floatp, c_int), c_int, c_float, c_char_p], c_int),
    "EvaluateLinearTransform": ([c_int, c_int], c_int),
    "DeleteLinearTransform": ([c_int], None),
    "GetLinearTransformRotationKeys": ([c_int], ArrayResultInt),
    "GenerateLinearTransformRotationKey": ([c_int], None),
    "GenerateAndSerializeRotationKey": ([c_int], ArrayResultByte),
    "LoadRotationKey": ([(ubytep, c_ulong), c_ulong], None),
    "SerializeDiagonal": ([c_int, c_int], ArrayResultByte),
    "LoadPlaintextDiagonal": ([(ubytep, c_ulong), c_int, c_ulong], None),
    "RemovePlaintextDiagonals": ([c_int], None),
    "RemoveRotationKeys": ([], None),
    "GenerateConsolidatedRotationKeys": ([(intp, c_int)], None),
    "NewPolynomialEvaluator": ([], None),
    "GenerateMonomial": ([(floatp, c_int)], c_int),
    "GenerateChebyshev": ([(floatp, c_int)], c_int),
    "EvaluatePolynomial": ([c_int, c_int, c_ulong], c_int),
    "GetPolyDepth": ([c_int], c_int),
    "B200_GeneratePolynomialF64": ([f64p, c_int, c_int], c_int),
    "NewBootstrapper": ([(intp, c_int), c_int], None),
    "Bootstrap": ([c_int, c_int], c_int),
    "DeleteBootstrappers": ([], None),
    # additions
    "B200_ImportCiphertext": ([u64p, c_int, c_int, c_double, c_int], c_int),
    "B200_ExportCiphertext": ([c_int, u64p], None),
    "B200_CiphertextDevicePtr": ([c_int], c_vp),
    "B200_NewCiphertext": ([c_int, c_int, c_double, c_double, c_int], c_int),
    "B200_CiphertextAcquire": ([c_int, c_int], c_vp),
    "B200_GetCiphertextScaleParts": ([c_int, f64p, f64p], None),
    "B200_GetCiphertextLogSlots": ([c_int], c_int),
    "B200_GetCiphertextScaleF64": ([c_int], c_double),
    "B200_SetCiphertextScaleF64": ([c_int, c_double], None),
    "B200_ImportPlaintext": ([u64p, c_int, c_double], c_int),
    "B200_ExportPlaintext": ([c_int, u64p], None),
    "B200_ExportSecretKey": ([i64p], None),
    "B200_ImportSecretKey": ([i64p], None),
    "B200_ExportRelinKey": ([u64p], None),
    "B200_ExportGaloisKey": ([c_u64, u64p], None),
    "B200_ExportPublicKey": ([u64p], None),
    "B200_SchemeInfo": ([intp, intp, intp, u64p], c_int),
    "B200_EncodeF64": ([f64p, c_int, c_int, c_double], c_int),
    "B200_DecodeF64": ([c_int, f64p, f64p], None),
    "B200_Conjugate": ([c_int], c_int),
    "B200_NoteEvaluatorKey": ([c_u64], None),
}


def _flat_argtypes(spec):
    out = []
    for a in spec:
        if isinstance(a, tuple):
            out.extend(a)
        else:
            out.append(a)
    return out


class B200Function:
    """ctypes call with the list <-> (pointer, length) marshalling of LattigoFunction (lattigo/bindings.py:9-88)."""

    free = None

    def __init__(self, cfunc, spec, restype):
        self.cfunc, self.spec, self.restype = cfunc, spec, restype
        cfunc.argtypes = _flat_argtypes(spec)
        cfunc.restype = restype

    def __call__(self, *args):
        assert len(args) == len(self.spec), "%s expects %d arguments" % (self.cfunc.__name__, len(self.spec))
        cargs, keep = [], []
        for a, t in zip(args, self.spec):
            if isinstance(t, tuple):
                ptr_t, len_t = t
                if isinstance(a, np.ndarray):
                    arr = np.ascontiguousarray(a, dtype={intp: np.int32, floatp: np.float32, ubytep: np.uint8}[ptr_t])
                    keep.append(arr)
                    cargs += [arr.ctypes.data_as(ptr_t), len_t(arr.size)]
                else:
                    seq = list(a)
                    buf = (ptr_t._type_ * len(seq))(*seq)
                    keep.append(buf)
                    cargs += [buf, len_t(len(seq))]
            elif t is c_char_p:
                cargs.append(a.encode("utf-8") if isinstance(a, str) else a)
            elif t in (u64p, i64p, f64p) and isinstance(a, np.ndarray):
                cargs.append(a.ctypes.data_as(t))
            else:
                cargs.append(a)
        res = self.cfunc(*cargs)
        if isinstance(res, ArrayResultByte):
            buf = ctypes.cast(res.Data, ctypes.POINTER(ctypes.c_ubyte * res.Length)).contents
            return np.frombuffer(buf, dtype=np.uint8), res.Data          # caller frees later (deferred, as the reference)
        if isinstance(res, (ArrayResultFloat, ArrayResultDouble, ArrayResultInt, ArrayResultUInt64)):
            n = int(res.Length)
            dt = {ArrayResultFloat: np.float32, ArrayResultDouble: np.float64, ArrayResultInt: np.int32,
                  ArrayResultUInt64: np.uint64}[type(res)]
            out = np.ctypeslib.as_array(res.Data, shape=(n,)).astype(dt).tolist() if n else []
            B200Function.free(ctypes.cast(res.Data, c_vp))
            return out
        return res


class B200Library:
    """The `backend` object for `orion.backend: b200` (see INTEGRATION.md for the one-branch hook in orion)."""

    def __init__(self):
        self.lib = raw.load()
        self.arithmeticoperator_handle = None       # attribute the drifted Python layer passes around (PY/evaluator.py:36-40)
        for name, (spec, res) in SCHEME_SIGNATURES.items():
            setattr(self, name, B200Function(getattr(self.lib, name), spec, res))   # a missing export fails loudly here
        # PY/poly_evaluator.py:23 calls GenerateChebyshev(coeffs, len(coeffs)) (heongpu-adapter signature): accept both
        cheb = self.GenerateChebyshev
        self.GenerateChebyshev = lambda coeffs, *_n: cheb(coeffs)
        B200Function.free = self.lib.FreeCArray
        self.lib.FreeCArray.argtypes = [c_vp]
        self.lib.FreeCArray.restype = None

    # same entry point as LattigoLibrary.setup_bindings (lattigo/bindings.py:126-179)
    def setup_bindings(self, orion_params):
        self.NewScheme(orion_params.get_logn(), orion_params.get_logq(), orion_params.get_logp(),
                       orion_params.get_logscale(), orion_params.get_hamming_weight(), orion_params.get_ringtype(),
                       orion_params.get_keys_path(), orion_params.get_io_mode())

    # host-only: composite minimax sign polynomials (lattigo/polyeval.go:91-167); same arguments, flat list out
    def GenerateMinimaxSignCoeffs(self, degrees, prec=128, logalpha=12, logerr=12, debug=0):
        from .minimax import generate_minimax_sign_coeffs
        return generate_minimax_sign_coeffs(list(degrees), prec, logalpha, logerr, bool(debug))

    # extras of the heongpu adapter used by the backend-agnostic wrappers
    def _ModDropCiphertext(self, _op_handle, ct, _opts=None):
        return self.ModDropCiphertext(ct)
