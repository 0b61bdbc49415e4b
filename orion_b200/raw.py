"""ctypes binding of the primitive-level C ABI (B200_*) declared in include/orion_b200.h.

No torch, no oracle: this module only marshals numpy buffers into liborion_b200.so.  Loading the library
works without a GPU (symbol checks); any compute call without a CUDA device aborts loudly in the library.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liborion_b200.so")

c_int, c_u64, c_vp, c_sz = ctypes.c_int, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_size_t
u64p = ctypes.POINTER(ctypes.c_uint64)
intp = ctypes.POINTER(ctypes.c_int)

RAW_SIGNATURES = {
    "B200_GenModuli": ([intp, c_int, intp, c_int, c_int, u64p, u64p], c_int),
    "B200_FindPsi": ([c_u64, c_u64], c_u64),
    "B200_DeviceCount": ([], c_int),
    "B200_NewContext": ([c_int, u64p, c_int, u64p, c_int], c_vp),
    "B200_DeleteContext": ([c_vp], None),
    "B200_ContextPsi": ([c_vp, c_int], c_u64),
    "B200_GaloisElement": ([c_vp, c_int], c_u64),
    "B200_Sync": ([], None),
    "B200_LaunchCount": ([], ctypes.c_ulonglong),
    "B200_AlgorithmicBytes": ([], ctypes.c_double),
    "B200_ReclaimedBytes": ([], ctypes.c_double),
    "B200_TrimPool": ([], None),
    "B200_DeviceBytes": ([c_int], ctypes.c_double),
    "B200_Stream": ([], c_vp),
    "B200_TimerStart": ([], None),
    "B200_TimerStopMs": ([], ctypes.c_float),
    "B200_HostProfDump": ([], None),
    "B200_ProbeEnable": ([ctypes.c_int], None),
    "B200_ProbeReadMs": ([ctypes.POINTER(ctypes.c_long)], ctypes.c_double),
    "B200_ProbeSlots": ([], c_int),
    "B200_ProbeSlotName": ([c_int], ctypes.c_char_p),
    "B200_ProbeReadSlotMs": ([c_int, ctypes.POINTER(ctypes.c_long)], ctypes.c_double),
    "B200_DevAlloc": ([c_sz], c_vp),
    "B200_DevFree": ([c_vp], None),
    "B200_Upload": ([c_vp, u64p, c_sz], None),
    "B200_UploadAsync": ([c_vp, u64p, c_sz], None),
    "B200_Download": ([u64p, c_vp, c_sz], None),
    "B200_HostAllocPinned": ([c_sz], c_vp),
    "B200_HostFreePinned": ([c_vp], None),
    "B200_NTT": ([c_vp, u64p, u64p, intp, c_int, c_int], None),
    "B200_NTTDev": ([c_vp, c_vp, c_vp, intp, c_int, c_int], None),
    "B200_BaseConv": ([c_vp, intp, c_int, intp, c_int, u64p, u64p, c_int], None),
    "B200_AutomorphismNTT": ([c_vp, c_u64, u64p, u64p, c_int], None),
    "B200_AutomorphismNTTDev": ([c_vp, c_u64, c_vp, c_vp, c_int], None),
    "B200_EvkToDevice": ([c_vp, u64p], c_vp),
    "B200_KeySwitch": ([c_vp, c_int, u64p, c_vp, u64p, u64p], None),
    "B200_KeySwitchDev": ([c_vp, c_int, c_vp, c_vp, c_vp, c_vp], None),
    "B200_KeySwitchBatch": ([c_vp, c_int, c_int, ctypes.POINTER(c_vp), c_int, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp)], None),
    "B200_KeySwitchBatchDev": ([c_vp, c_int, c_int, c_vp, c_int, ctypes.POINTER(c_vp), c_vp], None),
    "B200_Decompose": ([c_vp, c_int, u64p, u64p], None),
    "B200_ModDown": ([c_vp, c_int, u64p, u64p], None),
    "B200_Rescale": ([c_vp, c_int, u64p, u64p, c_int], None),
    "B200_RescaleDev": ([c_vp, c_int, c_vp, c_vp, c_int], None),
    "B200_RescaleDevNoAlloc": ([c_vp, c_int, c_vp, c_vp, c_int], c_int),
    "B200_NoAllocMaxWords": ([], c_sz),
    "B200_Elementwise": ([c_vp, c_int, u64p, u64p, u64p, c_int], None),
}

_lib = None


def load():
    """dlopen liborion_b200.so and declare every B200_* signature."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("liborion_b200.so is not built: run `python -m orion_b200.build` "
                           "(there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (args, res) in RAW_SIGNATURES.items():
        f = getattr(lib, name)
        f.argtypes = args
        f.restype = res
    _lib = lib
    return lib


def _p(a):
    return a.ctypes.data_as(u64p)


def _ints(xs):
    return (ctypes.c_int * len(xs))(*[int(x) for x in xs])


def gen_moduli(logq, logp, log_nthroot):
    lib = load()
    q = np.zeros(len(logq), dtype=np.uint64)
    p = np.zeros(max(1, len(logp)), dtype=np.uint64)
    lib.B200_GenModuli(_ints(logq), len(logq), _ints(logp), len(logp), log_nthroot, _p(q), _p(p))
    return [int(x) for x in q], [int(x) for x in p[: len(logp)]]


class RawContext:
    """A modulus chain with its tables resident in HBM, plus the primitive operations on host numpy buffers."""

    def __init__(self, logN, Q, P):
        self.lib = load()
        self.logN, self.N, self.nQ, self.nP = logN, 1 << logN, len(Q), len(P)
        self.Q, self.P = [int(x) for x in Q], [int(x) for x in P]
        q = np.array(self.Q, dtype=np.uint64)
        p = np.array(self.P if self.P else [0], dtype=np.uint64)
        self.h = self.lib.B200_NewContext(logN, _p(q), self.nQ, _p(p), self.nP)

    def close(self):
        if getattr(self, "h", None):
            self.lib.B200_DeleteContext(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def num_digits(self, level):
        return (level + 1 + self.nP - 1) // self.nP

    def psi(self, idx):
        return int(self.lib.B200_ContextPsi(self.h, idx))

    def galois_element(self, rot):
        return int(self.lib.B200_GaloisElement(self.h, rot))

    def ntt(self, a, mod_idx, inverse=False):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(mod_idx), self.N)
        out = np.empty_like(a)
        self.lib.B200_NTT(self.h, _p(a), _p(out), _ints(mod_idx), len(mod_idx), 1 if inverse else 0)
        return out

    def baseconv(self, a, src, dst, centered=False):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(src), self.N)
        out = np.empty((len(dst), self.N), dtype=np.uint64)
        self.lib.B200_BaseConv(self.h, _ints(src), len(src), _ints(dst), len(dst), _p(a), _p(out), 1 if centered else 0)
        return out

    def automorphism_ntt(self, a, gal_el):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n = a.size // self.N
        out = np.empty_like(a)
        self.lib.B200_AutomorphismNTT(self.h, gal_el, _p(a), _p(out), n)
        return out

    def evk_to_device(self, evk):
        evk = np.ascontiguousarray(evk, dtype=np.uint64)
        beta = self.num_digits(self.nQ - 1)
        assert evk.size == beta * 2 * (self.nQ + self.nP) * self.N
        return self.lib.B200_EvkToDevice(self.h, _p(evk))

    def keyswitch(self, level, c1, evk_dev, c0=None):
        c1 = np.ascontiguousarray(c1, dtype=np.uint64).reshape(level + 1, self.N)
        out = np.empty((2, level + 1, self.N), dtype=np.uint64)
        c0p = None
        if c0 is not None:
            c0 = np.ascontiguousarray(c0, dtype=np.uint64).reshape(level + 1, self.N)
            c0p = _p(c0)
        self.lib.B200_KeySwitch(self.h, level, _p(c1), evk_dev, c0p, _p(out))
        return out

    def keyswitch_batch(self, level, cts, evk_devs, has_c0=True):
        """Pipelined batch on host arrays: cts[i] is [2][L][N] (c0, c1) if has_c0 else [L][N]; returns a list of [2][L][N]."""
        n = len(cts)
        cts = [np.ascontiguousarray(c, dtype=np.uint64) for c in cts]
        outs = [np.empty((2, level + 1, self.N), dtype=np.uint64) for _ in range(n)]
        arr = ctypes.c_void_p * n
        self.lib.B200_KeySwitchBatch(self.h, level, n, arr(*[c.ctypes.data for c in cts]), 1 if has_c0 else 0,
                                     arr(*evk_devs), arr(*[o.ctypes.data for o in outs]))
        return outs

    def keyswitch_batch_dev(self, level, cts, evk_devs, has_c0=True):
        """Device-resident batch entry point (what bench.py times): one launch sequence for all ciphertexts.
        cts: [n][2][L][N] (c0, c1) if has_c0 else [n][L][N]; returns [n][2][L][N]."""
        cts = np.ascontiguousarray(cts, dtype=np.uint64)
        n = cts.shape[0]
        assert cts.size == n * (2 if has_c0 else 1) * (level + 1) * self.N
        out = np.empty((n, 2, level + 1, self.N), dtype=np.uint64)
        d_in, d_out = self.lib.B200_DevAlloc(cts.size), self.lib.B200_DevAlloc(out.size)
        self.lib.B200_Upload(d_in, _p(cts), cts.size)
        self.lib.B200_KeySwitchBatchDev(self.h, level, n, d_in, 1 if has_c0 else 0, (ctypes.c_void_p * n)(*evk_devs), d_out)
        self.lib.B200_Download(_p(out), d_out, out.size)
        self.lib.B200_DevFree(d_in); self.lib.B200_DevFree(d_out)
        return out

    def decompose(self, level, c1):
        c1 = np.ascontiguousarray(c1, dtype=np.uint64).reshape(level + 1, self.N)
        out = np.empty((self.num_digits(level), level + 1 + self.nP, self.N), dtype=np.uint64)
        self.lib.B200_Decompose(self.h, level, _p(c1), _p(out))
        return out

    def moddown(self, level, acc):
        acc = np.ascontiguousarray(acc, dtype=np.uint64).reshape(level + 1 + self.nP, self.N)
        out = np.empty((level + 1, self.N), dtype=np.uint64)
        self.lib.B200_ModDown(self.h, level, _p(acc), _p(out))
        return out

    def rescale(self, level, a, polys=1):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(polys, level + 1, self.N)
        out = np.empty((polys, level, self.N), dtype=np.uint64)
        self.lib.B200_Rescale(self.h, level, _p(a), _p(out), polys)
        return out

    def elementwise(self, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n = a.size // self.N
        out = np.empty_like(a)
        bp = None
        if b is not None:
            b = np.ascontiguousarray(b, dtype=np.uint64)
            bp = _p(b)
        self.lib.B200_Elementwise(self.h, {"add": 0, "sub": 1, "neg": 2, "mul": 3}[op], _p(a), bp, _p(out), n)
        return out
