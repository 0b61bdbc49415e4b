"""CPU oracle for the orion-b200 CKKS hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (orion_b200) never imports it.

`lib()` loads oracle/libckks_oracle.so (built by `make -C oracle`); `ref()` loads the reference's own
CPU NTT compiled in place (oracle/_ref/libgpuntt_ref.so) and returns None when it is not there.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_ref = None

u64p = ctypes.POINTER(ctypes.c_uint64)
i64p = ctypes.POINTER(ctypes.c_int64)
intp = ctypes.POINTER(ctypes.c_int)
u32p = ctypes.POINTER(ctypes.c_uint32)
f64p = ctypes.POINTER(ctypes.c_double)
c_u64 = ctypes.c_uint64
c_int = ctypes.c_int
c_vp = ctypes.c_void_p


def build(force=False):
    so = os.path.join(_HERE, "libckks_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("ckks_oracle.c", "ckks_oracle_ckks.c", "ckks_oracle.h")]
    stale = force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "libckks_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference") and (force or not os.path.exists(os.path.join(_HERE, "_ref", "libgpuntt_ref.so"))):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if os.path.isdir("/root/reference") and (force or not os.path.exists(os.path.join(_HERE, "_ref", "libgpuntt_refgpu.so"))):
        subprocess.check_call(["make", "-C", _HERE, "refgpu"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if os.path.isdir("/root/reference") and (force or not os.path.exists(os.path.join(_HERE, "_ref", "orion_pyc.zip"))):
        # the reference's Python layer as bytecode, for the live drop-in test (tests/test_gpu_live.py)
        subprocess.check_call([sys.executable, os.path.join(_HERE, "build_refpy.py")], stdout=subprocess.DEVNULL)
    return so


def P(a, typ=u64p):
    """numpy array -> ctypes pointer (array must stay alive in the caller)."""
    return a.ctypes.data_as(typ)


def ints(xs):
    return (ctypes.c_int * len(xs))(*xs)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    so = build()
    L = ctypes.CDLL(so)
    sig = {
        "orc_is_prime": ([c_u64], c_int),
        "orc_powmod": ([c_u64, c_u64, c_u64], c_u64),
        "orc_invmod": ([c_u64, c_u64], c_u64),
        "orc_mulmod": ([c_u64, c_u64, c_u64], c_u64),
        "orc_gen_ntt_primes": ([c_int, c_u64, c_int, u64p, c_int, u64p], c_int),
        "orc_gen_moduli": ([intp, c_int, intp, c_int, c_int, u64p, u64p], c_int),
        "orc_primitive_root": ([c_u64], c_u64),
        "orc_psi": ([c_u64, c_u64], c_u64),
        "orc_ctx_new": ([c_int, u64p, c_int, c_int], c_vp),
        "orc_ctx_new_with_psi": ([c_int, u64p, u64p, c_int, c_int], c_vp),
        "orc_ctx_free": ([c_vp], None),
        "orc_ctx_N": ([c_vp], c_int),
        "orc_ctx_modulus": ([c_vp, c_int], c_u64),
        "orc_ctx_psi": ([c_vp, c_int], c_u64),
        "orc_set_threads": ([c_int], None),
        "orc_get_threads": ([], c_int),
        "orc_ntt": ([c_vp, c_int, u64p], None),
        "orc_intt": ([c_vp, c_int, u64p], None),
        "orc_ntt_many": ([c_vp, intp, c_int, u64p], None),
        "orc_intt_many": ([c_vp, intp, c_int, u64p], None),
        "orc_schoolbook_negacyclic": ([u64p, u64p, u64p, c_int, c_u64], None),
        "orc_add": ([c_vp, c_int, u64p, u64p, u64p], None),
        "orc_sub": ([c_vp, c_int, u64p, u64p, u64p], None),
        "orc_neg": ([c_vp, c_int, u64p, u64p], None),
        "orc_mul": ([c_vp, c_int, u64p, u64p, u64p], None),
        "orc_mul_scalar": ([c_vp, c_int, u64p, c_u64, u64p], None),
        "orc_add_scalar": ([c_vp, c_int, u64p, c_u64, u64p], None),
        "orc_automorphism_index_ntt": ([c_vp, c_u64, u32p], None),
        "orc_automorphism_ntt": ([c_vp, c_u64, u64p, u64p], None),
        "orc_automorphism_coeff": ([c_vp, c_int, c_u64, u64p, u64p], None),
        "orc_galois_element": ([c_vp, c_int], c_u64),
        "orc_galois_element_conj": ([c_vp], c_u64),
        "orc_baseconv": ([c_vp, intp, c_int, intp, c_int, u64p, u64p, c_int], None),
        "orc_rescale": ([c_vp, c_int, u64p, u64p], None),
        "orc_num_digits": ([c_vp, c_int], c_int),
        "orc_decompose": ([c_vp, c_int, u64p, u64p], None),
        "orc_inner_product": ([c_vp, c_int, u64p, u64p, u64p, u64p], None),
        "orc_moddown": ([c_vp, c_int, u64p, u64p], None),
        "orc_keyswitch": ([c_vp, c_int, u64p, u64p, u64p, u64p], None),
        "orc_rng_u64": ([c_u64, c_u64, c_u64], c_u64),
        "orc_sample_uniform": ([c_vp, c_int, c_u64, c_u64, u64p], None),
        "orc_sample_gaussian_i64": ([c_int, c_u64, c_u64, i64p], None),
        "orc_sample_ternary_hw": ([c_int, c_int, c_u64, c_u64, i64p], None),
        "orc_small_to_rns": ([c_vp, c_int, i64p, u64p], None),
        "orc_gaussian_cdt": ([u64p, c_int], c_int),
        "orc_roots": ([c_int, f64p, f64p], None),
        "orc_special_ifft": ([f64p, f64p, c_int, c_int], None),
        "orc_special_fft": ([f64p, f64p, c_int, c_int], None),
        "orc_encode_round": ([c_vp, f64p, f64p, c_int, ctypes.c_double, intp, c_int, u64p], None),
        "orc_scaled_const_rns": ([c_vp, ctypes.c_double, ctypes.c_double, ctypes.c_double, intp, c_int, u64p], None),
        "orc_decode_crt": ([c_vp, u64p, c_int, ctypes.c_double, c_int, f64p, f64p], None),
    }
    for name, (args, res) in sig.items():
        f = getattr(L, name)
        f.argtypes = args
        f.restype = res
    _lib = L
    return L


def ref():
    """The reference's own CPU NTT (GPU-NTT ntt_cpu.cu) compiled in place, or None."""
    global _ref
    if _ref is not None:
        return _ref
    so = os.path.join(_HERE, "_ref", "libgpuntt_ref.so")
    if not os.path.exists(so):
        return None
    try:
        R = ctypes.CDLL(so)
    except OSError:
        return None
    ull = ctypes.c_ulonglong
    ullp = ctypes.POINTER(ull)
    R.ref_ntt_pool.argtypes = [c_int, ullp, ullp, ullp]
    R.ref_ntt_forward.argtypes = [c_int, ull, ull, u64p, u64p]
    R.ref_ntt_inverse.argtypes = [c_int, ull, ull, u64p, u64p]
    R.ref_schoolbook.argtypes = [c_int, ull, u64p, u64p, u64p]
    for f in (R.ref_ntt_pool, R.ref_ntt_forward, R.ref_ntt_inverse, R.ref_schoolbook):
        f.restype = None
    _ref = R
    return R


_refgpu = None


def ref_gpu():
    """The reference's own GPU merge NTT (GPU-NTT ntt.cu, gpuntt::GPU_NTT_Inplace) compiled in place for sm_100a, or
    None.  Needs a CUDA device to be called; loading does not."""
    global _refgpu
    if _refgpu is not None:
        return _refgpu
    so = os.path.join(_HERE, "_ref", "libgpuntt_refgpu.so")
    if not os.path.exists(so):
        return None
    try:
        R = ctypes.CDLL(so)
    except OSError:
        return None
    R.refgpu_create.argtypes = [c_int, c_int, u64p, u64p]
    R.refgpu_create.restype = c_vp
    R.refgpu_destroy.argtypes = [c_vp]
    R.refgpu_destroy.restype = None
    R.refgpu_ntt_host.argtypes = [c_vp, u64p, u64p, c_int, c_int]
    R.refgpu_ntt_host.restype = None
    R.refgpu_time_ms.argtypes = [c_vp, c_int, c_int, c_int, c_int]
    R.refgpu_time_ms.restype = ctypes.c_float
    _refgpu = R
    return R


def gen_moduli(logq, logp, log_nthroot):
    L = lib()
    Q = np.zeros(len(logq), dtype=np.uint64)
    Pp = np.zeros(max(len(logp), 1), dtype=np.uint64)
    rc = L.orc_gen_moduli(ints(logq), len(logq), ints(logp), len(logp), log_nthroot, P(Q), P(Pp))
    assert rc == 0
    return [int(x) for x in Q], [int(x) for x in Pp[: len(logp)]]


class Ctx:
    """RAII handle on an orc_ctx."""

    def __init__(self, logN, moduli, nQ, nP, psi=None):
        self.L = lib()
        self.logN, self.N, self.nQ, self.nP = logN, 1 << logN, nQ, nP
        self.moduli = [int(m) for m in moduli]
        m = np.array(self.moduli, dtype=np.uint64)
        if psi is None:
            self.h = self.L.orc_ctx_new(logN, P(m), nQ, nP)
        else:
            ps = np.array(psi, dtype=np.uint64)
            self.h = self.L.orc_ctx_new_with_psi(logN, P(m), P(ps), nQ, nP)

    def __del__(self):
        try:
            if self.h:
                self.L.orc_ctx_free(self.h)
                self.h = None
        except Exception:
            pass

    # ---- numpy-level conveniences (each is one call into the C oracle) ----
    def ntt(self, a, mod_idx, inverse=False):
        a = np.array(a, dtype=np.uint64).reshape(len(mod_idx), self.N).copy()
        f = self.L.orc_intt_many if inverse else self.L.orc_ntt_many
        f(self.h, ints(mod_idx), len(mod_idx), P(a))
        return a

    def baseconv(self, a, src, dst, centered=False):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(src), self.N)
        out = np.empty((len(dst), self.N), dtype=np.uint64)
        self.L.orc_baseconv(self.h, ints(src), len(src), ints(dst), len(dst), P(a), P(out), 1 if centered else 0)
        return out

    def num_digits(self, level):
        return int(self.L.orc_num_digits(self.h, level))

    def galois_element(self, rot):
        return int(self.L.orc_galois_element(self.h, rot))

    def automorphism_ntt(self, a, gal_el):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, self.N)
        out = np.empty_like(a)
        for i in range(a.shape[0]):
            self.L.orc_automorphism_ntt(self.h, gal_el, P(a[i]), P(out[i]))
        return out

    def decompose(self, level, c1):
        c1 = np.ascontiguousarray(c1, dtype=np.uint64).reshape(level + 1, self.N)
        out = np.empty((self.num_digits(level), level + 1 + self.nP, self.N), dtype=np.uint64)
        self.L.orc_decompose(self.h, level, P(c1), P(out))
        return out

    def inner_product(self, level, digits, evk):
        W = level + 1 + self.nP
        a0 = np.empty((W, self.N), dtype=np.uint64)
        a1 = np.empty((W, self.N), dtype=np.uint64)
        digits = np.ascontiguousarray(digits, dtype=np.uint64)
        evk = np.ascontiguousarray(evk, dtype=np.uint64)
        self.L.orc_inner_product(self.h, level, P(digits), P(evk), P(a0), P(a1))
        return a0, a1

    def moddown(self, level, acc):
        acc = np.ascontiguousarray(acc, dtype=np.uint64).reshape(level + 1 + self.nP, self.N)
        out = np.empty((level + 1, self.N), dtype=np.uint64)
        self.L.orc_moddown(self.h, level, P(acc), P(out))
        return out

    def keyswitch(self, level, c1, evk):
        c1 = np.ascontiguousarray(c1, dtype=np.uint64).reshape(level + 1, self.N)
        evk = np.ascontiguousarray(evk, dtype=np.uint64)
        o0 = np.empty((level + 1, self.N), dtype=np.uint64)
        o1 = np.empty((level + 1, self.N), dtype=np.uint64)
        self.L.orc_keyswitch(self.h, level, P(c1), P(evk), P(o0), P(o1))
        return np.stack([o0, o1])

    def rescale(self, level, a):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(level + 1, self.N)
        out = np.empty((level, self.N), dtype=np.uint64)
        self.L.orc_rescale(self.h, level, P(a), P(out))
        return out

    def elementwise(self, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, self.N)
        out = np.empty_like(a)
        if b is not None:
            b = np.ascontiguousarray(b, dtype=np.uint64).reshape(-1, self.N)
        for i in range(a.shape[0]):
            if op == "neg":
                self.L.orc_neg(self.h, i, P(a[i]), P(out[i]))
            else:
                getattr(self.L, "orc_" + op)(self.h, i, P(a[i]), P(b[i]), P(out[i]))
        return out

    def uniform_poly(self, rng, mod_idx):
        """[len(mod_idx)][N] with limb i uniform in [0, q_{mod_idx[i]}) from a numpy Generator."""
        return np.stack([rng.integers(0, self.moduli[m], size=self.N, dtype=np.uint64) for m in mod_idx])
