"""Generates tests/golden/ntt_kat.npz from the REFERENCE's own CPU NTT (GPU-NTT ntt_cpu.cu compiled in place into
oracle/_ref/libgpuntt_ref.so) on the in-tree known-answer (prime, psi) pools (nttparameters.cu:83-160).
Run in the build container (needs /root/reference):  python tests/golden/make_ntt_kat.py"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

R = O.ref()
assert R is not None, "oracle/_ref/libgpuntt_ref.so missing: run `make -C oracle ref`"
out = {}
rng = np.random.default_rng(20261104)
for logn in (10, 11, 12, 13):
    p, w, psi = ctypes.c_ulonglong(), ctypes.c_ulonglong(), ctypes.c_ulonglong()
    R.ref_ntt_pool(logn, p, w, psi)
    q = p.value
    n = 1 << logn
    a = rng.integers(0, q, size=n, dtype=np.uint64)
    a[:4] = [0, 1, q - 1, q // 2]
    f = np.zeros_like(a)
    R.ref_ntt_forward(logn, q, psi.value, O.P(a), O.P(f))
    i = np.zeros_like(a)
    R.ref_ntt_inverse(logn, q, psi.value, O.P(a), O.P(i))
    out["logn%d_prime_psi" % logn] = np.array([q, psi.value], dtype=np.uint64)
    out["logn%d_in" % logn] = a
    out["logn%d_fwd" % logn] = f
    out["logn%d_inv" % logn] = i
# schoolbook product from the reference (small n): pins NTT * pointwise * INTT
logn = 10
q, psi = [int(x) for x in out["logn10_prime_psi"]]
a = rng.integers(0, q, size=1 << logn, dtype=np.uint64)
b = rng.integers(0, q, size=1 << logn, dtype=np.uint64)
c = np.zeros_like(a)
R.ref_schoolbook(1 << logn, q, O.P(a), O.P(b), O.P(c))
out["school_a"], out["school_b"], out["school_c"] = a, b, c
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ntt_kat.npz"), **out)
print("wrote ntt_kat.npz", {k: v.shape for k, v in out.items()})
