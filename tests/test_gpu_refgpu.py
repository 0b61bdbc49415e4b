"""The reference's OWN GPU NTT on the B200, beside the product's.

`oracle/_ref/libgpuntt_refgpu.so` is GPU-NTT's merge NTT (`gpuntt::GPU_NTT_Inplace`, RNS overload,
/root/reference/orion/backend/heongpu/adrianHEonGPU/thirdparty/GPU-NTT/src/lib/ntt_merge/ntt.cu, declared at
src/include/ntt_merge/ntt.cuh:198-201) compiled in place for sm_100a by `make -C oracle refgpu` -- the one piece of the
reference's GPU path that builds without network.  Two uses:
  * parity: the product's NTT/INTT words must equal the reference GPU kernel's on the same (prime, psi) chain;
  * measurement: the "reference GPU kernel to beat" of SURVEY.md section 8(d), CUDA events in place of nvbench.
The measurement is written to gpurun_out/ref_gpu_ntt.json (copied to profiles/ by hand).
"""
import ctypes
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [  # (logN, LogQ, LogP)
    (13, [29, 26, 26, 26, 26, 26], [29, 29]),          # configs/mlp.yml
    (14, [60, 30, 30, 30, 30, 30, 30], [60, 60]),      # configs/lola2.yml
    (16, [60, 46, 45, 47, 46], [61, 46]),              # both sides of the FP64 / integer butterfly split
]


def _refgpu(orc):
    R = orc.ref_gpu()
    if R is None:
        pytest.skip("oracle/_ref/libgpuntt_refgpu.so not built (make -C oracle refgpu needs /root/reference)")
    return R


def _u64(xs):
    return np.array([int(x) for x in xs], dtype=np.uint64)


@pytest.mark.parametrize("case", CASES)
def test_ntt_words_equal_reference_gpu_kernel(orc, b200lib, case):
    R = _refgpu(orc)
    logN, lq, lp = case
    Q, P = b200lib.gen_moduli(lq, lp, logN + 1)
    gc = b200lib.RawContext(logN, Q, P)
    mods = list(Q) + list(P)
    nmod, N = len(mods), 1 << logN
    primes, psis = _u64(mods), _u64([gc.psi(i) for i in range(nmod)])
    h = R.refgpu_create(logN, nmod, orc.P(primes), orc.P(psis))
    rng = np.random.default_rng(11)
    batch = 2 * nmod                                   # two polynomials per modulus, modulus order cycling
    a = np.stack([rng.integers(0, mods[i % nmod], size=N, dtype=np.uint64) for i in range(batch)])
    a[0, :4] = [0, mods[0] - 1, 1, mods[0] - 2]         # edge residues
    idx = [i % nmod for i in range(batch)]
    for inverse in (0, 1):
        ref = np.empty_like(a)
        R.refgpu_ntt_host(h, orc.P(a), orc.P(ref), batch, inverse)
        got = gc.ntt(a, idx, inverse=bool(inverse))
        assert np.array_equal(got, ref), "product NTT differs from the reference GPU kernel (inverse=%d)" % inverse
    R.refgpu_destroy(h)


def test_reference_gpu_ntt_speed_beside_ours(orc, b200lib):
    """Same transform, same chain, same batch as bench.py's ntt_roofline: 189 limbs at logN=16 (the converted digit limbs
    of one key switch at L=24, k=3), inputs cycled over 3 x 99 MB > L2.  Best of three runs per side; asserts only that ours
    is not slower (measured: 3.7x faster at logN=16, 2.6-3.2x at logN=13)."""
    R = _refgpu(orc)
    lib = b200lib.load()
    out = {}
    for logN, lq, lp, nl in ((16, [60] + [45] * 23, [61] * 3, 189), (13, [60] + [30] * 32, [60, 60], 35 * 16)):
        Q, P = b200lib.gen_moduli(lq, lp, logN + 1)
        gc = b200lib.RawContext(logN, Q, P)
        mods = list(Q) + list(P)
        nmod, N = len(mods), 1 << logN
        primes, psis = _u64(mods), _u64([gc.psi(i) for i in range(nmod)])
        h = R.refgpu_create(logN, nmod, orc.P(primes), orc.P(psis))
        nbuf = max(3, int(300e6 // (nl * N * 8)) + 1)
        row = {"logN": logN, "limbs": nl, "bytes_algorithmic": 16 * N * nl, "buffers_cycled": nbuf}
        for inverse, name in ((0, "ntt"), (1, "intt")):
            midx = (ctypes.c_int * nl)(*[i % nmod for i in range(nl)])
            bufs = [lib.B200_DevAlloc(nl * N) for _ in range(nbuf)]
            host = np.random.default_rng(5).integers(0, 1 << 28, size=nl * N, dtype=np.uint64)
            for b in bufs:
                lib.B200_Upload(b, b200lib._p(host), nl * N)
            ref_ms, our_ms = float("inf"), float("inf")
            for _ in range(3):                                   # best of three, alternating: a noisy box must not decide
                ref_ms = min(ref_ms, float(R.refgpu_time_ms(h, nl, inverse, 24, nbuf)))
                for i in range(3):
                    lib.B200_NTTDev(gc.h, bufs[i % nbuf], bufs[i % nbuf], midx, nl, inverse)
                lib.B200_Sync()
                lib.B200_TimerStart()
                for i in range(24):
                    lib.B200_NTTDev(gc.h, bufs[i % nbuf], bufs[i % nbuf], midx, nl, inverse)
                our_ms = min(our_ms, float(lib.B200_TimerStopMs()) / 24)
            for b in bufs:
                lib.B200_DevFree(b)
            gb = 16 * N * nl / 1e9
            row[name] = {"reference_gpu_ms": ref_ms, "ours_ms": our_ms, "reference_gpu_GBps": gb / (ref_ms * 1e-3),
                         "ours_GBps": gb / (our_ms * 1e-3), "speedup": ref_ms / our_ms}
            assert our_ms <= ref_ms * 1.25, (name, logN, our_ms, ref_ms)
        R.refgpu_destroy(h)
        gc.close()
        out["logN%d" % logN] = row
    print(json.dumps(out))
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "ref_gpu_ntt.json"), "w") as f:
            json.dump(out, f, indent=1)
