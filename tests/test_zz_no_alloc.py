"""B200_RescaleDevNoAlloc (csrc/noalloc.cu): the allocation-free rescale -- scratch and launch descriptors in one static
device array -- against the existing B200_RescaleDev, through the C ABI, on device buffers the caller owns.

No counterpart in the reference.  Collected last on purpose:
the file name sorts after every other test file.  Tolerance: the one the integer tests use -- bit-exact.
Passes alone in a fresh process and after the rest of the suite: it creates its own context and touches no global state
but the shared scratch array, which needs no initial contents."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dev(lib, raw, a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    p = lib.B200_DevAlloc(a.size)
    lib.B200_Upload(p, raw._p(a), a.size)
    return p


def _down(lib, raw, p, shape):
    out = np.empty(shape, dtype=np.uint64)
    lib.B200_Download(raw._p(out), p, out.size)
    return out


@pytest.mark.parametrize("case", [(13, [55, 40, 40, 40, 40], [61, 61]),           # small ring, FP64-path primes
                                  (16, [60, 46, 45, 52, 47, 46], [61]),           # large ring, both arithmetic paths
                                  (14, [60] + [30] * 20, [60, 60])])              # many limbs
def test_no_alloc_rescale_equals_the_pooled_rescale_twice_in_a_row(b200lib, orc, case):
    raw = b200lib
    lib = raw.load()
    logN, lq, lp = case
    Q, P = raw.gen_moduli(lq, lp, logN + 1)
    gc = raw.RawContext(logN, Q, P)
    oc = orc.Ctx(logN, Q + P, len(Q), len(P))
    N = gc.N
    rng = np.random.default_rng(21)
    assert lib.B200_NoAllocMaxWords() == 1 << 22
    for call, (level, polys) in enumerate([(len(Q) - 1, 2), (len(Q) - 1, 2), (2, 1), (len(Q) - 2, 2)]):
        # two calls in a row at the same shape with DIFFERENT inputs (the second must not see the first's scratch), then
        # other shapes over the same array
        a = np.stack([oc.uniform_poly(rng, list(range(level + 1))) for _ in range(polys)])
        d_in = _dev(lib, raw, a)
        d_ref, d_new = lib.B200_DevAlloc(polys * level * N), lib.B200_DevAlloc(polys * level * N)
        lib.B200_RescaleDev(gc.h, level, d_in, d_ref, polys)
        rc = lib.B200_RescaleDevNoAlloc(gc.h, level, d_in, d_new, polys)
        assert rc == 0
        ref = _down(lib, raw, d_ref, (polys, level, N))
        new = _down(lib, raw, d_new, (polys, level, N))
        assert np.array_equal(new, ref), "call %d (level %d, polys %d)" % (call, level, polys)
        for p in range(polys):                                                    # and both are the oracle's rescale
            assert np.array_equal(new[p], oc.rescale(level, a[p]))
        # the input is left untouched
        assert np.array_equal(_down(lib, raw, d_in, a.shape), a)
        for p in (d_in, d_ref, d_new):
            lib.B200_DevFree(p)
    gc.close()


def test_no_alloc_rescale_rejects_what_does_not_fit(b200lib):
    """polys * (level + 1) * N beyond the static array, or arguments out of range: error code, nothing launched, output
    untouched."""
    raw = b200lib
    lib = raw.load()
    logN, lq, lp = 16, [60] + [45] * 39, [61]
    Q, P = raw.gen_moduli(lq, lp, logN + 1)
    gc = raw.RawContext(logN, Q, P)
    N, level, polys = gc.N, 39, 2                                                 # 2 * 40 * 65536 = 5.2 M words > 2^22
    assert polys * (level + 1) * N > lib.B200_NoAllocMaxWords()
    d_in = lib.B200_DevAlloc(polys * (level + 1) * N)
    sentinel = np.full(polys * level * N, 0xA5A5A5A5A5A5A5A5, dtype=np.uint64)
    d_out = _dev(lib, raw, sentinel)
    l0 = lib.B200_LaunchCount()
    assert lib.B200_RescaleDevNoAlloc(gc.h, level, d_in, d_out, polys) == -1
    assert lib.B200_RescaleDevNoAlloc(gc.h, 0, d_in, d_out, 1) == -1              # cannot rescale at level 0
    assert lib.B200_RescaleDevNoAlloc(gc.h, 3, d_in, d_out, 3) == -1              # 1 or 2 polynomials
    assert lib.B200_LaunchCount() == l0
    assert np.array_equal(_down(lib, raw, d_out, sentinel.shape), sentinel)
    # the largest shape that fits at this ring: 2 polynomials of 32 limbs
    assert lib.B200_RescaleDevNoAlloc(gc.h, 31, d_in, d_out, 2) == 0
    lib.B200_Sync()
    lib.B200_DevFree(d_in); lib.B200_DevFree(d_out)
    gc.close()
