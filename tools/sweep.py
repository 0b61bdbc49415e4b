#!/usr/bin/env python
"""Primitive sweep (BASELINE.json configs[4]): NTT, INTT, automorphism, rescale, key switch (single
and batched) on device-resident operands at logN = 13..16, L in {6, 12, 24} Q-limbs (45-bit, first 60-bit), k = 3
special primes; hoisted BSGS linear transform with 128 diagonals and bootstrap through the Orion-facing C ABI.
Times are CUDA events on the library stream after warm-up; GB/s = algorithmic bytes of SURVEY.md 8(d) / time.

  python tools/sweep.py [--quick] [--no-scheme] > profiles/rN_sweep.txt
"""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import raw  # noqa: E402

lib = raw.load()
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0


def timeit(fn, reps):
    for _ in range(3):
        fn()
    lib.B200_Sync()
    lib.B200_TimerStart()
    for _ in range(reps):
        fn()
    return lib.B200_TimerStopMs() * 1e3 / reps          # us


def row(name, logN, L, batch, us, nbytes):
    """us = time per op (per ciphertext for batched calls); nbytes = algorithmic bytes per op."""
    gbs = nbytes / (us * 1e-6) / 1e9
    print("%-26s logN=%2d L=%2d batch=%3d  %9.1f us/op  %9.0f op/s  %7.0f GB/s algorithmic  %.2f of measured HBM" %
          (name, logN, L, batch, us, 1e6 / us, gbs, gbs / PEAK), flush=True)


def cpu_row(logN, Q, P, L, K):
    """The CPU oracle (OpenMP, all host cores) on the same key switch: a bounded sample, reported beside the GPU rows."""
    import oracle
    oc = oracle.Ctx(logN, list(Q) + list(P), len(Q), len(P))
    oc.L.orc_set_threads(os.cpu_count() or 1)
    rng = np.random.default_rng(1234)
    nmod, level = L + K, L - 1
    beta = oc.num_digits(level)
    evk = rng.integers(0, 1 << 44, size=(beta, 2, nmod, oc.N), dtype=np.uint64)
    c1 = oc.uniform_poly(rng, list(range(L)))
    oc.keyswitch(level, c1, evk)
    n, t0 = 0, time.perf_counter()
    while n < 16 and time.perf_counter() - t0 < 2.0:
        oc.keyswitch(level, c1, evk); n += 1
    us = (time.perf_counter() - t0) / n * 1e6
    print("%-26s logN=%2d L=%2d batch=  1  %9.1f us/op  %9.1f op/s  (CPU oracle, %d threads, %d key switches timed)" %
          ("key switch, CPU oracle", logN, L, us, 1e6 / us, os.cpu_count() or 1, n), flush=True)


cpu_rows = False


def primitives(logN, L, K=3):
    Q, P = raw.gen_moduli([60] + [45] * (L - 1), [61] * K, logN + 1)
    ctx = raw.RawContext(logN, Q, P)
    N = ctx.N
    level, nmod, beta = L - 1, L + K, ctx.num_digits(L - 1)
    LN = L * N
    rng = np.random.default_rng(1234)
    nrot = 3                                               # rotate buffers so consecutive reps do not hit L2-resident data only
    bufs = [lib.B200_DevAlloc(2 * LN) for _ in range(nrot)]
    outs = [lib.B200_DevAlloc(2 * LN) for _ in range(nrot)]
    h = rng.integers(0, 1 << 44, size=2 * LN, dtype=np.uint64)
    for b in bufs:
        lib.B200_Upload(b, raw._p(h), 2 * LN)
    mods = raw._ints(list(range(L)) * 2)
    st = {"i": 0}

    def nxt():
        st["i"] = (st["i"] + 1) % nrot
        return st["i"]
    reps = 30
    us = timeit(lambda: lib.B200_NTTDev(ctx.h, bufs[nxt()], outs[st["i"]], mods, 2 * L, 0), reps)
    row("NTT (ciphertext, 2L limbs)", logN, L, 1, us, 16 * N * 2 * L)
    us = timeit(lambda: lib.B200_NTTDev(ctx.h, bufs[nxt()], outs[st["i"]], mods, 2 * L, 1), reps)
    row("INTT (ciphertext)", logN, L, 1, us, 16 * N * 2 * L)
    us = timeit(lambda: lib.B200_RescaleDev(ctx.h, level, bufs[nxt()], outs[st["i"]], 2), reps)
    row("rescale", logN, L, 1, us, 8 * N * (2 * L + 2 * (L - 1)))
    gal = ctx.galois_element(1)
    us = timeit(lambda: lib.B200_AutomorphismNTTDev(ctx.h, gal, bufs[nxt()], outs[st["i"]], 2 * L), reps)
    row("automorphism (NTT domain)", logN, L, 1, us, 2 * 2 * L * N * 8)
    evk = ctx.evk_to_device(rng.integers(0, 1 << 44, size=beta * 2 * nmod * N, dtype=np.uint64))
    ks_bytes = 8 * N * (4 * L + 2 * beta * (L + K))
    us = timeit(lambda: lib.B200_KeySwitchDev(ctx.h, level, bufs[nxt()] + LN * 8, evk, bufs[st["i"]], outs[st["i"]]), reps)
    row("key switch (= rotation)", logN, L, 1, us, ks_bytes)
    if cpu_rows:
        cpu_row(logN, Q, P, L, K)
    for batch in (8, 64, 256):
        if batch * 4 * LN * 8 > (60 << 30):
            continue
        big_in = lib.B200_DevAlloc(batch * 2 * LN); big_out = lib.B200_DevAlloc(batch * 2 * LN)
        for b in range(batch):
            lib.B200_Upload(big_in + b * 2 * LN * 8, raw._p(h), 2 * LN)
        ek = (ctypes.c_void_p * batch)(*([evk] * batch))
        us = timeit(lambda: lib.B200_KeySwitchBatchDev(ctx.h, level, batch, big_in, 1, ek, big_out), max(3, reps // batch))
        row("key switch, batched", logN, L, batch, us / batch, ks_bytes)
        lib.B200_DevFree(big_in); lib.B200_DevFree(big_out)
    for b in bufs + outs + [evk]:
        lib.B200_DevFree(b)
    ctx.close()


def scheme_level(logN, L, K=3, nd=128, boot=True):
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.B200_SetSeed(7)
    logq = [55] + [40] * (L - 1)
    B.NewScheme(logN, logq, [61] * K, 40, 192, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator(); B.NewLinearTransformEvaluator()
    N = 1 << logN
    slots = N // 2
    rng = np.random.default_rng(5)
    level = L - 1
    x = rng.uniform(-1, 1, slots)
    ct = B.Encrypt(B.B200_EncodeF64(x, slots, level, float(1 << 40)))
    diag_idx = list(range(nd))
    flat = rng.uniform(-1, 1, nd * slots).astype(np.float32)
    t0 = time.perf_counter()
    lt = B.GenerateLinearTransform(diag_idx, flat, level, 2.0, "none")
    for g in sorted(B.GetLinearTransformRotationKeys(lt)):
        B.GenerateLinearTransformRotationKey(g)
    lib.B200_Sync()
    setup = time.perf_counter() - t0
    outs = []

    def run():
        outs.append(B.EvaluateLinearTransform(lt, ct))
        if len(outs) > 2:
            B.DeleteCiphertext(outs.pop(0))
    us = timeit(run, 5)
    beta = (L + K - 1) // K
    n1 = int(round(np.sqrt(nd)))
    lt_bytes = 8 * N * (2 * L + nd * (L + K) + 2 * n1 * 2 * beta * (L + K) + 2 * L)
    row("hoisted BSGS LT, %d diagonals (setup %.1f s)" % (nd, setup), logN, L, 1, us, lt_bytes)
    B.DeleteLinearTransform(lt)
    if boot:
        t0 = time.perf_counter()
        B.NewBootstrapper([61] * 4, slots)
        lib.B200_Sync()
        setup = time.perf_counter() - t0
        low = B.Encrypt(B.B200_EncodeF64(x, slots, 1, float(1 << 40)))
        res = []

        def runb():
            res.append(B.Bootstrap(low, slots))
            if len(res) > 2:
                B.DeleteCiphertext(res.pop(0))
        l0 = lib.B200_LaunchCount()
        us = timeit(runb, 3)
        nl = (lib.B200_LaunchCount() - l0) // 6
        print("%-26s logN=%2d L=%2d batch=  1  %9.1f us/op  (%.2f ms; %d launches; bootstrapper setup %.1f s; full slots)" %
              ("bootstrap", logN, L, us, us / 1e3, nl, setup), flush=True)
        B.DeleteBootstrappers()
    B.DeleteScheme()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--no-scheme", action="store_true")
    ap.add_argument("--cpu", action="store_true", help="time the CPU oracle's key switch beside every (logN, L) row group")
    a = ap.parse_args()
    global cpu_rows
    cpu_rows = a.cpu
    print("# primitive sweep on %s; measured HBM peak %.0f GB/s; times are CUDA-event averages" % (time.strftime("%Y-%m-%d"), PEAK))
    logNs = (13, 16) if a.quick else (13, 14, 15, 16)
    Ls = (24,) if a.quick else (6, 12, 24)
    for logN in logNs:
        for L in Ls:
            primitives(logN, L)
    if not a.no_scheme:
        for logN in ((13,) if a.quick else (13, 14, 15, 16)):
            scheme_level(logN, 12, boot=True)


if __name__ == "__main__":
    main()
