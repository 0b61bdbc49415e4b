"""Is the small-N regime CPU-launch-bound or GPU-latency-bound?  Enqueue time vs completion time of a chain of
key switches / rescales at logN=13 (the ResNet-20 shipped parameter set)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
lib = raw.load()
logN = 13
Q, P = raw.gen_moduli([60] + [30] * 32, [60, 60], logN + 1)
ctx = raw.RawContext(logN, Q, P)
N, L, K = ctx.N, len(Q), len(P)
level, nmod, beta = L - 1, L + K, ctx.num_digits(L - 1)
rng = np.random.default_rng(1)
evk = ctx.evk_to_device(rng.integers(0, 1 << 29, size=beta * 2 * nmod * N, dtype=np.uint64))
LN = L * N
a = lib.B200_DevAlloc(2 * LN); b = lib.B200_DevAlloc(2 * LN)
lib.B200_Upload(a, raw._p(rng.integers(0, 1 << 29, size=2 * LN, dtype=np.uint64)), 2 * LN)
for name, fn, reps in (("keyswitch L=33", lambda: lib.B200_KeySwitchDev(ctx.h, level, a + LN * 8, evk, a, b), 300),
                       ("rescale L=33", lambda: lib.B200_RescaleDev(ctx.h, level, a, b, 2), 300)):
    for _ in range(20): fn()
    lib.B200_Sync()
    l0 = lib.B200_LaunchCount(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    t1 = time.perf_counter(); lib.B200_Sync(); t2 = time.perf_counter()
    nl = lib.B200_LaunchCount() - l0
    print("%-16s %d launches/op: enqueue %.1f us/op (%.2f us/launch), complete %.1f us/op (%.2f us/launch)" %
          (name, nl // reps, 1e6 * (t1 - t0) / reps, 1e6 * (t1 - t0) / nl, 1e6 * (t2 - t0) / reps, 1e6 * (t2 - t0) / nl))
