"""Cluster (one launch) vs two-launch NTT/INTT at a small ring: time per transform of n limbs, back to back on one stream
(a dependent chain, as in the ResNet-20 replay).  Run with ORION_B200_CLUSTER_NTT=0/1."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
lib = raw.load()
logN = int(sys.argv[1]) if len(sys.argv) > 1 else 13
Q, P = raw.gen_moduli([60] + [30] * 32, [61] * 8, logN + 1)
ctx = raw.RawContext(logN, Q, P)
N, nmod = ctx.N, len(Q) + len(P)
out = []
for nl in (2, 8, 16, 48, 96, 144, 288, 864):
    midx = (ctypes.c_int * nl)(*[i % nmod for i in range(nl)])
    buf = lib.B200_DevAlloc(nl * N)
    host = np.random.default_rng(5).integers(0, 1 << 28, size=nl * N, dtype=np.uint64)
    lib.B200_Upload(buf, raw._p(host), nl * N)
    row = [nl]
    for inverse in (0, 1):
        for i in range(5):
            lib.B200_NTTDev(ctx.h, buf, buf, midx, nl, inverse)
        lib.B200_Sync()
        reps = 200
        lib.B200_TimerStart()
        for i in range(reps):
            lib.B200_NTTDev(ctx.h, buf, buf, midx, nl, inverse)
        row.append(1e3 * float(lib.B200_TimerStopMs()) / reps)
    lib.B200_DevFree(buf)
    out.append(row)
print("cluster=%s logN=%d  limbs: us/NTT us/INTT" % (os.environ.get("ORION_B200_CLUSTER_NTT", "1"), logN))
for r in out:
    print("  %4d  %7.2f  %7.2f" % tuple(r))
