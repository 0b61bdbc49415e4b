"""Small driver for ncu: forward + inverse NTT over 192 limbs at logN=16 (the key-switch digit NTT shape)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
lib = raw.load()
LOGN = 16
Q, P = raw.gen_moduli([60] + [45] * 23, [61, 61, 61], LOGN + 1)
ctx = raw.RawContext(LOGN, Q, P)
nl = 192
mods = raw._ints([i % 27 for i in range(nl)])
bufs = [lib.B200_DevAlloc(nl * ctx.N) for _ in range(3)]
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    lib.B200_NTTDev(ctx.h, bufs[it % 3], bufs[it % 3], mods, nl, 0)
    lib.B200_NTTDev(ctx.h, bufs[it % 3], bufs[it % 3], mods, nl, 1)
lib.B200_Sync()
print("done", lib.B200_LaunchCount())
