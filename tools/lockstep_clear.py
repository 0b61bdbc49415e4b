"""Debug helper (not a test): replay a trace on the GPU and on plain float vectors (tests/clear_backend.py) in lockstep and
print, per ciphertext-producing call, how far the decrypted GPU result is from the cleartext one.

  python tools/lockstep_clear.py resnet20 [max_reports]
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from clear_backend import ClearBackend
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
name = sys.argv[1] if len(sys.argv) > 1 else "resnet20"
max_rep = int(sys.argv[2]) if len(sys.argv) > 2 else 40
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-3
G = os.path.join(ROOT, "tests", "golden")
t = Trace(os.path.join(G, "trace_%s.json" % name), os.path.join(G, "trace_%s.npz" % name))
B = B200Library(); B.B200_SetSeed(42)
C = ClearBackend()
rg, rc = Replayer(B, t, seed=1234), Replayer(C, t, seed=1234)


def gvals(ct):
    p = B.Decrypt(ct)
    v = np.asarray(B.Decode(p), dtype=np.float64)
    B.DeletePlaintext(p)
    return v


rep = 0
worst_so_far = 0.0
for i, op in enumerate(t.ops):
    nm, args, ret = op
    rg._run(i, op); rc._run(i, op)
    if i < t.n_compile or ret is None or ret not in rg.ct or ret not in rc.ct:
        continue
    if nm in ("Encode", "Decrypt", "Decode", "GenerateLinearTransform", "GenerateChebyshev", "GenerateMonomial"):
        continue
    try:
        g = gvals(rg.ct[ret])
    except Exception as e:
        print(i, nm, "decrypt failed", e); continue
    c = np.asarray(C.obj[rc.ct[ret]], dtype=np.float64)
    err = float(np.max(np.abs(g - c)))
    lvl, sc = B.GetCiphertextLevel(rg.ct[ret]), B.GetCiphertextScale(rg.ct[ret])
    if err > thr or err > 4 * worst_so_far:
        print("%5d %-24s %-28s -> ct %d  level %2d  log2(scale) %.3f  max|clear| %.4g  max err %.3e" %
              (i, nm, str(args)[:28], ret, lvl, np.log2(float(sc)), float(np.max(np.abs(c))), err))
        rep += err > thr
        worst_so_far = max(worst_so_far, err)
        if rep >= max_rep:
            break
print("done")
