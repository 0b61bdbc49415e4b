"""Debug helper (not a test): replay a trace on the GPU and on the oracle in lockstep; print the first op whose result differs."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)
from oracle_backend import OracleBackend
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
name = sys.argv[1] if len(sys.argv) > 1 else "mlp"
t = Trace(os.path.join(HERE, "golden", "trace_%s.json" % name), os.path.join(HERE, "golden", "trace_%s.npz" % name))
logN = t.meta["config"]["ckks_params"]["LogN"]
B = B200Library(); B.B200_SetSeed(42)
O = OracleBackend(42)
rg, ro = Replayer(B, t, seed=1234), Replayer(O, t, seed=1234)

def gwords(ct):
    level, deg = B.GetCiphertextLevel(ct), B.GetCiphertextDegree(ct)
    out = np.empty((deg + 1, level + 1, 1 << logN), dtype=np.uint64)
    B.B200_ExportCiphertext(ct, out)
    return out

bad = 0
for i, op in enumerate(t.ops):
    rg._run(i, op); ro._run(i, op)
    nm, args, ret = op
    if nm == "Encode":
        lv = B.GetPlaintextLevel(rg.pt[ret])
        gp = np.empty((lv + 1, 1 << logN), dtype=np.uint64); B.B200_ExportPlaintext(rg.pt[ret], gp)
        print(i, "Encode", ret, "plaintext same:", np.array_equal(gp, O.pt.d[ro.pt[ret]].data))
    if nm == "Encrypt":
        g, o = gwords(rg.ct[ret]), O.words(ro.ct[ret])
        print(i, "Encrypt: c0 same", np.array_equal(g[0], o[0]), "c1 same", np.array_equal(g[1], o[1]), "oracle events so far", O.O.next_event,
              "oracle galois keys", len(O.O.galois_keys))
    if ret is not None and ret in rg.ct and ret in ro.ct and nm not in ("Encode", "Decrypt", "GenerateLinearTransform") and "Plaintext" not in nm[:6]:
        try:
            g, o = gwords(rg.ct[ret]), O.words(ro.ct[ret])
        except Exception as e:
            continue
        same = g.shape == o.shape and np.array_equal(g, o)
        gs, os_ = B.GetCiphertextScale(rg.ct[ret]), float(ro.B.ct.d[ro.ct[ret]].scale)
        if not same or bad:
            print(i, nm, str(args)[:60], "->", ret, "same" if same else "DIFF", g.shape, o.shape, "scale", gs, os_)
            bad += 1
            if bad > 6: break
print("done, mismatches:", bad)
