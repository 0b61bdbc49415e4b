"""ResNet-20 (or another trace) with K images in flight per GPU (orion_b200.replay.BatchReplayer): s/image and images/s as a
function of K.   python tools/batch_images.py [trace] [K ...]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import BatchReplayer, Trace
name = sys.argv[1] if len(sys.argv) > 1 else "resnet20"
ks = [int(a) for a in sys.argv[2:]] or [1, 2, 4]
lib = raw.load()
G = os.path.join(ROOT, "tests", "golden")
t = Trace(os.path.join(G, "trace_%s.json" % name), os.path.join(G, "trace_%s.npz" % name))
clear = np.array(t.meta["clear_output"])
for k in ks:
    B = B200Library(); B.B200_SetSeed(42)
    r = BatchReplayer(B, t, k)
    r.setup(); lib.B200_Sync()
    outs = r.infer(); lib.B200_Sync()
    r.infer(); lib.B200_Sync()
    best = 1e9
    for _ in range(3):
        lib.B200_Sync(); lib.B200_TimerStart()
        outs = r.infer()
        best = min(best, lib.B200_TimerStopMs() / 1e3)
    same = all(np.array_equal(np.asarray(o), np.asarray(outs[0])) for o in outs)
    err = max(float(np.max(np.abs(np.asarray(o, dtype=np.float64)[: clear.size] - clear))) for o in outs)
    print("%s K=%d: %.4f s per batch, %.4f s/image, %.2f images/s, all images identical: %s, max err vs cleartext %.2e"
          % (name, k, best, best / k, k / best, same, err), flush=True)
    B.DeleteScheme()
