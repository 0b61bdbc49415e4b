"""ResNet-20 at logN=16 (BASELINE configs[3]) through the drop-in boundary: setup, REPS inferences, numerics against the recorded cleartext
logits, memory high-water mark and how much regenerable state was dropped under memory pressure.  One JSON line."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
lib = raw.load()
B = B200Library(); B.B200_SetSeed(42)
gold = os.path.join(ROOT, "tests", "golden")
t = Trace(os.path.join(gold, "trace_resnet20_logn16.json"), os.path.join(gold, "trace_resnet20_logn16.npz"))
r = Replayer(B, t)
t0 = time.perf_counter(); r.setup(); lib.B200_Sync(); setup_s = time.perf_counter() - t0
print("setup %.1f s, device bytes %.1f GB (peak %.1f), reclaimed %.1f GB" % (setup_s, lib.B200_DeviceBytes(0) / 2**30, lib.B200_DeviceBytes(1) / 2**30,
                                                                             lib.B200_ReclaimedBytes() / 2**30), file=sys.stderr, flush=True)
clear = np.array(t.meta["clear_output"])
times, outs = [], []
for i in range(reps):
    rc0 = lib.B200_ReclaimedBytes(); l0 = lib.B200_LaunchCount()
    lib.B200_Sync(); t0 = time.perf_counter()
    out = np.asarray(r.infer(), dtype=np.float64)[: clear.size]
    lib.B200_Sync(); dt = time.perf_counter() - t0
    times.append(dt); outs.append(out)
    print("inference %d: %.2f s, %d launches, max|FHE - clear| %.2e, reclaimed %.1f GB, device bytes %.1f GB" % (
        i, dt, lib.B200_LaunchCount() - l0, float(np.max(np.abs(out - clear))), (lib.B200_ReclaimedBytes() - rc0) / 2**30, lib.B200_DeviceBytes(0) / 2**30),
        file=sys.stderr, flush=True)
print(json.dumps({"s_per_image": min(times), "times": times, "setup_s": setup_s, "max_abs_err_vs_cleartext": float(np.max(np.abs(outs[-1] - clear))),
                  "mae": float(np.mean(np.abs(outs[-1] - clear))), "argmax_equal": bool(int(np.argmax(outs[-1])) == int(np.argmax(clear))),
                  "peak_device_gb": lib.B200_DeviceBytes(1) / 2**30, "reclaimed_gb": lib.B200_ReclaimedBytes() / 2**30,
                  "counts": {k: v // reps for k, v in r.counts.items() if k in ("Bootstrap", "EvaluateLinearTransform", "EvaluatePolynomial", "RotateNew")}}))
