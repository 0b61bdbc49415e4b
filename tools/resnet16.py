"""ResNet-20 at logN=16 (BASELINE configs[3]: configs/resnet.yml with LogN forced to 16; 32768 slots, sparse-slot bootstraps at
16384 / 8192 / 4096) through the drop-in boundary, in a process of its own (bench.py runs it as a child: 160 GB of keys, encoded
diagonals and bootstrap factors).  usage: resnet16.py [reps] [seed].  Prints one JSON line: seconds per image (encode + encrypt of
the image and decrypt + decode of the logits inside the timed region, CUDA events on the library stream), numerics against the
recorded cleartext logits, launches, SURVEY 8(d) algorithmic bytes, memory high-water mark, bytes dropped under memory pressure."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 42
lib = raw.load()
B = B200Library(); B.B200_SetSeed(seed)
gold = os.path.join(ROOT, "tests", "golden")
t = Trace(os.path.join(gold, "trace_resnet20_logn16.json"), os.path.join(gold, "trace_resnet20_logn16.npz"))
r = Replayer(B, t, seed=1234)
t0 = time.perf_counter(); r.setup(); lib.B200_Sync(); setup_s = time.perf_counter() - t0
clear = np.array(t.meta["clear_output"])
out = np.asarray(r.infer(), dtype=np.float64)[: clear.size]; lib.B200_Sync()           # warm-up (plans, pool growth, DFT factor encodings)
times = []
l0, a0 = lib.B200_LaunchCount(), lib.B200_AlgorithmicBytes()
for i in range(reps):
    lib.B200_Sync(); lib.B200_TimerStart()
    out = np.asarray(r.infer(), dtype=np.float64)[: clear.size]
    times.append(lib.B200_TimerStopMs() / 1e3)
n_calls = reps + 1
print(json.dumps({
    "s_per_image": min(times), "times": times, "setup_s": setup_s, "launches_per_image": int((lib.B200_LaunchCount() - l0) // reps),
    "algorithmic_bytes_per_image": (lib.B200_AlgorithmicBytes() - a0) / reps,
    "counts": {k: r.counts.get(k, 0) // n_calls for k in ("EvaluateLinearTransform", "EvaluatePolynomial", "Bootstrap", "MulRelinCiphertextNew", "RotateNew")},
    "numerics": {"mae_vs_cleartext_network": float(np.mean(np.abs(out - clear))), "max_abs_err": float(np.max(np.abs(out - clear))),
                 "argmax_equal": bool(int(np.argmax(out)) == int(np.argmax(clear)))},
    "peak_device_gb": lib.B200_DeviceBytes(1) / 2**30, "reclaimed_under_pressure_gb": lib.B200_ReclaimedBytes() / 2**30}))
