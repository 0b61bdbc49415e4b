"""Diagnostic: per-inference wall time and the slowest single backend call (host time) of the ResNet-20 replay."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
lib = raw.load()
gold = os.path.join(ROOT, "tests", "golden")
B = B200Library(); B.B200_SetSeed(42)
t = Trace(os.path.join(gold, "trace_resnet20.json"), os.path.join(gold, "trace_resnet20.npz"))
r = Replayer(B, t, seed=1234)
r.setup(); lib.B200_Sync()
orig = r._run
worst = [0.0, None, 0]
def timed(i, op):
    t0 = time.perf_counter(); orig(i, op); dt = time.perf_counter() - t0
    if dt > worst[0]: worst[0], worst[1], worst[2] = dt, op[0], i
r._run = timed
import ctypes
free = ctypes.c_size_t(); tot = ctypes.c_size_t()
cudart = ctypes.CDLL("libcudart.so.12")
for k in range(int(sys.argv[1]) if len(sys.argv) > 1 else 8):
    worst[0] = 0.0
    t0 = time.perf_counter(); r.infer(); lib.B200_Sync(); dt = time.perf_counter() - t0
    cudart.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(tot))
    print("infer %d: %.3f s; slowest call %s #%d %.1f ms; device memory in use %.2f GB" % (k, dt, worst[1], worst[2], worst[0] * 1e3, (tot.value - free.value) / 2**30), flush=True)
