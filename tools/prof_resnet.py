"""Driver for profiling the ResNet-20 trace replay: setup + warm-up inference + one timed inference."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
lib = raw.load()
gold = os.path.join(ROOT, "tests", "golden")
B = B200Library(); B.B200_SetSeed(42)
name = os.environ.get("TRACE", "resnet20")             # TRACE=resnet20_logn16 for the logN=16 variant
t = Trace(os.path.join(gold, "trace_%s.json" % name), os.path.join(gold, "trace_%s.npz" % name))
r = Replayer(B, t, seed=1234)
r.setup(); lib.B200_Sync()
print("setup_launches", lib.B200_LaunchCount(), flush=True)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for i in range(n):
    l0 = lib.B200_LaunchCount(); t0 = time.perf_counter()
    r.infer(); lib.B200_Sync()
    print("infer %d: %.3f s, %d launches" % (i, time.perf_counter() - t0, lib.B200_LaunchCount() - l0), flush=True)
    if i == 0:
        import collections
        c = collections.Counter()
        print("op counts", dict(r.counts))

if "--by-op" in sys.argv:
    import collections
    tsum, lsum, cnt = collections.Counter(), collections.Counter(), collections.Counter()
    orig = r._run
    def timed(i, op):
        lib.B200_Sync(); l0 = lib.B200_LaunchCount(); t0 = time.perf_counter()
        orig(i, op)
        lib.B200_Sync()
        tsum[op[0]] += time.perf_counter() - t0; lsum[op[0]] += lib.B200_LaunchCount() - l0; cnt[op[0]] += 1
    r._run = timed
    r.infer()
    tot = sum(tsum.values())
    for k, v in tsum.most_common(14):
        print("%-26s n=%4d  %8.1f ms  %5.1f%%  %7d launches  (%.1f us/launch)" % (k, cnt[k], v * 1e3, 100 * v / tot, lsum[k], 1e6 * v / max(1, lsum[k])))
    print("total %.3f s" % tot)

if "--host-time" in sys.argv:
    # host-side enqueue time per op type (no synchronisation added): how far the launch rate is from the GPU time
    import collections
    tsum, cnt = collections.Counter(), collections.Counter()
    base_run = r._run if "--by-op" not in sys.argv else orig
    def timed(i, op):
        t0 = time.perf_counter()
        base_run(i, op)
        tsum[op[0]] += time.perf_counter() - t0; cnt[op[0]] += 1
    r._run = timed
    lib.B200_Sync(); t0 = time.perf_counter()
    r.infer(); lib.B200_Sync()
    wall = time.perf_counter() - t0
    for k, v in tsum.most_common(8):
        print("host %-26s n=%4d  %8.1f ms" % (k, cnt[k], v * 1e3))
    print("host total %.3f s of wall %.3f s (Decrypt/Decode include the final synchronisation)" % (sum(tsum.values()), wall))

if "--hostprof" in sys.argv:
    # ORION_B200_HOSTPROF=1 python tools/prof_resnet.py 2 --hostprof : C++-side enqueue timers for one inference
    lib.B200_Sync(); lib.B200_HostProfDump()
    t0 = time.perf_counter(); r.infer(); lib.B200_Sync()
    print("hostprof inference wall %.3f s" % (time.perf_counter() - t0), flush=True)
    lib.B200_HostProfDump()
