"""CUDA-graph replay of Bootstrap against the eager path: same words, and the time per call (ResNet-20 chain, logN=13)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
from orion_b200.bindings import B200Library
lib = raw.load()
B = B200Library(); B.B200_SetSeed(11)
logq = [60] + [30] * 32
B.NewScheme(13, logq, [60, 60], 30, 192, "standard", "", "none")
B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
slots = 4096
B.NewBootstrapper([61] * 8, slots); lib.B200_Sync()
rng = np.random.default_rng(8)
N = 8192
def words(ct):
    lvl = B.GetCiphertextLevel(ct)
    out = np.empty(2 * (lvl + 1) * N, dtype=np.uint64)
    B.B200_ExportCiphertext(ct, out)
    return out
outs = []
cts = [B.Encrypt(B.B200_EncodeF64(rng.uniform(-1, 1, slots), slots, 1, float(1 << 30))) for _ in range(2)]
for rep in range(12):                      # call 0 eager; then every lane the scheduler picks captures once and replays; two inputs
    outs.append([words(B.Bootstrap(c, slots)) for c in cts])
for rep in range(1, 12):
    for k in range(2):
        assert np.array_equal(outs[rep][k], outs[0][k]), "graph replay %d differs from the eager bootstrap (input %d)" % (rep, k)
assert not np.array_equal(outs[0][0], outs[0][1])
lib.B200_Sync(); l0 = lib.B200_LaunchCount(); t0 = time.perf_counter()
n = 16
for i in range(n):
    B.DeleteCiphertext(B.Bootstrap(cts[i & 1], slots))
t_host = time.perf_counter() - t0
lib.B200_Sync()
dt = (time.perf_counter() - t0) / n
print("graphs=%s: words identical over eager/capture/replay; %.2f ms/bootstrap (host enqueue %.2f ms), %d launches counted per call" % (
    os.environ.get("ORION_B200_GRAPHS", "default"), dt * 1e3, t_host / n * 1e3, (lib.B200_LaunchCount() - l0) // n))
