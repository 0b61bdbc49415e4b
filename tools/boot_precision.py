"""Bootstrap precision and latency at the shipped ResNet-20 parameter set (configs/resnet.yml: logN=13, LogQ 60+32x30,
LogP 60x2, LogScale 30, H=192, boot LogP 8x61) for the CoeffsToSlots / SlotsToCoeffs depths given by
ORION_B200_CTS_DEPTH / ORION_B200_STC_DEPTH (defaults 4 / 3)."""
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
t0 = time.perf_counter(); B.NewBootstrapper([61] * 8, slots); lib.B200_Sync(); t_new = time.perf_counter() - t0
rng = np.random.default_rng(8)
x = rng.uniform(-1, 1, slots)
pt = B.B200_EncodeF64(x, slots, 1, float(1 << 30))
ct = B.Encrypt(pt)
errs = []
for i in range(3):
    out = B.Bootstrap(ct, slots)
    p = B.Decrypt(out)
    re, im = np.empty(slots), np.empty(slots)
    B.B200_DecodeF64(p, re, im)
    errs.append(float(np.max(np.abs(re - x))))
lib.B200_Sync()
l0 = lib.B200_LaunchCount(); t0 = time.perf_counter()
n = 8
for i in range(n):
    out = B.Bootstrap(ct, slots); B.DeleteCiphertext(out)
lib.B200_Sync()
dt = (time.perf_counter() - t0) / n
print("cts=%s stc=%s: max err %.3e (-log2 %.1f), mean err %.3e, %.2f ms/bootstrap, %d launches, level out %d, NewBootstrapper %.1f s" % (
    os.environ.get("ORION_B200_CTS_DEPTH", "4"), os.environ.get("ORION_B200_STC_DEPTH", "3"), max(errs), -np.log2(max(errs)),
    float(np.mean(np.abs(re - x))), dt * 1e3, (lib.B200_LaunchCount() - l0) // n, B.GetCiphertextLevel(out) if False else len(logq) - 1, t_new))
