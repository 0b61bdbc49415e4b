"""Bootstrap latency at the ResNet-20 chain for a given ring: python tools/boot16.py [logN=16] [slots=16384] [reps=3].
Prints the launch count before the timed bootstraps (for ncu -s) and per bootstrap (for ncu -c)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
from orion_b200.bindings import B200Library
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 16
slots = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
lib = raw.load()
B = B200Library(); B.B200_SetSeed(11)
B.NewScheme(logn, [60] + [30] * 32, [60, 60], 30, 192, "standard", "", "none")
B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
B.NewBootstrapper([61] * 8, slots); lib.B200_Sync()
full = 1 << (logn - 1)
rng = np.random.default_rng(8)
x = np.zeros(full); x[:slots] = rng.uniform(-1, 1, slots)          # zero-padded, as Orion hands it over
ct = B.Encrypt(B.B200_EncodeF64(x, full, 1, float(1 << 30)))
out = B.Bootstrap(ct, slots); lib.B200_Sync()                       # warm-up: DFT factors encoded here
re, im = np.empty(full), np.empty(full)
B.B200_DecodeF64(B.Decrypt(out), re, im)
err = float(np.max(np.abs(re[:slots] - x[:slots])))
lib.B200_Sync(); l0 = lib.B200_LaunchCount()
lib.B200_TimerStart()
for _ in range(reps):
    B.DeleteCiphertext(B.Bootstrap(ct, slots))
ms = lib.B200_TimerStopMs() / reps
print("logN=%d slots=%d: %.2f ms/bootstrap, max err %.2e (%.1f bits), launches before %d, per bootstrap %d, device %.1f GB" % (
    logn, slots, ms, err, -np.log2(err), l0, (lib.B200_LaunchCount() - l0) // reps, lib.B200_DeviceBytes(1) / 2**30))
