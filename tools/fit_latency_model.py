"""Re-fit Orion's bootstrap-placement latency model to B200 timings.

The reference's placer (orion/core/level_dag.py:164-249) weighs a bootstrap as  t_boot = a * exp(b * l_eff) + c  seconds
(a, b, c = 3.41, 0.18, 4.81: its own CPU fit) and a linear transform as  alpha * #diagonals * level  (alpha = 0.001).
This tool measures both on the device through the drop-in boundary at the shipped ResNet-20 ring (logN = 13, 30-bit
levels, boot LogP 8 x 61) and prints constants in the same form, so a maintainer can paste them next to the CPU ones.
Latencies are per call with a synchronisation after each call (what the placer's critical path sees)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orion_b200 import raw
from orion_b200.bindings import B200Library

lib = raw.load()
LOGN, SLOTS = 13, 4096
rng = np.random.default_rng(3)


def scheme(l_eff):
    B = B200Library(); B.B200_SetSeed(5)
    B.NewScheme(LOGN, [60] + [30] * l_eff, [60, 60], 30, 192, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator(); B.NewLinearTransformEvaluator()
    return B


def timed(fn, reps):
    lib.B200_Sync(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); lib.B200_Sync(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts))


boot, lt = [], []
for l_eff in (8, 16, 24, 32):
    B = scheme(l_eff)
    B.NewBootstrapper([61] * 8, SLOTS)
    x = rng.uniform(-1, 1, SLOTS).astype(np.float32)
    ct = B.Encrypt(B.Encode(x.tolist(), l_eff, 1 << 30))

    def one():
        o = B.Bootstrap(ct, SLOTS); B.DeleteCiphertext(o)
    one(); one()
    boot.append((l_eff, timed(one, 5)))
    if l_eff == 32:
        for level in (4, 12, 20, 28):
            for nd in (16, 64):
                idx = list(range(nd))
                flat = rng.uniform(-1, 1, nd * SLOTS).astype(np.float32)
                t = B.GenerateLinearTransform(idx, flat, level, 2.0, "none")
                for g in B.GetLinearTransformRotationKeys(t):
                    B.GenerateLinearTransformRotationKey(g)
                c = B.Encrypt(B.Encode(x.tolist(), level, 1 << 30))

                def ev():
                    o = B.EvaluateLinearTransform(t, c); B.DeleteCiphertext(o)
                ev(); ev()
                lt.append((level, nd, timed(ev, 7)))
                B.DeleteLinearTransform(t); B.DeleteCiphertext(c)
    B.DeleteBootstrappers(); B.DeleteScheme()

le = np.array([b[0] for b in boot], float); tb = np.array([b[1] for b in boot])
lin = np.polyfit(le, tb, 1)
out = {"ring": "logN=13, LogQ 60+30xl_eff, LogP 60x2, boot LogP 61x8, full slots", "bootstrap_s": {int(l): t for l, t in boot},
       "bootstrap_linear_fit_s": {"per_level": float(lin[0]), "const": float(lin[1])},
       "linear_transform_s": [{"level": l, "diagonals": nd, "s": t} for l, nd, t in lt]}
try:
    from scipy.optimize import curve_fit
    (a, b, c), _ = curve_fit(lambda l, a, b, c: a * np.exp(b * l) + c, le, tb, p0=(1e-3, 0.05, 5e-3), maxfev=20000)
    out["bootstrap_exp_fit_s"] = {"a": float(a), "b": float(b), "c": float(c), "reference_cpu": {"a": 3.41, "b": 0.18, "c": 4.81}}
except Exception as e:
    out["bootstrap_exp_fit_s"] = "fit failed: %s" % e
A = np.array([[l * nd, 1.0] for l, nd, _ in lt]); y = np.array([t for _, _, t in lt])
alpha, const = np.linalg.lstsq(A, y, rcond=None)[0]
out["linear_transform_fit_s"] = {"alpha_per_diagonal_level": float(alpha), "const": float(const), "reference_cpu_alpha": 0.001}
print(json.dumps(out, indent=1))
d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
if os.path.isdir(d):
    json.dump(out, open(os.path.join(d, "latency_model_b200.json"), "w"), indent=1)
