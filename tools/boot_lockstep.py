"""Debug helper (not a test): one bootstrap on the GPU with ORION_B200_BOOT_DUMP and the same on oracle/bootstrap.py;
prints the first stage whose ciphertext words (or host tables) differ.   python tools/boot_lockstep.py [num_slots]"""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
d = tempfile.mkdtemp()
os.environ["ORION_B200_BOOT_DUMP"] = d
from oracle.bootstrap import OBootstrapper
from oracle.scheme import OScheme
from orion_b200.bindings import B200Library
num_slots = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
logn, logq, logp, logscale, h, seed = 12, [55, 40, 40, 40, 40], [56, 56], 40, 64, 11
B = B200Library(); B.B200_SetSeed(seed)
B.NewScheme(logn, logq, logp, logscale, h, "standard", "", "none")
B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
O = OScheme(logn, logq, logp, logscale, h, seed)
O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
i = 1
while i < (1 << O.log_max_slots):
    O.galois_key(O.ctx.galois_element(i)); i *= 2
rng = np.random.default_rng(8)
slots = 1 << (logn - 1)
xin = np.zeros(slots); xin[:num_slots] = rng.uniform(-1, 1, num_slots)
ct = B.Encrypt(B.B200_EncodeF64(xin, slots, 1, float(1 << logscale)))
oct_ = O.encrypt(O.encode(xin, 1, 2.0 ** logscale))
B.NewBootstrapper([61] * 4, num_slots)
ob = OBootstrapper(O, [61] * 4, num_slots)
out = B.Bootstrap(ct, num_slots)
dump = {}
ob.bootstrap(oct_, dump)
chain = np.fromfile(os.path.join(d, "chain.bin"), dtype=np.uint64)
print("chain equal:", [int(x) for x in chain] == ob.bs.Q + ob.bs.Pm, len(chain))
cheb = np.fromfile(os.path.join(d, "cheb.bin"))
print("cheb equal:", np.array_equal(cheb, np.array(ob.cheb)), "max diff", np.max(np.abs(cheb - np.array(ob.cheb))))
for name, ms in (("cts_m", ob.cts_m), ("stc_m", ob.stc_m)):
    for k, m in enumerate(ms):
        flat = np.fromfile(os.path.join(d, "%s%d.bin" % (name, k)))
        mine = []
        for key in m.keys():
            re, im = m.d[key]
            mine.append(np.concatenate([[float(key)], np.stack([re, im], 1).ravel()]))
        mine = np.concatenate(mine)
        same = mine.shape == flat.shape and np.array_equal(mine, flat)
        print(name, k, "equal:", same, "" if same else ("shapes %s %s maxdiff %g" % (mine.shape, flat.shape, np.max(np.abs(mine - flat)) if mine.shape == flat.shape else -1)))
for tag in sorted(dump):
    f = os.path.join(d, tag + ".bin")
    if not os.path.exists(f):
        print(tag, "no GPU dump"); continue
    g = np.fromfile(f, dtype=np.uint64)
    meta = np.fromfile(os.path.join(d, tag + ".meta.bin"))
    o = dump[tag]
    same = g.size == o.data.size and np.array_equal(g, o.data.ravel())
    print("%-12s words equal: %-5s level %d/%d  scale equal: %s (%.17g vs %.17g)" % (tag, same, int(meta[0]), o.level,
          np.longdouble(meta[2]) + np.longdouble(meta[3]) == o.scale, meta[2], float(o.scale)))
    if not same and g.size == o.data.size:
        gd = g.reshape(o.data.shape); bad = np.argwhere(gd != o.data)
        print("     first differing (poly, limb, coeff):", bad[0], "count", len(bad), "limbs touched", sorted(set(bad[:, 1]))[:12])
