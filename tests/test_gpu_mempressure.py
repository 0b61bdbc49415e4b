"""Memory pressure (csrc/context.cu alloc_under_pressure): when the device cannot hold every Galois key and bootstrap factor set,
the library drops least-recently-used regenerable state and rebuilds it on next use.  ResNet-20 at logN=16 needs this on a 180 GB
B200; here the same path is forced on the logN=13 network with ORION_B200_MEM_LIMIT_MB, and the logits must stay within the
tolerance of the unconstrained run (regenerated keys carry fresh randomness, so the words differ but the values do not)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")

CODE = """
import os, sys, json, numpy as np
sys.path.insert(0, %r)
from orion_b200 import raw
from orion_b200.bindings import B200Library
from orion_b200.replay import Replayer, Trace
lib = raw.load()
B = B200Library(); B.B200_SetSeed(42)
t = Trace(%r, %r)
r = Replayer(B, t); r.setup()
outs = [np.asarray(r.infer(), dtype=np.float64)[:10].tolist() for _ in range(2)]
lib.B200_Sync()
print('OUT' + json.dumps({'outs': outs, 'peak': lib.B200_DeviceBytes(1), 'now': lib.B200_DeviceBytes(0), 'reclaimed': lib.B200_ReclaimedBytes()}))
"""


def run(limit_mb=None):
    env = dict(os.environ)
    if limit_mb:
        env["ORION_B200_MEM_LIMIT_MB"] = str(int(limit_mb))
        env["ORION_B200_DEBUG"] = "1"
    code = CODE % (ROOT, os.path.join(GOLD, "trace_resnet20.json"), os.path.join(GOLD, "trace_resnet20.npz"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, "\n".join(l for l in r.stderr.splitlines() if not l.startswith("[boot]"))[-4000:]
    return json.loads([l for l in r.stdout.splitlines() if l.startswith("OUT")][-1][3:])


def test_resnet20_survives_memory_pressure(b200lib):
    free = run()
    assert free["reclaimed"] == 0
    clear = np.array(json.load(open(os.path.join(GOLD, "trace_resnet20.json")))["meta"]["clear_output"])
    tight = run(limit_mb=0.9 * free["peak"] / 2**20)
    print("resnet20 logN=13: peak %.2f GB unconstrained; limit %.2f GB -> %.2f GB of keys / factor sets dropped and rebuilt over 2 inferences"
          % (free["peak"] / 2**30, 0.9 * free["peak"] / 2**30, tight["reclaimed"] / 2**30))
    # (10.6 GB of the 20 GB are the network's encoded diagonals, which are the caller's and never dropped; one linear transform pins
    # the ~60 keys it uses, 4.7 GB, and a bootstrap in progress holds its
    # encoded DFT factors, 0.8 GB: 16.1 GB cannot be given back, so the limit cannot go much lower than 0.9 x peak on this network)
    assert tight["reclaimed"] > 0, "the limit did not bite"
    assert tight["peak"] <= 0.9 * free["peak"] + 1
    for o in tight["outs"]:
        assert np.max(np.abs(np.array(o) - np.array(free["outs"][0]))) < 1e-3, (o, free["outs"][0])
        assert int(np.argmax(o)) == int(np.argmax(clear))
