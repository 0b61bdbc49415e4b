"""CPU tests of the drop-in boundary: the shared library loads without a GPU and exports every symbol the headers
declare; host-only entry points agree with the oracle; the product never imports the oracle."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in ("orion_b200.h", "orion_b200_scheme.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for m in re.finditer(r"^[A-Za-z_][\w \*]*?[\s\*]+(\w+)\s*\([^;{]*\)\s*;", src, flags=re.M):
            names.append(m.group(1))
    return sorted(set(names))


def test_library_exports_every_declared_symbol(b200lib):
    lib = ctypes.CDLL(b200lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) > 110
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and the two Python signature tables cover exactly the declared surface
    from orion_b200.bindings import SCHEME_SIGNATURES
    bound = set(SCHEME_SIGNATURES) | set(b200lib.RAW_SIGNATURES)
    assert set(names) <= bound, sorted(set(names) - bound)


def test_host_only_entry_points_match_oracle(b200lib, orc):
    for logN, lq, lp in ((13, [29, 26, 26, 26, 26, 26], [29, 29]), (14, [60, 30, 30, 30], [60, 60]),
                         (16, [60] + [45] * 23, [61, 61, 61])):
        assert b200lib.gen_moduli(lq, lp, logN + 1) == orc.gen_moduli(lq, lp, logN + 1)
    lib = b200lib.load()
    Q, _ = orc.gen_moduli([45, 30], [], 14)
    for q in Q:
        assert lib.B200_FindPsi(q, 1 << 14) == orc.lib().orc_psi(q, 1 << 14)
    assert lib.B200_DeviceCount() >= 0


def test_compute_without_gpu_fails_loudly(b200lib):
    if b200lib.load().B200_DeviceCount() > 0:
        return
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from orion_b200 import raw\n"
            "raw.RawContext(12, [1099511480321], [])\n" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CUDA device" in r.stderr and "no CPU path" in r.stderr


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "orion_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "ckks_oracle" not in src.replace("oracle/ckks_oracle.c orc_rng_u64", ""), f
    r = subprocess.run(["ldd", os.path.join(pkg, "liborion_b200.so")], capture_output=True, text=True)
    assert "oracle" not in r.stdout


def test_binding_marshalling_shapes():
    from orion_b200.bindings import SCHEME_SIGNATURES, _flat_argtypes
    # NewScheme(logN, logQ*, lenQ, logP*, lenP, logScale, h, ringType, keysPath, ioMode): lattigo/scheme.go:35-44
    assert len(_flat_argtypes(SCHEME_SIGNATURES["NewScheme"][0])) == 10
    assert len(_flat_argtypes(SCHEME_SIGNATURES["GenerateLinearTransform"][0])) == 7
    assert len(_flat_argtypes(SCHEME_SIGNATURES["Encode"][0])) == 4


def test_reference_gpu_shim_loads_and_exports(orc):
    """oracle/_ref/libgpuntt_refgpu.so (the reference's GPU merge NTT compiled in place for sm_100a): loads without a GPU
    and exports the four entry points tests/test_gpu_refgpu.py drives.  Skipped when the reference tree was not there to
    build it from."""
    import pytest
    R = orc.ref_gpu()
    if R is None:
        pytest.skip("oracle/_ref/libgpuntt_refgpu.so not built")
    for name in ("refgpu_create", "refgpu_destroy", "refgpu_ntt_host", "refgpu_time_ms"):
        assert hasattr(R, name)


def test_bench_reference_arm_prints_exactly_one_json_line():
    """`bench.py --impl reference` (the driver's reference arm): the CPU oracle on the host cores, one JSON line on stdout,
    the keys of the contract, nothing else on stdout."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "keyswitches_per_s_logN16" and d["unit"] == "keyswitch/s"
    assert d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_late_deletes_without_a_scheme_are_no_ops(b200lib):
    """The reference's Python objects free their handles from __del__ in any order, also after DeleteScheme and at interpreter
    exit (orion/backend/python/tensors.py:16-22,112-118); the Go shim's Delete* then delete from an empty map
    (lattigo/tensors.go:35-42).  Same here: no abort, no GPU needed."""
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.DeleteScheme()
    B.DeleteCiphertext(3); B.DeletePlaintext(0); B.DeleteLinearTransform(7)
    B.DeleteScheme()
