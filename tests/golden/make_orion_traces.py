"""Records the sequence of backend calls the REFERENCE Orion makes for a model, as a replayable trace.

Runs in the build container only (imports the unmodified reference from /root/reference, which does not exist on the
GPU box).  The reference's `Scheme.setup_backend` is pointed at a *symbolic* backend that tracks ids, levels, scales
and slot counts but does no cryptography, so `init_scheme -> fit -> compile -> encode -> encrypt -> net.he() ->
net(ct) -> decrypt -> decode` runs on the CPU exactly as the reference drives its Lattigo backend.  Every call is
logged; float payloads (encoded vectors, diagonals, polynomial coefficients) go to an .npz next to the .json op list.

  python tests/golden/make_orion_traces.py mlp        -> tests/golden/trace_mlp.{json,npz}      (full payloads)
  python tests/golden/make_orion_traces.py resnet20   -> tests/golden/trace_resnet20.json       (structure only)
  python tests/golden/make_orion_traces.py lola_ci    -> configs/lola.yml as shipped (ConjugateInvariant ring)

`orion_b200/replay.py` replays a trace through B200Library on the GPU (tests/test_gpu_models.py, bench.py).
"""
import json
import math
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def stub_missing_modules():
    for name in ("h5py", "matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                m.File = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("h5py stub"))
                sys.modules[name] = m
    if "matplotlib" in sys.modules and not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules.get("matplotlib.pyplot")


def patch_scipy_for_torch_indices():
    """The reference indexes scipy sparse matrices with torch tensors (orion/core/packing.py:94,100); the scipy in this
    image only accepts numpy indices.  Convert at the scipy boundary instead of touching the reference."""
    import scipy.sparse as sp
    import torch

    def conv(k):
        if isinstance(k, tuple):
            return tuple(conv(x) for x in k)
        return k.numpy() if isinstance(k, torch.Tensor) else k

    for cls in (sp.lil_matrix, sp.csc_matrix, sp.csr_matrix):
        g, st = cls.__getitem__, cls.__setitem__
        cls.__getitem__ = (lambda g: lambda self, k: g(self, conv(k)))(g)
        cls.__setitem__ = (lambda st: lambda self, k, v: st(self, conv(k), v.numpy() if isinstance(v, torch.Tensor) else v))(st)


class SymbolicBackend:
    """Tracks metadata only.  Scale / level rules are those of include/orion_b200_scheme.h (= the Go shim's)."""

    def __init__(self, keep_payloads=True):
        self.ops, self.payload, self.keep = [], {}, keep_payloads
        self.pt, self.ct, self.lt, self.poly = {}, {}, {}, {}
        self.next = {"pt": 0, "ct": 0, "lt": 0, "poly": 0}
        self.free = {"pt": [], "ct": [], "lt": [], "poly": []}
        self.arithmeticoperator_handle = None
        self.recording = True

    # ---- plumbing ----
    def _new(self, kind, meta):
        table = getattr(self, kind)
        if self.free[kind]:
            i = min(self.free[kind]); self.free[kind].remove(i)
        else:
            i = self.next[kind]; self.next[kind] += 1
        table[i] = meta
        return i

    def _del(self, kind, i):
        if i in getattr(self, kind):
            del getattr(self, kind)[i]; self.free[kind].append(i)

    def _log(self, name, *args, ret=None):
        if self.recording:
            self.ops.append([name, list(args), ret])
        return ret

    def _store(self, arr, dtype=np.float32):
        key = "p%d" % len(self.payload)
        a = np.asarray(arr, dtype=dtype)
        self.payload[key] = a if (self.keep or a.size <= 512) else np.array([a.size], dtype=np.int64)   # polynomials are always kept
        return key

    # ---- scheme ----
    def setup_bindings(self, params):
        from orion_b200 import raw
        self.logn = params.get_logn()
        self.logq, self.logp = params.get_logq(), params.get_logp()
        self.logscale = params.get_logscale()
        ci = params.get_ringtype() != "standard"            # conjugate-invariant ring: N slots, primes = 1 mod 4N
        self.Q, self.P = raw.gen_moduli(self.logq, self.logp, self.logn + (2 if ci else 1))
        self.max_slots = 1 << (self.logn if ci else self.logn - 1)
        self.max_level = len(self.Q) - 1
        self._log("NewScheme", self.logn, self.logq, self.logp, self.logscale, params.get_hamming_weight(),
                  params.get_ringtype(), "", "none")

    def DeleteScheme(self):
        pass

    def __getattr__(self, name):
        if name.startswith("New") or name in ("GenerateSecretKey", "GeneratePublicKey", "GenerateRelinearizationKey",
                                              "GenerateEvaluationKeys", "AddRotationKey", "RemoveRotationKeys",
                                              "DeleteBootstrappers"):
            def f(*args):
                args = [list(a) if isinstance(a, (list, tuple)) else a for a in args]
                self._log(name, *args)
            return f
        raise AttributeError("SymbolicBackend: reference called an unmodelled backend function: " + name)

    # ---- tensors ----
    def DeletePlaintext(self, i): self._del("pt", i); self._log("DeletePlaintext", i)
    def DeleteCiphertext(self, i): self._del("ct", i); self._log("DeleteCiphertext", i)
    def GetPlaintextScale(self, i): return int(self.pt[i]["scale"])
    def GetCiphertextScale(self, i): return int(self.ct[i]["scale"])
    def SetPlaintextScale(self, i, s): self.pt[i]["scale"] = float(s); self._log("SetPlaintextScale", i, int(s))
    def SetCiphertextScale(self, i, s): self.ct[i]["scale"] = float(s); self._log("SetCiphertextScale", i, int(s))
    def GetPlaintextLevel(self, i): return self.pt[i]["level"]
    def GetCiphertextLevel(self, i): return self.ct[i]["level"]
    def GetPlaintextSlots(self, i): return self.max_slots
    def GetCiphertextSlots(self, i): return self.ct[i].get("slots", self.max_slots)
    def GetCiphertextDegree(self, i): return 1
    def GetModuliChain(self): return list(self.Q)
    def GetLivePlaintexts(self): return list(self.pt)
    def GetLiveCiphertexts(self): return list(self.ct)

    # ---- encoder / encryptor ----
    def Encode(self, values, level, scale):
        i = self._new("pt", dict(level=int(level), scale=float(scale)))
        return self._log("Encode", self._store(values), int(level), int(scale), ret=i)

    def Decode(self, pt):
        self._log("Decode", pt)
        return [0.0] * self.max_slots

    def Encrypt(self, pt):
        i = self._new("ct", dict(self.pt[pt]))
        return self._log("Encrypt", pt, ret=i)

    def Decrypt(self, ct):
        i = self._new("pt", dict(level=self.ct[ct]["level"], scale=self.ct[ct]["scale"]))
        return self._log("Decrypt", ct, ret=i)

    # ---- evaluator ----
    def _unary(self, name, ct, new, level=None, scale=None, extra=()):
        m = dict(self.ct[ct])
        if level is not None: m["level"] = level
        if scale is not None: m["scale"] = scale
        if new:
            i = self._new("ct", m)
        else:
            self.ct[ct] = m; i = ct
        return self._log(name, ct, *extra, ret=i)

    def Negate(self, ct): return self._unary("Negate", ct, True)
    def Rotate(self, ct, k): return self._unary("Rotate", ct, False, extra=(int(k),))
    def RotateNew(self, ct, k): return self._unary("RotateNew", ct, True, extra=(int(k),))
    def CloneCiphertext(self, ct): return self._unary("CloneCiphertext", ct, True)

    def _rescaled(self, ct):
        m = self.ct[ct]
        assert m["level"] >= 1, "symbolic backend: rescale at level 0"
        return m["level"] - 1, m["scale"] / self.Q[m["level"]]

    def Rescale(self, ct):
        l, s = self._rescaled(ct); return self._unary("Rescale", ct, False, l, s)

    def RescaleNew(self, ct):
        l, s = self._rescaled(ct)
        self.ct[ct].update(level=l, scale=s)
        return self._unary("RescaleNew", ct, True)

    def _ModDropCiphertext(self, _h, ct, _o=None):
        return self._unary("ModDropCiphertext", ct, False, self.ct[ct]["level"] - 1)

    def AddScalar(self, ct, c): return self._unary("AddScalar", ct, False, extra=(float(c),))
    def AddScalarNew(self, ct, c): return self._unary("AddScalarNew", ct, True, extra=(float(c),))
    def SubScalar(self, ct, c): return self._unary("SubScalar", ct, False, extra=(float(c),))
    def SubScalarNew(self, ct, c): return self._unary("SubScalarNew", ct, True, extra=(float(c),))
    def MulScalarInt(self, ct, k): return self._unary("MulScalarInt", ct, False, extra=(int(k),))
    def MulScalarIntNew(self, ct, k): return self._unary("MulScalarIntNew", ct, True, extra=(int(k),))

    def _mulf(self, name, ct, c, new):
        m = self.ct[ct]
        s = m["scale"] if float(c) == math.floor(float(c)) else m["scale"] * self.Q[m["level"]]
        return self._unary(name, ct, new, scale=s, extra=(float(c),))

    def MulScalarFloat(self, ct, c): return self._mulf("MulScalarFloat", ct, c, False)
    def MulScalarFloatNew(self, ct, c): return self._mulf("MulScalarFloatNew", ct, c, True)

    def _binary(self, name, a, b, new, b_is_pt, mul):
        ma, mb = self.ct[a], (self.pt[b] if b_is_pt else self.ct[b])
        level = min(ma["level"], mb["level"])
        scale = ma["scale"] * mb["scale"] if mul else max(ma["scale"], mb["scale"])
        return self._unary(name, a, new, level, scale, extra=(b,))

    def AddPlaintext(self, a, b): return self._binary("AddPlaintext", a, b, False, True, False)
    def AddPlaintextNew(self, a, b): return self._binary("AddPlaintextNew", a, b, True, True, False)
    def SubPlaintext(self, a, b): return self._binary("SubPlaintext", a, b, False, True, False)
    def SubPlaintextNew(self, a, b): return self._binary("SubPlaintextNew", a, b, True, True, False)
    def MulPlaintext(self, a, b): return self._binary("MulPlaintext", a, b, False, True, True)
    def MulPlaintextNew(self, a, b): return self._binary("MulPlaintextNew", a, b, True, True, True)
    def AddCiphertext(self, a, b): return self._binary("AddCiphertext", a, b, False, False, False)
    def AddCiphertextNew(self, a, b): return self._binary("AddCiphertextNew", a, b, True, False, False)
    def SubCiphertext(self, a, b): return self._binary("SubCiphertext", a, b, False, False, False)
    def SubCiphertextNew(self, a, b): return self._binary("SubCiphertextNew", a, b, True, False, False)
    def MulRelinCiphertext(self, a, b): return self._binary("MulRelinCiphertext", a, b, False, False, True)
    def MulRelinCiphertextNew(self, a, b): return self._binary("MulRelinCiphertextNew", a, b, True, False, True)

    # ---- linear transforms ----
    def GenerateLinearTransform(self, idxs, data, level, bsgs_ratio, io_mode):
        i = self._new("lt", dict(level=int(level), idxs=[int(x) for x in idxs]))
        return self._log("GenerateLinearTransform", [int(x) for x in idxs], self._store(data), int(level),
                         float(bsgs_ratio), "none", ret=i)

    def GetLinearTransformRotationKeys(self, lt):
        from oracle.scheme import OScheme
        idxs = sorted({k & (self.max_slots - 1) for k in self.lt[lt]["idxs"]})
        n1 = OScheme.find_best_bsgs_ratio(idxs, self.max_slots, 0)
        _, giants, babies = OScheme.bsgs_index(idxs, self.max_slots, n1)
        nth_root = 2 * self.max_slots * 2                   # 2N (standard ring) / 4N (conjugate-invariant ring: N slots)
        return sorted({pow(5, r, nth_root) for r in giants + babies if r})

    def GenerateConsolidatedRotationKeys(self, keys):
        self._log("GenerateConsolidatedRotationKeys", [int(k) for k in keys])

    def GenerateLinearTransformRotationKey(self, g):
        self._log("GenerateLinearTransformRotationKey", int(g))

    def EvaluateLinearTransform(self, lt, ct):
        m = self.ct[ct]
        level = min(m["level"], self.lt[lt]["level"])
        i = self._new("ct", dict(m, level=level, scale=m["scale"] * self.Q[self.lt[lt]["level"]]))
        return self._log("EvaluateLinearTransform", lt, ct, ret=i)

    def DeleteLinearTransform(self, lt): self._del("lt", lt); self._log("DeleteLinearTransform", lt)

    # ---- polynomials ----
    def GenerateMonomial(self, coeffs, *_):
        i = self._new("poly", dict(deg=len(coeffs) - 1))
        return self._log("GenerateMonomial", self._store(coeffs), ret=i)

    def GenerateChebyshev(self, coeffs, *_):
        i = self._new("poly", dict(deg=len(coeffs) - 1))
        return self._log("GenerateChebyshev", self._store(coeffs), ret=i)

    def GetPolyDepth(self, p): return max(1, int(self.poly[p]["deg"]).bit_length())

    def EvaluatePolynomial(self, ct, poly, out_scale):
        m = self.ct[ct]
        depth = self.GetPolyDepth(poly)
        assert m["level"] - depth >= 0, "symbolic backend: polynomial deeper than the level"
        i = self._new("ct", dict(m, level=m["level"] - depth, scale=float(out_scale)))
        return self._log("EvaluatePolynomial", ct, poly, int(out_scale), ret=i)

    def GenerateMinimaxSignCoeffs(self, degrees, prec, logalpha, logerr, debug):
        from orion_b200.minimax import generate_minimax_sign_coeffs
        return generate_minimax_sign_coeffs(list(degrees), prec, logalpha, logerr)

    # ---- bootstrapping ----
    def NewBootstrapper(self, logps, slots):
        self._log("NewBootstrapper", [int(x) for x in logps], int(slots))

    def Bootstrap(self, ct, slots):
        m = self.ct[ct]
        i = self._new("ct", dict(m, level=self.max_level, scale=float(1 << self.logscale)))
        return self._log("Bootstrap", ct, int(slots), ret=i)


def build_model(name):
    import torch
    import orion.models as models
    torch.manual_seed(42)
    if name == "mlp":
        net = models.MLP()
        x = (torch.rand(1, 1, 28, 28) - 0.1307) / 0.3081
        cfg = os.path.join(REF, "configs", "mlp.yml")
    elif name == "lola":
        net = models.LoLA()
        x = (torch.rand(1, 1, 28, 28) - 0.1307) / 0.3081
        cfg = os.path.join(REF, "configs", "lola2.yml")
    elif name == "mlp_ci":                                   # the reference's own test config: tests/configs/mlp.yml
        net = models.MLP()
        x = (torch.rand(1, 1, 28, 28) - 0.1307) / 0.3081
        cfg = os.path.join(REF, "tests", "configs", "mlp.yml")
    elif name == "lola_ci":                                  # configs/lola.yml as shipped: ConjugateInvariant ring, logN=13
        net = models.LoLA()
        x = (torch.rand(1, 1, 28, 28) - 0.1307) / 0.3081
        cfg = os.path.join(REF, "configs", "lola.yml")
    elif name == "resnet20":
        net = models.ResNet20()
        mean = torch.tensor([0.4914, 0.4822, 0.4465]).view(1, 3, 1, 1)
        std = torch.tensor([0.2023, 0.1994, 0.2010]).view(1, 3, 1, 1)
        x = (torch.rand(1, 3, 32, 32) - mean) / std
        cfg = os.path.join(REF, "configs", "resnet.yml")
    else:
        raise SystemExit("unknown model " + name)
    return net, x, cfg


def main():
    import yaml
    name = sys.argv[1] if len(sys.argv) > 1 else "mlp"
    keep = "--structure-only" not in sys.argv
    stub_missing_modules()
    patch_scipy_for_torch_indices()
    sys.path.insert(0, REF)
    import torch
    import orion
    import orion.core.orion as core
    net, x, cfg_path = build_model(name)
    cfg = yaml.safe_load(open(cfg_path))
    cfg["orion"]["debug"] = False
    cfg["orion"]["io_mode"] = "none"                       # run with TORCHDYNAMO_DISABLE=1 where inductor cannot compile
    if name not in ("lola_ci", "mlp_ci"):
        cfg["ckks_params"]["RingType"] = "standard"
    for a in sys.argv[2:]:
        if a.startswith("--logn="):
            cfg["ckks_params"]["LogN"] = int(a.split("=")[1])
    sb = SymbolicBackend(keep_payloads=keep)

    def setup_backend(self, params):
        sb.setup_bindings(params)
        return sb
    core.Scheme.setup_backend = setup_backend
    orion.init_scheme(cfg)
    net.eval()
    with torch.no_grad():
        clear = net(x).flatten().numpy().astype(np.float64)
    orion.fit(net, x)
    input_level = orion.compile(net)
    n_compile_ops = len(sb.ops)
    vec_ptxt = orion.encode(x, input_level)
    vec_ctxt = orion.encrypt(vec_ptxt)
    net.he()
    out = net(vec_ctxt)
    out_ids = list(out.ids)
    out.decrypt().decode()
    meta = dict(model=name, config=cfg, input_level=int(input_level), n_compile_ops=n_compile_ops,
                output_ct_ids=[int(i) for i in out_ids], payloads="full" if keep else "sizes",
                Q=[int(q) for q in sb.Q], P=[int(p) for p in sb.P],
                clear_output=clear.tolist(), n_ops=len(sb.ops))
    suffix = "" if cfg["ckks_params"]["LogN"] == yaml.safe_load(open(cfg_path))["ckks_params"]["LogN"] else "_logn%d" % cfg["ckks_params"]["LogN"]
    base = os.path.join(ROOT, "tests", "golden", "trace_" + name + suffix)
    def plain(o):
        if isinstance(o, (list, tuple)):
            return [plain(v) for v in o]
        if isinstance(o, np.integer):
            return int(o)
        if isinstance(o, np.floating):
            return float(o)
        return o
    json.dump(dict(meta=meta, ops=plain(sb.ops)), open(base + ".json", "w"))
    np.savez_compressed(base + ".npz", **sb.payload)
    counts = {}
    for o in sb.ops:
        counts[o[0]] = counts.get(o[0], 0) + 1
    print(json.dumps(dict(model=name, ops=len(sb.ops), counts=counts, input_level=int(input_level))))
    print("sizes:", os.path.getsize(base + ".json"), os.path.getsize(base + ".npz"))


if __name__ == "__main__":
    main()
