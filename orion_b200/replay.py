"""Replays a recorded Orion backend-call trace (tests/golden/trace_*.json + .npz) through B200Library.

A trace is the exact sequence of `backend.X(...)` calls the reference Orion made for a model (recorded by
tests/golden/make_orion_traces.py with a symbolic backend).  Replaying it drives this backend the way Orion would:
same calls, same order, same handles -- without needing the reference's Python on the GPU box.
Payloads are either stored in full (MLP / LoLA) or as sizes only (ResNet-20: seeded synthetic values of that shape).
"""
import json
import time

import numpy as np

DEBUG_OPS = ("Decrypt", "Decode")


class Trace:
    def __init__(self, json_path, npz_path=None):
        d = json.load(open(json_path))
        self.meta, self.ops = d["meta"], d["ops"]
        self.payload = dict(np.load(npz_path)) if npz_path else {}
        self.n_compile = self.meta["n_compile_ops"]

    def values(self, key, rng):
        a = self.payload[key]
        if a.dtype != np.int64:
            return a
        n = int(a[0])
        return (rng.standard_normal(n) * 0.05).astype(np.float32)       # synthetic payload of the recorded size


class Replayer:
    """Maps recorded handles to live handles and issues the calls."""

    def __init__(self, backend, trace, seed=1234, skip_debug=True):
        self.B, self.t, self.rng, self.skip_debug = backend, trace, np.random.default_rng(seed), skip_debug
        self.ct, self.pt, self.lt, self.poly = {}, {}, {}, {}
        self.last_decrypt = max((i for i, o in enumerate(trace.ops) if o[0] == "Decrypt"), default=-1)
        self.output = None
        self.counts = {}

    def _run(self, i, op):
        B, (name, args, ret) = self.B, op
        ct, pt, lt, poly = self.ct, self.pt, self.lt, self.poly
        self.counts[name] = self.counts.get(name, 0) + 1
        if name == "NewScheme":
            B.NewScheme(*args)
        elif name == "Encode":
            pt[ret] = B.Encode(self.t.values(args[0], self.rng), args[1], args[2])
        elif name == "Encrypt":
            ct[ret] = B.Encrypt(pt[args[0]])
        elif name == "Decrypt":
            if self.skip_debug and i != self.last_decrypt:
                return
            pt[ret] = B.Decrypt(ct[args[0]])
        elif name == "Decode":
            if args[0] in pt:
                vals = B.Decode(pt[args[0]])
                if i > self.last_decrypt:
                    self.output = np.array(vals)
        elif name == "DeletePlaintext":
            if args[0] in pt:
                B.DeletePlaintext(pt.pop(args[0]))
        elif name == "DeleteCiphertext":
            if args[0] in ct:
                B.DeleteCiphertext(ct.pop(args[0]))
        elif name == "DeleteLinearTransform":
            if args[0] in lt:
                B.DeleteLinearTransform(lt.pop(args[0]))
        elif name == "SetCiphertextScale":
            B.SetCiphertextScale(ct[args[0]], args[1])
        elif name == "SetPlaintextScale":
            B.SetPlaintextScale(pt[args[0]], args[1])
        elif name == "GenerateLinearTransform":
            lt[ret] = B.GenerateLinearTransform(args[0], self.t.values(args[1], self.rng), args[2], args[3], args[4])
        elif name == "EvaluateLinearTransform":
            ct[ret] = B.EvaluateLinearTransform(lt[args[0]], ct[args[1]])
        elif name in ("GenerateMonomial", "GenerateChebyshev"):
            poly[ret] = getattr(B, name)(self.t.payload[args[0]])
        elif name == "EvaluatePolynomial":
            ct[ret] = B.EvaluatePolynomial(ct[args[0]], poly[args[1]], args[2])
        elif name == "Bootstrap":
            ct[ret] = B.Bootstrap(ct[args[0]], args[1])
        elif name == "ModDropCiphertext":
            ct[ret] = B.ModDropCiphertext(ct[args[0]])
        elif name in ("AddPlaintext", "AddPlaintextNew", "SubPlaintext", "SubPlaintextNew", "MulPlaintext", "MulPlaintextNew"):
            ct[ret] = getattr(B, name)(ct[args[0]], pt[args[1]])
        elif name in ("AddCiphertext", "AddCiphertextNew", "SubCiphertext", "SubCiphertextNew",
                      "MulRelinCiphertext", "MulRelinCiphertextNew"):
            ct[ret] = getattr(B, name)(ct[args[0]], ct[args[1]])
        elif name in ("Negate", "Rescale", "RescaleNew", "CloneCiphertext"):
            ct[ret] = getattr(B, name)(ct[args[0]])
        elif name in ("Rotate", "RotateNew", "AddScalar", "AddScalarNew", "SubScalar", "SubScalarNew", "MulScalarInt",
                      "MulScalarIntNew", "MulScalarFloat", "MulScalarFloatNew"):
            ct[ret] = getattr(B, name)(ct[args[0]], args[1])
        else:                                   # New*, Generate*Key*, AddRotationKey, NewBootstrapper, ...
            getattr(B, name)(*args)

    def run(self, start, stop):
        for i in range(start, stop):
            self._run(i, self.t.ops[i])

    def setup(self):
        """init_scheme + compile: everything up to (not including) the encode of the input."""
        self.run(0, self.t.n_compile)
        self._snap = (dict(self.ct), dict(self.pt))

    def reset_inference(self):
        """Release the handles a previous inference created; back to the state right after setup()."""
        ct0, pt0 = self._snap
        for k, v in list(self.ct.items()):
            if k not in ct0 or ct0[k] != v:
                self.B.DeleteCiphertext(v)
        for k, v in list(self.pt.items()):
            if k not in pt0 or pt0[k] != v:
                self.B.DeletePlaintext(v)
        self.ct, self.pt = dict(ct0), dict(pt0)

    def infer(self):
        """One encrypted inference: encode + encrypt the input, net(ct), decrypt + decode.  Returns the decoded slots.
        Repeatable: handles created by a previous inference are released first."""
        self.reset_inference()
        self.run(self.t.n_compile, len(self.t.ops))
        return self.output


class BatchReplayer:
    """K independent inferences (K images) over ONE scheme -- keys, encoded diagonals, polynomials and bootstrappers are
    shared -- issued call by call in lockstep: call i of image 0, call i of image 1, ...  The images' ciphertexts are
    independent, so the library's dataflow lanes (csrc/capi.cu) run them on different streams and the small-ring launches of
    one image fill the SMs the others leave idle.  This is how a server would use the boundary for a batch: nothing in the
    per-call semantics changes, every image's result is bit-identical to its single replay."""

    def __init__(self, backend, trace, k, seed=1234):
        self.rs = [Replayer(backend, trace, seed=seed) for _ in range(k)]
        self.t = trace

    def setup(self):
        r0 = self.rs[0]
        r0.setup()
        for r in self.rs[1:]:
            r.ct, r.pt, r.lt, r.poly = dict(r0.ct), dict(r0.pt), r0.lt, r0.poly
            r._snap = (dict(r0._snap[0]), dict(r0._snap[1]))

    def infer(self):
        for r in self.rs:
            r.reset_inference()
        ops = self.t.ops
        for i in range(self.t.n_compile, len(ops)):
            op = ops[i]
            for r in self.rs:
                r._run(i, op)
        return [r.output for r in self.rs]

    @property
    def counts(self):
        return self.rs[0].counts


def replay_model(backend, json_path, npz_path, sync=None):
    t = Trace(json_path, npz_path)
    r = Replayer(backend, t)
    t0 = time.perf_counter(); r.setup()
    if sync: sync()
    t1 = time.perf_counter(); out = r.infer()
    if sync: sync()
    t2 = time.perf_counter()
    return dict(output=out, setup_s=t1 - t0, infer_s=t2 - t1, meta=t.meta, counts=r.counts)
