"""Runs the REFERENCE's own Python layer (byte-compiled into oracle/_ref/orion_pyc by oracle/build_refpy.py) against a backend
object -- TEST INFRASTRUCTURE ONLY.  `run_flow` is the flow of the reference's tests/models/test_mlp.py:14-48 and
examples/run_*.py: init_scheme -> cleartext forward -> fit -> compile -> encode -> encrypt -> net.he() -> net(ct) ->
decrypt -> decode, with `Scheme.setup_backend` (orion/core/orion.py:90-106) returning the given backend: the one-branch hook
of INTEGRATION.md, applied by monkey-patching instead of editing the reference."""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PYC = os.path.join(ROOT, "oracle", "_ref", "orion_pyc.zip")

# configs/mlp.yml:3-19 of the reference with `backend` replaced (values restated here: the yml does not travel)
MLP_CONFIG = {
    "ckks_params": {"LogN": 13, "LogQ": [29, 26, 26, 26, 26, 26], "LogP": [29, 29], "LogScale": 26, "H": 8192, "RingType": "standard"},
    "orion": {"margin": 2, "embedding_method": "hybrid", "backend": "b200", "fuse_modules": True, "debug": False,
              "diags_path": "", "keys_path": "", "io_mode": "none"},
}
# configs/lola2.yml (logN=14, 60 + 6x30, LogP 60x2)
LOLA_CONFIG = {
    "ckks_params": {"LogN": 14, "LogQ": [60, 30, 30, 30, 30, 30, 30], "LogP": [60, 60], "LogScale": 30, "H": 8192, "RingType": "standard"},
    "orion": dict(MLP_CONFIG["orion"]),
}


def available():
    return os.path.exists(PYC)


def _helpers():
    spec = importlib.util.spec_from_file_location("make_orion_traces", os.path.join(HERE, "golden", "make_orion_traces.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def import_orion():
    """The byte-compiled reference package; /root/reference is NOT on sys.path (it does not exist on the GPU box)."""
    h = _helpers()
    h.stub_missing_modules()
    h.patch_scipy_for_torch_indices()
    if PYC not in sys.path:
        sys.path.insert(0, PYC)
    os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
    import orion
    import orion.core.orion as core
    assert os.path.abspath(orion.__file__).startswith(PYC), orion.__file__
    return orion, core


def run_flow(make_backend, model="mlp", config=None, seed=42):
    """Returns (cleartext logits, FHE logits, backend)."""
    import torch
    orion, core = import_orion()
    import orion.models as models
    box = {}

    def setup_backend(self, params):                       # `if backend == "b200": ...`
        b = make_backend()
        b.setup_bindings(params)
        box["b"] = b
        return b
    core.Scheme.setup_backend = setup_backend
    torch.manual_seed(seed)
    cfg = config or (MLP_CONFIG if model == "mlp" else LOLA_CONFIG)
    orion.init_scheme(cfg)
    net = models.MLP() if model == "mlp" else models.LoLA()
    x = (torch.rand(1, 1, 28, 28) - 0.1307) / 0.3081       # synthetic MNIST-normalised image
    net.eval()
    with torch.no_grad():
        clear = net(x).flatten().numpy().astype(np.float64)
    orion.fit(net, x)
    level = orion.compile(net)
    ct = orion.encrypt(orion.encode(x, level))
    net.he()
    out = net(ct)
    fhe = np.asarray(out.decrypt().decode(), dtype=np.float64).flatten()[: clear.size]
    return clear, fhe, box["b"]
