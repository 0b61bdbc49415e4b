"""The live drop-in run: the UNMODIFIED reference Python layer (orion.init_scheme / fit / compile / encode / encrypt /
net.he() / net(ct) / decrypt / decode, through the real orion/backend/python/* wrappers) drives B200Library.

This is the reference's own end-to-end test (tests/models/test_mlp.py:14-48) with `backend: b200` hooked in as INTEGRATION.md
describes (one branch in Scheme.setup_backend, applied here by monkey-patching) and a synthetic MNIST-shaped image in place
of the dataset download.  Gate = the reference's: MAE(clear, FHE) < 0.005.  It is the test that shows the recorded traces
(tests/golden/, symbolic level / scale rules) are what the live layer really does: the live logits equal the replayed ones.

The reference package is used as bytecode built from /root/reference by oracle/build_refpy.py into oracle/_ref/orion_pyc.zip
(git-ignored build output that travels to the GPU box, like oracle/_ref/*.so); skipped when it is absent."""
import contextlib
import io
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import live_orion  # noqa: E402


def _flow(model):
    from orion_b200.bindings import B200Library

    def make():
        b = B200Library()
        b.B200_SetSeed(42)
        return b
    buf = io.StringIO()                                         # the reference prints every intermediate tensor
    with contextlib.redirect_stdout(buf):
        clear, fhe, b = live_orion.run_flow(make, model)
    return clear, fhe, b


@pytest.mark.skipif(not live_orion.available(), reason="oracle/_ref/orion_pyc.zip not built (python oracle/build_refpy.py needs /root/reference)")
@pytest.mark.parametrize("model", ["mlp", "lola"])
def test_reference_python_layer_runs_on_the_b200_backend(b200lib, model):
    clear, fhe, b = _flow(model)
    try:
        mae = float(np.mean(np.abs(clear - fhe)))
        print("live %s: MAE %.2e, max |err| %.2e" % (model, mae, float(np.max(np.abs(clear - fhe)))))
        assert mae < 0.005, (mae, clear, fhe)                    # tests/models/test_mlp.py:45-48
        assert int(np.argmax(fhe)) == int(np.argmax(clear))
        # the recorded trace of the same flow (same seed, same image) decodes to the same logits when replayed
        from orion_b200.bindings import B200Library
        from orion_b200.replay import replay_model
        gold = os.path.join(HERE, "golden")
        b.DeleteScheme()
        B2 = B200Library(); B2.B200_SetSeed(42)
        res = replay_model(B2, os.path.join(gold, "trace_%s.json" % model), os.path.join(gold, "trace_%s.npz" % model))
        assert np.max(np.abs(np.array(res["meta"]["clear_output"]) - clear)) < 1e-6          # same network, same input
        rep = np.asarray(res["output"], dtype=np.float64)[: clear.size]
        assert np.max(np.abs(rep - fhe)) < 2e-3, (rep, fhe)       # same calls; only the encryption randomness differs
        B2.DeleteScheme()
    finally:
        b.DeleteScheme()
