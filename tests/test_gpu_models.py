"""End-to-end encrypted inference through the drop-in boundary, driven by call traces recorded from the reference
Orion (tests/golden/make_orion_traces.py).  Gate = the reference's own: MAE(clear, FHE) < 0.005 (tests/models/test_mlp.py:45-48)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def run_trace(name):
    from orion_b200.bindings import B200Library
    from orion_b200.replay import replay_model
    B = B200Library()
    B.B200_SetSeed(42)
    res = replay_model(B, os.path.join(GOLD, "trace_%s.json" % name), os.path.join(GOLD, "trace_%s.npz" % name))
    B.DeleteScheme()
    return res


def test_mlp_encrypted_inference_matches_cleartext(b200lib):
    res = run_trace("mlp")
    clear = np.array(res["meta"]["clear_output"])
    fhe = res["output"][: len(clear)]
    mae = float(np.mean(np.abs(clear - fhe)))
    assert mae < 0.005, (mae, clear, fhe)
    assert int(np.argmax(fhe)) == int(np.argmax(clear))
    assert np.max(np.abs(clear - fhe)) < 1e-2
    assert res["counts"]["EvaluateLinearTransform"] == 3 and res["counts"]["MulRelinCiphertextNew"] == 2
