"""Multi-GPU data path on real hardware (needs >= 2 GPUs; skipped on a 1-GPU box): SURVEY.md 8(e).

Two ranks over NCCL (tests/multi_worker.py under torch.distributed.run):
  * the gather of a layer's output ciphertexts moves DEVICE limbs (B200_CiphertextAcquire + torch.distributed broadcast on
    the library's stream): every gathered ciphertext equals, word for word, the one a single rank computes;
  * one ResNet-20 inference sharded across the ranks (block rows of the linear layers, per-ciphertext activations and
    bootstraps where the operand lives) decodes to exactly the logits of the single-rank replay."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    from orion_b200 import raw
    return raw.load().B200_DeviceCount()


def test_two_rank_nccl_gather_and_sharded_inference(b200lib, tmp_path):
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    out = tmp_path / "multi.json"
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "multi_worker.py"), str(out)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    assert len(res) == 2
    for rr in res:
        assert rr["gather_equal"] is True, rr
        assert rr["sharded_equal"] is True, rr
        assert rr["gather_bytes"] > 0 and rr["sharded_transfers"] > 0
    # both ranks did real work: the sharding is not "everything on rank 0"
    assert min(rr["sharded_ops_run"] for rr in res) > 0.25 * max(rr["sharded_ops_run"] for rr in res), res
