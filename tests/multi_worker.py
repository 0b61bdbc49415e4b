"""Worker of tests/test_gpu_multi.py: one process per GPU (torchrun), NCCL for ciphertext limbs, gloo for metadata.

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tests/multi_worker.py <out.json>
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    os.environ["ORION_B200_DEVICE"] = str(local)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    meta = dist.new_group(backend="gloo")
    from orion_b200 import raw
    from orion_b200.bindings import B200Library
    from orion_b200.parallel import DeviceExchange, ShardedReplayer, shard
    from orion_b200.replay import Replayer, Trace
    lib = raw.load()
    res = {"rank": rank, "world": world}

    # ---- 1. gather of a layer's output ciphertexts on device limbs: gathered words == the words computed locally ----
    B = B200Library()
    B.B200_SetSeed(99)                                           # same keys and same sampling events on every rank
    logN, logq = 13, [55, 40, 40, 40]
    B.NewScheme(logN, logq, [56, 56], 40, 64, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
    N, n_items = 1 << logN, 5
    rng = np.random.default_rng(3)
    xs = [rng.uniform(-1, 1, N // 2).astype(np.float32) for _ in range(n_items)]
    base = [B.Encrypt(B.Encode(x.tolist(), 3, 1 << 40)) for x in xs]            # replicated inputs (deterministic encryption)

    def layer(i):                                                # some per-item work: square, rescale, rotate
        c = B.Rescale(B.MulRelinCiphertextNew(base[i], base[i]))
        return B.RotateNew(c, 1 << (i % 3))

    def words(ct):
        a = np.empty((B.GetCiphertextDegree(ct) + 1, B.GetCiphertextLevel(ct) + 1, N), dtype=np.uint64)
        B.B200_ExportCiphertext(ct, a)
        return a
    everything = {i: layer(i) for i in range(n_items)}          # what a single rank would compute
    mine = {i: everything[i] for i in shard(n_items, rank, world)}
    x = DeviceExchange(B, N, rank, world, meta_group=meta)
    got = x.all_gather(dict(mine), n_items)
    ok = True
    for i in range(n_items):
        same = np.array_equal(words(got[i]), words(everything[i]))
        same = same and B.GetCiphertextScale(got[i]) == B.GetCiphertextScale(everything[i]) and B.GetCiphertextLevel(got[i]) == B.GetCiphertextLevel(everything[i])
        ok = ok and same
    # a gathered remote ciphertext is a first-class handle: use it
    j = (rank + 1) % n_items
    s1 = words(B.AddCiphertextNew(got[j], got[j])); s2 = words(B.AddCiphertextNew(everything[j], everything[j]))
    res["gather_equal"] = bool(ok and np.array_equal(s1, s2))
    res["gather_bytes"] = int(x.bytes_moved)
    B.DeleteScheme()

    # ---- 2. one ResNet-20 inference sharded over the ranks: logits bit-identical to the single-rank replay ----
    name = os.environ.get("ORION_B200_MULTI_TRACE", "resnet20")
    gold = os.path.join(ROOT, "tests", "golden")
    t = Trace(os.path.join(gold, "trace_%s.json" % name), os.path.join(gold, "trace_%s.npz" % name))
    B = B200Library(); B.B200_SetSeed(42)
    r1 = Replayer(B, t, seed=1234); r1.setup(); single = np.asarray(r1.infer(), dtype=np.float32)
    B.DeleteScheme()
    B = B200Library(); B.B200_SetSeed(42)
    x = DeviceExchange(B, 1 << t.meta["config"]["ckks_params"]["LogN"], rank, world, meta_group=meta)
    rs = ShardedReplayer(B, t, x, seed=1234); rs.setup()
    out = rs.infer()
    lib.B200_Sync()
    flag = [bool(np.array_equal(np.asarray(out, dtype=np.float32), single))] if rank == 0 else [None]
    dist.broadcast_object_list(flag, src=0, group=meta)
    res["sharded_equal"] = flag[0]
    res["sharded_ops_run"] = rs.ops_run
    res["sharded_transfers"] = x.transfers
    res["sharded_bytes"] = int(x.bytes_moved)
    B.DeleteScheme()
    allres = [None] * world
    dist.all_gather_object(allres, res, group=meta)
    if rank == 0:
        json.dump(allres, open(sys.argv[1], "w"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
