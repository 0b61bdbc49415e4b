"""world_size-2 gloo test of the ciphertext-sharding / gather logic (orion_b200/parallel.py) on CPU."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_items, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from orion_b200 import parallel
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    mine = parallel.shard(n_items, rank, world)
    local = {i: (np.arange(2 * 3 * 64, dtype=np.uint64).reshape(2, 3, 64) + (i << 40) + (1 << 63)) for i in mine}
    allv = parallel.gather_limbs(local, n_items, rank, world)
    ok = all(np.array_equal(allv[i], np.arange(2 * 3 * 64, dtype=np.uint64).reshape(2, 3, 64) + (i << 40) + (1 << 63))
             for i in range(n_items))
    q.put((rank, mine, ok))
    dist.destroy_process_group()


def test_shard_is_a_partition():
    from orion_b200 import parallel
    for n in (0, 1, 5, 8, 17):
        for w in (1, 2, 4, 8):
            got = sorted(i for r in range(w) for i in parallel.shard(n, r, w))
            assert got == list(range(n))
            assert all(parallel.owner(i, w) == r for r in range(w) for i in parallel.shard(n, r, w))


def test_gather_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 500
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 5, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[2] for r in res), res
    assert sorted(sum((r[1] for r in res), [])) == [0, 1, 2, 3, 4]


def _sharded_worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from clear_backend import ClearBackend
    from orion_b200.parallel import HostExchange, ShardedReplayer
    from orion_b200.replay import Replayer, Trace
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    gold = os.path.join(ROOT, "tests", "golden")
    t = Trace(os.path.join(gold, "trace_resnet20.json"), os.path.join(gold, "trace_resnet20.npz"))
    B = ClearBackend()
    x = HostExchange(rank, world, get=lambda h: B.obj[h], put=lambda a: B._new(np.array(a)))
    r = ShardedReplayer(B, t, x, seed=1234)
    r.setup()
    out = r.infer()
    out2 = r.infer()                                            # repeatable: handles of the first inference are released
    single = None
    if rank == 0:
        r1 = Replayer(ClearBackend(), t, seed=1234); r1.setup(); single = np.asarray(r1.infer())
    ok = rank != 0 or (np.array_equal(np.asarray(out), single) and np.array_equal(np.asarray(out2), single))
    q.put((rank, bool(ok), r.ops_run, x.transfers, len([h for h in B.obj]) ))
    dist.destroy_process_group()


def test_sharded_replay_world2_gloo_matches_single_rank():
    """One ResNet-20 inference sharded over 2 CPU ranks (gloo; tests/clear_backend.py as the backend): the placement rules of
    orion_b200/parallel.py ShardedReplayer -- block rows of a linear layer round-robin, per-ciphertext calls where the operand
    lives, broadcast on first remote use, stale copies dropped after in-place calls -- give exactly the single-rank result,
    and both ranks do a real share of the calls."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 90
    procs = [ctx.Process(target=_sharded_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] for r in res), res
    ops = [r[2] for r in res]
    assert min(ops) > 0.25 * max(ops), res
    assert res[0][3] == res[1][3] and res[0][3] > 0            # both ranks saw the same broadcasts


def test_plan_rows_reads_the_block_structure_of_the_linear_layers():
    from orion_b200.parallel import plan_rows
    from orion_b200.replay import Trace
    gold = os.path.join(ROOT, "tests", "golden")
    t = Trace(os.path.join(gold, "trace_resnet20.json"), os.path.join(gold, "trace_resnet20.npz"))
    rows = plan_rows(t.ops, t.n_compile)
    assert len(rows) == 147                                      # every EvaluateLinearTransform call has a row
    # ResNet-20 at logN=13: 16-channel stage = 4 ciphertexts per tensor (rows 0..3), 32-channel = 2, 64-channel = 1
    assert max(rows.values()) == 3
    from collections import Counter
    c = Counter(rows.values())
    assert c[0] > c[1] > c[2] == c[3]
