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
