"""Ciphertext-sharded multi-GPU execution: one process per GPU, context + keys replicated (regenerated
from the seed), independent ciphertexts / images assigned round-robin to ranks, no collective on the data path; the
only exchange is a gather of output ciphertexts so that every rank (or rank 0) holds a layer's full output.

The collective runs on raw RNS limbs ([degree+1][level+1][N] uint64) through torch.distributed: NCCL over NVLink on
the GPU box, gloo in the CPU tests.  Nothing here touches cryptography."""
import numpy as np


def shard(n_items, rank, world):
    """Indices of the items rank `rank` owns (round-robin: item i -> rank i mod world)."""
    return list(range(rank, n_items, world))


def owner(item, world):
    return item % world


def gather_limbs(local, n_items, rank, world, group=None, device=None):
    """All-gather of per-item limb arrays.  `local` maps item index -> uint64 ndarray (all items share one shape).
    Returns the list of all n_items arrays, in item order, on every rank."""
    import torch
    import torch.distributed as dist
    mine = shard(n_items, rank, world)
    assert sorted(local) == mine, "each rank must contribute exactly the items it owns"
    if world == 1:
        return [np.asarray(local[i]) for i in range(n_items)]
    shape = np.asarray(local[mine[0]]).shape if mine else None
    shapes = [None] * world
    dist.all_gather_object(shapes, shape, group=group)
    shape = next(s for s in shapes if s is not None)
    per_rank = (n_items + world - 1) // world
    buf = np.zeros((per_rank,) + tuple(shape), dtype=np.int64)
    for k, i in enumerate(mine):
        buf[k] = np.asarray(local[i]).view(np.int64)
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    res = []
    for i in range(n_items):
        r, k = owner(i, world), i // world
        res.append(out[r][k].cpu().numpy().view(np.uint64))
    return res


def gather_ciphertexts(backend, ct_ids, n_items, rank, world, N, device=None):
    """Export the locally-owned output ciphertexts, all-gather their limbs, import the remote ones.
    ct_ids: item index -> local ciphertext handle.  Returns item index -> handle for all items."""
    local, meta = {}, {}
    for i, ct in ct_ids.items():
        level, deg = backend.GetCiphertextLevel(ct), backend.GetCiphertextDegree(ct)
        a = np.empty((deg + 1, level + 1, N), dtype=np.uint64)
        backend.B200_ExportCiphertext(ct, a)
        local[i] = a
        meta[i] = (level, deg, backend.B200_GetCiphertextScaleF64(ct))
    import torch.distributed as dist
    metas = [None] * world
    if world > 1:
        dist.all_gather_object(metas, meta)
    else:
        metas = [meta]
    allmeta = {}
    for m in metas:
        allmeta.update(m)
    limbs = gather_limbs(local, n_items, rank, world, device=device)
    out = dict(ct_ids)
    for i in range(n_items):
        if i not in out:
            level, deg, scale = allmeta[i]
            out[i] = backend.B200_ImportCiphertext(np.ascontiguousarray(limbs[i]), level, deg, scale, 0)
    return out
