"""Ciphertext-sharded multi-GPU execution: one process per GPU, context + keys + encoded diagonals
replicated, independent ciphertexts assigned to ranks, and ONE kind of exchange -- a ciphertext's RNS limbs go from the rank
that computed it to the ranks that need it (the gather of layer outputs before the next linear layer).

The exchange runs on DEVICE memory: the limbs are read and written where the library keeps them
(`B200_CiphertextAcquire` -> device pointer, ordered on the library's main stream with event waits only), wrapped as torch
tensors through `__cuda_array_interface__`, and moved with torch.distributed -- NCCL over NVLink / NVSwitch on the GPU box
-- with the library's stream as torch's current stream, so producers, transfer and consumers are ordered on the device and
the host never waits.  Ciphertext metadata (level, degree, exact scale, slots: 40 bytes) travels host-to-host over a gloo
group, which lets a receiver allocate its buffer as soon as the owner's HOST has enqueued the producing call.

`ShardedReplayer` shards ONE inference across ranks (strong scaling): block rows of a linear layer
(orion/backend/python/lt_evaluator.py:172-193) go to rank row % G, every per-ciphertext call (EvaluatePolynomial, Bootstrap,
element-wise ops: orion/backend/python/poly_evaluator.py:28-31, tensors.py:294-304) runs where its operand lives, and a
ciphertext is broadcast the first time another rank needs it.  Nothing here touches cryptography; results are bit-identical
to the single-rank replay (tests/test_gpu_multi.py).

Images of a batch (bench.py resnet20 / lola2 images/s) need none of this: one image per rank, no exchange.
"""
import numpy as np

from .replay import Replayer


def shard(n_items, rank, world):
    """Indices of the items rank `rank` owns (round-robin: item i -> rank i mod world)."""
    return list(range(rank, n_items, world))


def owner(item, world):
    return item % world


# ---------------------------------------------------------------------------------------------------------------------
# host-staged gather of plain limb arrays (gloo on CPU: the logic test of tests/test_parallel_cpu.py)
# ---------------------------------------------------------------------------------------------------------------------
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


# ---------------------------------------------------------------------------------------------------------------------
# device-resident exchange
# ---------------------------------------------------------------------------------------------------------------------
class _DevWords:
    """A device buffer of `n` 64-bit words at `ptr`, as the CUDA array interface describes it (zero-copy into torch)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<i8", "data": (int(ptr), False), "version": 2}


class DeviceExchange:
    """Moves ciphertexts between the ranks of a torch.distributed job without leaving device memory."""

    def __init__(self, backend, N, rank, world, data_group=None, meta_group=None):
        import ctypes
        import torch
        self.B, self.N, self.rank, self.world = backend, int(N), rank, world
        self.data_group, self.meta_group = data_group, meta_group
        self.torch = torch
        self.stream = torch.cuda.ExternalStream(int(backend.lib.B200_Stream())) if torch.cuda.is_available() else None
        self._hi, self._lo = ctypes.c_double(0), ctypes.c_double(0)
        self.bytes_moved = 0            # payload bytes this rank sent or received
        self.transfers = 0

    def meta_of(self, ct):
        import ctypes
        B = self.B
        B.lib.B200_GetCiphertextScaleParts(ct, ctypes.byref(self._hi), ctypes.byref(self._lo))
        return (int(B.GetCiphertextLevel(ct)), int(B.GetCiphertextDegree(ct)), float(self._hi.value), float(self._lo.value),
                int(B.lib.B200_GetCiphertextLogSlots(ct)))

    def _tensor(self, ct, words, for_write):
        ptr = self.B.lib.B200_CiphertextAcquire(ct, 1 if for_write else 0)
        return self.torch.as_tensor(_DevWords(ptr, words), device="cuda")

    def broadcast(self, ct, src):
        """Rank `src` holds ciphertext handle `ct`; every other rank passes ct=None and gets a new handle with the same limbs
        and metadata.  Returns the local handle."""
        import torch.distributed as dist
        if self.world == 1:
            return ct
        meta = [self.meta_of(ct) if self.rank == src else None]
        dist.broadcast_object_list(meta, src=src, group=self.meta_group)
        level, degree, hi, lo, log_slots = meta[0]
        words = (degree + 1) * (level + 1) * self.N
        if self.rank != src:
            ct = self.B.lib.B200_NewCiphertext(level, degree, hi, lo, log_slots)
        t = self._tensor(ct, words, for_write=self.rank != src)
        with self.torch.cuda.stream(self.stream):
            dist.broadcast(t, src=src, group=self.data_group)
        self.bytes_moved += words * 8
        self.transfers += 1
        return ct

    def all_gather(self, ct_ids, n_items):
        """ct_ids: item index -> local handle for the items this rank owns (item i lives on rank i mod world).  Returns
        item index -> local handle for ALL items (the gather of a layer's output ciphertexts, on device limbs)."""
        out = dict(ct_ids)
        for i in range(n_items):
            out[i] = self.broadcast(out.get(i), owner(i, self.world))
        return out


class HostExchange:
    """The same interface over host objects and any torch.distributed backend (gloo): the placement logic of
    ShardedReplayer is tested with it on CPU ranks, with tests/clear_backend.py standing in for the library."""

    def __init__(self, rank, world, get, put, group=None):
        self.rank, self.world, self.get, self.put, self.group = rank, world, get, put, group
        self.bytes_moved = 0
        self.transfers = 0

    def broadcast(self, ct, src):
        import torch.distributed as dist
        if self.world == 1:
            return ct
        box = [self.get(ct) if self.rank == src else None]
        dist.broadcast_object_list(box, src=src, group=self.group)
        self.transfers += 1
        self.bytes_moved += int(np.asarray(box[0]).nbytes)
        return ct if self.rank == src else self.put(box[0])


def gather_ciphertexts(backend, ct_ids, n_items, rank, world, N, device=None, meta_group=None):
    """Gather of a layer's output ciphertexts on device limbs (NCCL); kept under the name round 1 used."""
    return DeviceExchange(backend, N, rank, world, meta_group=meta_group).all_gather(ct_ids, n_items)


# ---------------------------------------------------------------------------------------------------------------------
# one inference sharded across ranks
# ---------------------------------------------------------------------------------------------------------------------
_LAYER_OPS = ("EvaluateLinearTransform", "AddCiphertext", "AddCiphertextNew", "RescaleNew", "Rescale", "DeleteCiphertext", "Decrypt",
              "Decode", "DeletePlaintext")
_UNARY_NEW = ("Negate", "RescaleNew", "CloneCiphertext", "RotateNew", "AddScalarNew", "SubScalarNew", "MulScalarIntNew", "MulScalarFloatNew",
              "EvaluatePolynomial", "Bootstrap", "AddPlaintextNew", "SubPlaintextNew", "MulPlaintextNew")
_UNARY_INPLACE = ("Rescale", "Rotate", "AddScalar", "SubScalar", "MulScalarInt", "MulScalarFloat", "AddPlaintext", "SubPlaintext",
                  "MulPlaintext", "ModDropCiphertext", "SetCiphertextScale")
_BINARY_NEW = ("AddCiphertextNew", "SubCiphertextNew", "MulRelinCiphertextNew")
_BINARY_INPLACE = ("AddCiphertext", "SubCiphertext", "MulRelinCiphertext")


def plan_rows(ops, start):
    """Static pass over the inference part of a trace: for every EvaluateLinearTransform call, the block row it belongs to.
    A linear layer is a run of the calls lt_evaluator.py:172-193 issues (transforms, row accumulation, one rescale per row,
    frees, debug decrypts); inside it a new row starts whenever the FIRST input ciphertext of the layer comes round again."""
    rows, layer_first, row, in_layer = {}, None, 0, False
    for i in range(start, len(ops)):
        name, args, _ = ops[i]
        if name == "EvaluateLinearTransform":
            if not in_layer:
                in_layer, layer_first, row = True, args[1], 0
            elif args[1] == layer_first:
                row += 1
            rows[i] = row
        elif name not in _LAYER_OPS:
            in_layer = False
    return rows


class ShardedReplayer(Replayer):
    """Replays the inference part of a trace with its ciphertexts spread over the ranks.  Every rank walks the whole op list
    (so all ranks agree on who computes what without talking); a rank issues only the calls assigned to it."""

    def __init__(self, backend, trace, exchange, seed=1234):
        super().__init__(backend, trace, seed=seed)
        self.x = exchange
        self.rank, self.world = exchange.rank, exchange.world
        self.rows = plan_rows(trace.ops, trace.n_compile)
        self.home = {}           # recorded ciphertext handle -> rank that computes on it
        self.holders = {}        # recorded ciphertext handle -> set of ranks holding a valid copy
        self.ops_run = 0

    # -- data movement --
    def _need(self, h, on_rank):
        """Make recorded ciphertext h available on `on_rank` (a broadcast from its home the first time another rank needs it)."""
        if on_rank in self.holders[h]:
            return
        src = self.home[h]
        self.ct[h] = self.x.broadcast(self.ct.get(h) if self.rank == src else None, src)
        self.holders[h] = set(range(self.world))

    def _produced(self, h, on_rank):
        """Recorded ciphertext h was (over)written on `on_rank`: copies elsewhere are stale."""
        if self.rank != on_rank and h in self.ct and h in self.holders and self.rank in self.holders[h]:
            self.B.DeleteCiphertext(self.ct.pop(h))
        self.home[h], self.holders[h] = on_rank, {on_rank}

    def _run(self, i, op):
        name, args, ret = op
        if i < self.t.n_compile or self.world == 1:
            return super()._run(i, op)
        me = self.rank
        if name in ("Encode", "DeletePlaintext", "Decode"):
            if name == "Decode" and me != 0:
                return
            return super()._run(i, op)
        if name == "Encrypt":                                    # replicated: same seed, same sampling event on every rank
            super()._run(i, op)
            self.home[ret], self.holders[ret] = 0, set(range(self.world))
            return
        if name == "Decrypt":
            if self.skip_debug and i != self.last_decrypt:
                return
            self._need(args[0], 0)
            if me == 0:
                super()._run(i, op)
            return
        if name == "DeleteCiphertext":
            h = args[0]
            if h in self.holders:
                if me in self.holders[h] and h in self.ct:
                    self.B.DeleteCiphertext(self.ct.pop(h))
                self.ct.pop(h, None)
                del self.holders[h]; del self.home[h]
            return
        if name == "EvaluateLinearTransform":
            e = self.rows[i] % self.world
            self._need(args[1], e)
            if me == e:
                super()._run(i, op); self.ops_run += 1
            self._produced(ret, e)
            return
        if name in _BINARY_NEW or name in _BINARY_INPLACE:
            a, b = args[0], args[1]
            e = self.home[a]
            self._need(b, e)
            if me == e:
                super()._run(i, op); self.ops_run += 1
            self._produced(ret, e)
            return
        if name in _UNARY_NEW or name in _UNARY_INPLACE:
            a = args[0]
            e = self.home[a]
            if me == e:
                super()._run(i, op); self.ops_run += 1
            if name == "RescaleNew":                             # rescales its input in place AND returns a copy (LG/evaluator.go:91-99)
                self._produced(a, e)
            self._produced(a if name in _UNARY_INPLACE else ret, e)
            return
        raise RuntimeError("ShardedReplayer: no placement rule for backend call " + name)

    def infer(self):
        ct0, pt0 = self._snap
        for k, v in list(self.ct.items()):
            if k not in ct0 or ct0[k] != v:
                self.B.DeleteCiphertext(v)
        for k, v in list(self.pt.items()):
            if k not in pt0 or pt0[k] != v:
                self.B.DeletePlaintext(v)
        self.ct, self.pt = dict(ct0), dict(pt0)
        self.home, self.holders = {}, {}
        self.run(self.t.n_compile, len(self.t.ops))
        return self.output
