"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch).

Batched molecular graphs shard by graph (every op on the hot path is confined to one graph:
kgcnn/ops/partition.py:140-155, kgcnn/layers/pooling.py:215-218), so the data path has NO collective; the only
exchange is ONE bucketed all-reduce of the weight gradients per step (~0.5-1 M parameters = 2-4 MB fp32,
latency-bound), captured inside the step's CUDA graph.  The reference has no distributed code at all.
"""
import torch


def world_info():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n_items, rank, world):
    """Contiguous block of graphs [lo, hi) owned by `rank` (balanced to within one graph)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_graphs_by_cost(costs, world):
    """Contiguous blocks of graphs with balanced total cost (edge or triple count): prefix-sum split points.
    Returns a list of (lo, hi) per rank."""
    import numpy as np
    costs = np.asarray(costs, dtype=np.float64)
    csum = np.concatenate([[0.0], np.cumsum(costs)])
    total = csum[-1]
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(csum, total * r / world, side="left")))
    cuts.append(len(costs))
    cuts = [min(max(c, 0), len(costs)) for c in cuts]
    for i in range(1, len(cuts)):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def allreduce_gradients(grads, group=None, average=True):
    """Single-bucket all-reduce of a list of gradient tensors (returns new tensors, same shapes)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return list(grads)
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat = flat / dist.get_world_size(group)
    out, off = [], 0
    for g in grads:
        n = g.numel()
        out.append(flat[off:off + n].view_as(g))
        off += n
    return out


def allreduce_flat(flat, group=None, average=True):
    """All-reduce of the already flat gradient buffer (one NCCL call, captured inside the step's CUDA graph)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return flat
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat / dist.get_world_size(group) if average else flat


def broadcast_parameters(module, src=0, group=None):
    """Make every replica start from rank `src`'s weights."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    for p in module.parameters():
        if p.data.is_contiguous():
            dist.broadcast(p.data, src=src, group=group)
        else:                                  # strided views of a flat / padded buffer (FlatAdam, relational kernels)
            t = p.data.contiguous()
            dist.broadcast(t, src=src, group=group)
            p.data.copy_(t)


def broadcast_flat(flat, src=0, group=None):
    """One broadcast of a flat parameter buffer (FlatAdam.flat; the parameters are views of it)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    dist.broadcast(flat, src=src, group=group)


class PeerAllReduce:
    """All-reduce of the flat gradient buffer over NVLink peer memory instead of an NCCL collective.

    The gradient buffer itself lives in symmetric (peer-mapped) memory (`torch.distributed._symmetric_memory`: plumbing
    only -- allocation, the handle exchange and the device-side barrier); the reduction is OUR kernel
    (`kgcnn_allreduce_peer`, csrc/collective.cu): after a barrier every rank reads all ranks' buffers directly and adds
    them in rank order, so the replicas stay bitwise identical.  A second barrier keeps the buffers unchanged until every
    peer has read them.  Both barriers and the kernel are stream-ordered and captured inside the step's CUDA graph.

    `create()` returns None when the process group, the device or the torch build cannot provide symmetric memory
    (one rank, gloo, KGCNN_PEER_ALLREDUCE=0) or when the group is larger than MAX_WORLD_AUTO ranks: the caller then
    keeps the NCCL all-reduce."""

    MAX_WORLD_AUTO = 4

    def __init__(self, numel, device, group):
        import torch.distributed._symmetric_memory as symm
        self.group = group
        self.buffer = symm.empty(int(numel), dtype=torch.float32, device=device)
        self.buffer.zero_()
        self.handle = symm.rendezvous(self.buffer, group.group_name)          # collective
        self.world = int(self.handle.world_size)
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.out = torch.empty(int(numel), dtype=torch.float32, device=device)
        self.numel = int(numel)

    @classmethod
    def create(cls, numel, device, group=None):
        import os
        import torch.distributed as dist
        choice = os.environ.get("KGCNN_PEER_ALLREDUCE", "auto")
        if choice == "0":
            return None
        if not (dist.is_available() and dist.is_initialized()):
            return None
        group = group or dist.group.WORLD
        if dist.get_world_size(group) < 2 or dist.get_backend(group) != "nccl" or torch.device(device).type != "cuda":
            return None
        # every rank reads world x n bytes and the two barriers are all-to-all: measured on C2 (1.9 MB) 2.371 -> 2.342 ms
        # at 2 GPUs but 2.376 -> 2.413 ms at 8, where NCCL reduces inside the switch.  "1" forces it for any world size.
        if choice != "1" and dist.get_world_size(group) > cls.MAX_WORLD_AUTO:
            return None
        ok = torch.zeros(1, device=device)
        peer = None
        try:
            peer = cls(numel, device, group)
            ok.fill_(1.0)
        except Exception:
            peer = None
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)                # every rank or none
        return peer if float(ok) == 1.0 else None

    def reduce(self, average=True):
        """Sum (mean) over the ranks of `self.buffer` -> `self.out` (the buffer itself stays untouched)."""
        import ctypes
        from . import _lib
        arr = (ctypes.c_void_p * self.world)(*self.ptrs)
        self.handle.barrier(channel=0)            # every rank's gradients are complete
        _lib.check(_lib.lib().kgcnn_allreduce_peer(arr, self.world, self.numel, (1.0 / self.world) if average else 1.0,
                                                   self.out.data_ptr(), torch.cuda.current_stream().cuda_stream),
                   "kgcnn_allreduce_peer")
        self.handle.barrier(channel=1)            # every rank has read them: the buffers may be overwritten
        return self.out
