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
