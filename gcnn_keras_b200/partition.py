"""Single large graph across GPUs (BASELINE config 5: GCN on 2 M nodes / 60 M edges over 8 x B200).

1-D partition by contiguous RECEIVER range: rank r owns nodes [r*P, (r+1)*P), their output rows and all their
in-edges (a slice of the receiver-sorted edge list).  Per layer there is exactly one exchange:
  forward : halo all-gather of the sender features (every rank contributes its [P, F] block)
  backward: reduce-scatter(sum) of the partial d(features) -- the exact transpose
and no [E, F] tensor is ever materialised (the fused gather*weight->segment-sum kernel reads the gathered rows
straight from the all-gathered block).  The reference has no partitioning at all (kgcnn/layers/conv/gcn_conv.py:85-90
materialises [E, F] three times on one device).
"""
import numpy as np
import torch

from . import functional as F
from .plan import Relation


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


class _HaloAllGather(torch.autograd.Function):
    """[P, F] block per rank -> [world*P, F] on every rank; backward = reduce-scatter(sum)."""

    @staticmethod
    def forward(ctx, local):
        dist = _dist()
        if dist is None or dist.get_world_size() == 1:
            return local
        world = dist.get_world_size()
        out = local.new_empty((world * local.shape[0],) + tuple(local.shape[1:]))
        dist.all_gather_into_tensor(out, local.contiguous())
        return out

    @staticmethod
    def backward(ctx, g):
        return _HaloReduceScatter.apply(g)


class _HaloReduceScatter(torch.autograd.Function):
    """[world*P, F] partial sums -> this rank's [P, F] block of the total; backward = all-gather."""

    @staticmethod
    def forward(ctx, full):
        dist = _dist()
        if dist is None or dist.get_world_size() == 1:
            return full
        world, rank = dist.get_world_size(), dist.get_rank()
        per = full.shape[0] // world
        if dist.get_backend() == "nccl":
            out = full.new_empty((per,) + tuple(full.shape[1:]))
            dist.reduce_scatter_tensor(out, full.contiguous(), op=dist.ReduceOp.SUM)
            return out
        tmp = full.clone()                       # gloo (CPU tests): all-reduce + slice
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM)
        return tmp[rank * per:(rank + 1) * per].clone()

    @staticmethod
    def backward(ctx, g):
        return _HaloAllGather.apply(g)


def halo_all_gather(local):
    return _HaloAllGather.apply(local)


def partition_bounds(n_nodes, world):
    """Rows per rank (padded so that every rank owns the same count) and the [lo, hi) of real rows."""
    per = -(-n_nodes // world)
    return per, [(r * per, min(n_nodes, (r + 1) * per)) for r in range(world)]


def slice_edges_for_rank(recv_sorted, rank, per):
    """Edge range [e0, e1) whose receiver lies in rank's node range (recv ascending; host-side NumPy)."""
    lo, hi = rank * per, (rank + 1) * per
    return int(np.searchsorted(recv_sorted, lo, side="left")), int(np.searchsorted(recv_sorted, hi, side="left"))


class PartitionedGraph:
    """This rank's shard of one large receiver-sorted graph (device tensors + CSR / CSC relations)."""

    def __init__(self, n_nodes, recv, send, weights, rank, world, device):
        recv = np.asarray(recv)
        self.n_nodes, self.rank, self.world = int(n_nodes), rank, world
        self.per, bounds = partition_bounds(n_nodes, world)
        self.lo, self.hi = bounds[rank]
        e0, e1 = slice_edges_for_rank(recv, rank, self.per)
        self.e0, self.e1 = e0, e1
        self.n_edges = e1 - e0
        recv_local = torch.as_tensor(recv[e0:e1] - rank * self.per, dtype=torch.int32, device=device)
        send_global = torch.as_tensor(np.asarray(send)[e0:e1], dtype=torch.int32, device=device)
        self.rel_recv = Relation(recv_local, self.per, is_sorted=None)
        self.rel_send = Relation(send_global, self.per * world)
        self.weights = None if weights is None else torch.as_tensor(
            np.asarray(weights)[e0:e1], dtype=torch.float32, device=device).reshape(-1, 1)

    def pad_rows(self, x_local):
        """[hi-lo, F] -> [per, F] (zero rows for the padding of the last rank)."""
        if x_local.shape[0] == self.per:
            return x_local
        pad = x_local.new_zeros((self.per - x_local.shape[0],) + tuple(x_local.shape[1:]))
        return torch.cat([x_local, pad], dim=0)

    def aggregate(self, h_local):
        """y_i = sum_{e -> i} a_e h[s_e] for the rows this rank owns; h_local is this rank's [per, F] block."""
        h_full = halo_all_gather(h_local)
        return F.gather_mul_segment_sum(h_full, self.weights, self.rel_recv, self.rel_send)


def gcn_layer_partitioned(graph, x_local, kernel, bias, activation="relu"):
    """GCN.call (kgcnn/layers/conv/gcn_conv.py:85-90) on a partitioned graph: Dense on the local rows,
    halo exchange, fused aggregation, activation."""
    h = F.dense(x_local, kernel, bias, "linear")
    y = graph.aggregate(h)
    return F.bias_act(y, None, activation)
