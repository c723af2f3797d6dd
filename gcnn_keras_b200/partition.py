"""Single large graph across GPUs (BASELINE config 5: GCN on 2 M nodes / 60 M edges over 8 x B200).

1-D partition by contiguous RECEIVER range: rank r owns nodes [r*P, (r+1)*P), their output rows and all their
in-edges (a slice of the receiver-sorted edge list).  Per layer there is exactly one exchange, and it moves only the
rows that are actually referenced across the cut (HaloPlan, built once per graph):
  forward : every rank sends each peer the rows of its block that peer's edges read (row gather -> all-to-all);
            the edges whose sender is LOCAL are aggregated while that exchange is in flight, the edges with a
            remote sender afterwards, from the compact halo buffer;
  backward: the exact transpose -- partial d(features) of the halo rows go back to their owners (all-to-all) while
            the local edges' transpose runs, and are summed into the owner's rows in a fixed order (CSR segment sum).
No [E, F] tensor is ever materialised.  `exchange="allgather"` keeps the round-1 scheme (all-gather of the whole
[N, F] feature matrix / reduce-scatter of the whole gradient) as the comparand of the tests and of bench.py.
The reference has no partitioning at all (kgcnn/layers/conv/gcn_conv.py:85-90 materialises [E, F] three times on
one device).
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


def _all_to_all_rows(send_buf, send_counts, recv_counts):
    """Rows of send_buf, grouped by destination rank (send_counts rows each), to their destinations; returns
    (work handle or None, receive buffer grouped by source rank).  NCCL: one asynchronous all_to_all_single.
    gloo (the CPU tests of the host logic): emulated with an all-gather of the padded send buffers."""
    dist = _dist()
    world, rank = dist.get_world_size(), dist.get_rank()
    tail = tuple(send_buf.shape[1:])
    out = send_buf.new_empty((int(sum(recv_counts)),) + tail)
    if dist.get_backend() == "nccl":
        work = dist.all_to_all_single(out, send_buf, output_split_sizes=list(recv_counts),
                                      input_split_sizes=list(send_counts), async_op=True)
        return work, out
    sizes = [None] * world
    dist.all_gather_object(sizes, [int(c) for c in send_counts])
    longest = max(max(sum(sz) for sz in sizes), 1)
    padded = send_buf.new_zeros((longest,) + tail)
    padded[:send_buf.shape[0]] = send_buf
    everyone = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(everyone, padded)
    off = 0
    for q in range(world):
        start = sum(sizes[q][:rank])
        n = sizes[q][rank]
        out[off:off + n] = everyone[q][start:start + n]
        off += n
    return None, out


class HaloPlan:
    """Which rows cross the cut, computed ONCE per graph from this rank's sender column.

    halo_ids   : sorted unique GLOBAL ids of the remote senders of this rank's edges = the row order of the halo buffer
    need_counts: rows of the halo buffer owned by each rank (receive split sizes)
    send_ids   : LOCAL row ids this rank sends, grouped by destination rank (int32, device); send_counts per rank
    rel_back   : CSR relation send_ids -> local row (sums the returned partial gradients in a fixed order)
    exchange_needed: True if ANY rank has rows crossing the cut (MAX over the group at plan build, so identical on
                 every rank): the all-to-all is a collective, and a rank with nothing to send or receive -- an isolated
                 shard, a last rank that is all padding -- still has to join it with zero counts."""

    def __init__(self, send_global, rank, world, per, device):
        send_global = np.asarray(send_global, dtype=np.int64)
        lo = rank * per
        self.rank, self.world, self.per = rank, world, per
        self.is_local = (send_global >= lo) & (send_global < lo + per)
        self.halo_ids = np.unique(send_global[~self.is_local])
        owner = self.halo_ids // per
        self.need_counts = [int(c) for c in np.bincount(owner, minlength=world)[:world]]
        dist = _dist()
        if dist is None or world == 1:
            self.send_counts = [0] * world
            send_ids = np.zeros((0,), dtype=np.int64)
        else:
            # tell every owner which of its rows this rank reads: counts first, then the ids themselves
            counts_out = torch.tensor(self.need_counts, dtype=torch.int64, device=device)
            counts_in = self._a2a_sync(counts_out.reshape(-1, 1), [1] * world, [1] * world)
            self.send_counts = [int(c) for c in counts_in.reshape(-1).tolist()]
            ids_out = torch.as_tensor(self.halo_ids, dtype=torch.int64, device=device).reshape(-1, 1)
            ids_in = self._a2a_sync(ids_out, self.need_counts, self.send_counts)
            send_ids = ids_in.reshape(-1).cpu().numpy() - lo
            if len(send_ids) and (send_ids.min() < 0 or send_ids.max() >= per):
                raise IndexError("halo plan: a peer asked for a row outside this rank's block")
        self.n_halo, self.n_send = int(len(self.halo_ids)), int(len(send_ids))
        self.send_ids = torch.as_tensor(send_ids, dtype=torch.int32, device=device)
        self.rel_back = Relation(self.send_ids, per) if self.n_send else None
        self.exchange_needed = False
        if dist is not None and world > 1:
            crossing = torch.tensor([self.n_halo + self.n_send], dtype=torch.int64, device=device)
            dist.all_reduce(crossing, op=dist.ReduceOp.MAX)
            self.exchange_needed = bool(int(crossing.item()) > 0)

    def exchange_rows(self, rows, reverse=False):
        """The per-layer all-to-all: forward sends `rows` (grouped by destination, send_counts) and returns the halo
        buffer; reverse=True sends halo-shaped rows back to their owners.  Returns (work handle or None, buffer), or
        (None, None) when no rank of the group has a row crossing the cut.  Every rank calls it or none does."""
        if not self.exchange_needed:
            return None, None
        if reverse:
            return _all_to_all_rows(rows, self.need_counts, self.send_counts)
        return _all_to_all_rows(rows, self.send_counts, self.need_counts)

    @staticmethod
    def _a2a_sync(buf, send_counts, recv_counts):
        work, out = _all_to_all_rows(buf, send_counts, recv_counts)
        if work is not None:
            work.wait()
        return out

    def halo_index(self, send_global):
        """Remote senders -> row of the halo buffer (int64 NumPy; only meaningful where ~is_local)."""
        return np.searchsorted(self.halo_ids, send_global)


class _Aggregate(torch.autograd.Function):
    """y = A h on a partitioned graph (transpose=False: rows = this rank's receivers) or dx = A^T g (transpose=True),
    with the halo exchange overlapped with the local part.  Each is the other's derivative."""

    @staticmethod
    def forward(ctx, h, graph, transpose):
        ctx.graph, ctx.transpose = graph, transpose
        g, plan = graph, graph.halo
        h = h.contiguous()
        if not transpose:
            work = halo = None
            if plan.exchange_needed:                                       # a group-wide decision, never rank-local
                send_buf = F.raw_gather_rows(h, plan.send_ids) if plan.n_send else h.new_empty((0, h.shape[1]))
                work, halo = plan.exchange_rows(send_buf)
            y = F.raw_gms(h, g.w_loc, g.recv_loc, g.send_loc)              # local senders, while the rows travel
            if work is not None:
                work.wait()
            if plan.n_halo and g.recv_rem.n_elements:
                y += F.raw_gms(halo, g.w_rem, g.recv_rem, g.send_rem)
            return y
        work = back = None
        if plan.exchange_needed:
            g_halo = F.raw_gms(h, g.w_rem, g.send_rem, g.recv_rem) if plan.n_halo else h.new_empty((0, h.shape[1]))
            work, back = plan.exchange_rows(g_halo, reverse=True)
        dx = F.raw_gms(h, g.w_loc, g.send_loc, g.recv_loc)                 # local senders, while the rows travel
        if work is not None:
            work.wait()
        if plan.n_send:
            dx += F.raw_gms(back, None, plan.rel_back, None)               # fixed order: by source rank, then id
        return dx

    @staticmethod
    def backward(ctx, grad):
        return _Aggregate.apply(grad, ctx.graph, not ctx.transpose), None, None


class PartitionedGraph:
    """This rank's shard of one large receiver-sorted graph (device tensors + CSR / CSC relations)."""

    def __init__(self, n_nodes, recv, send, weights, rank, world, device, exchange="halo"):
        recv = np.asarray(recv)
        self.n_nodes, self.rank, self.world = int(n_nodes), rank, world
        self.per, bounds = partition_bounds(n_nodes, world)
        self.lo, self.hi = bounds[rank]
        e0, e1 = slice_edges_for_rank(recv, rank, self.per)
        self.e0, self.e1 = e0, e1
        self.n_edges = e1 - e0
        self.set_shard(recv[e0:e1] - rank * self.per, np.asarray(send)[e0:e1],
                       None if weights is None else np.asarray(weights)[e0:e1], device, exchange)

    @classmethod
    def from_shard(cls, n_nodes, recv_local, send_global, weights, rank, world, device, exchange="halo"):
        """For callers that generate only their own in-edges (receiver ids already local to the rank's block)."""
        self = cls.__new__(cls)
        self.n_nodes, self.rank, self.world = int(n_nodes), rank, world
        self.per, bounds = partition_bounds(n_nodes, world)
        self.lo, self.hi = bounds[rank]
        self.n_edges = len(recv_local)
        self.set_shard(recv_local, send_global, weights, device, exchange)
        return self

    def set_shard(self, recv_local, send_global, weights, device, exchange):
        self.exchange = exchange
        i32 = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.int32, device=device)
        f32 = lambda a: None if a is None else torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32,
                                                               device=device).reshape(-1, 1)
        recv_local, send_global = np.asarray(recv_local), np.asarray(send_global)
        self.rel_recv = Relation(i32(recv_local), self.per, is_sorted=None)
        self.rel_send = Relation(i32(send_global), self.per * self.world)
        self.weights = f32(weights)
        self.halo = None
        if exchange == "halo":
            plan = self.halo = HaloPlan(send_global, self.rank, self.world, self.per, device)
            loc, rem = plan.is_local, ~plan.is_local
            w = None if weights is None else np.asarray(weights)
            self.recv_loc = Relation(i32(recv_local[loc]), self.per, is_sorted=None)
            self.send_loc = Relation(i32(send_global[loc] - self.rank * self.per), self.per)
            self.w_loc = None if w is None else f32(w[loc])
            self.recv_rem = Relation(i32(recv_local[rem]), self.per, is_sorted=None)
            self.send_rem = Relation(i32(plan.halo_index(send_global[rem])), max(plan.n_halo, 1))
            self.w_rem = None if w is None else f32(w[rem])

    def pad_rows(self, x_local):
        """[hi-lo, F] -> [per, F] (zero rows for the padding of the last rank)."""
        if x_local.shape[0] == self.per:
            return x_local
        pad = x_local.new_zeros((self.per - x_local.shape[0],) + tuple(x_local.shape[1:]))
        return torch.cat([x_local, pad], dim=0)

    def aggregate(self, h_local):
        """y_i = sum_{e -> i} a_e h[s_e] for the rows this rank owns; h_local is this rank's [per, F] block."""
        if self.halo is not None:
            return _Aggregate.apply(h_local, self, False)
        h_full = halo_all_gather(h_local)
        return F.gather_mul_segment_sum(h_full, self.weights, self.rel_recv, self.rel_send)

    def warm(self):
        """Build every CSR / CSC plan now (outside a timed region)."""
        rels = [self.rel_recv, self.rel_send] if self.halo is None else \
            [self.recv_loc, self.send_loc, self.recv_rem, self.send_rem]
        for a, b in zip(rels[0::2], rels[1::2]):
            if a.n_elements:
                a.rowptr, b.rowptr, a.permuted_ids(b), b.permuted_ids(a)
        if self.halo is not None and self.halo.rel_back is not None:
            self.halo.rel_back.rowptr


def gcn_layer_partitioned(graph, x_local, kernel, bias, activation="relu"):
    """GCN.call (kgcnn/layers/conv/gcn_conv.py:85-90) on a partitioned graph: Dense on the local rows,
    halo exchange, fused aggregation, activation."""
    h = F.dense(x_local, kernel, bias, "linear")
    y = graph.aggregate(h)
    return F.bias_act(y, None, activation)
