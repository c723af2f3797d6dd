"""Per-batch graph plan.

The reference recomputes ``partition_row_indexing`` (kgcnn/ops/partition.py:97-161) in every gather and
every pooling call and ``tf.argsort`` + permutation gathers (kgcnn/layers/pooling.py:63-68) in every
pooling call, although the index tensor is the same object for all layers of a model.  Here the batch
indexed int32 indices, the receiver CSR, the sender CSC and the permutations are built ONCE per index
tensor and cached on it (keyed by tensor identity + node partition identity).
"""
import weakref

import torch

from . import functional as F
from ._lib import KgcnnError


class PlanMode:
    """How relations decide sortedness.
    None (default)  : check on the device and read the flag on the host (one sync per relation and batch).
    'record'        : same, and remember the answers in creation order (eager warm-up before a capture).
    'capture'       : no host sync (CUDA-graph capture): trust the recorded answers, write the device-side
                      check flags into `flags` so the host can validate the assumption after each replay."""

    MAX_FLAGS = 64

    def __init__(self, kind, hints=None, device=None):
        self.kind = kind
        self.hints = [] if hints is None else list(hints)
        self.cursor = 0
        self.flags = []          # (assumed_sorted, device int32[2]) per relation, capture mode
        # capture mode: ONE guard buffer [MAX_FLAGS, 2] (zeroed at the start of every replay by the captured step)
        # that the relations write their check flags into, and a mask of the words whose being non-zero violates
        # the recorded assumptions; the optimizer kernel reads both and skips the update when they intersect.
        self.flag_buf = self.mask_buf = None
        if kind == "capture" and device is not None:
            self.flag_buf = torch.zeros((self.MAX_FLAGS, 2), dtype=torch.int32, device=device)
            self.mask_buf = torch.zeros((self.MAX_FLAGS, 2), dtype=torch.int32, device=device)

    def next_flags(self):
        if self.flag_buf is None:
            return None
        if len(self.flags) >= self.MAX_FLAGS:
            raise RuntimeError("more than %d relations in one captured step" % self.MAX_FLAGS)
        return self.flag_buf[len(self.flags)]

    def finalize_mask(self):
        """After the capture: which guard words mean 'violated' (unsorted only where sortedness was assumed)."""
        if self.mask_buf is not None and self.flags:
            m = torch.tensor([[1 if a else 0, 1] for a, _ in self.flags], dtype=torch.int32)
            self.mask_buf[:len(self.flags)].copy_(m)

    def violated(self, host_flags):
        """host_flags: list of int32[2] tensors on the host, same order as self.flags."""
        for (assumed_sorted, _), f in zip(self.flags, host_flags):
            if int(f[1]) or (assumed_sorted and int(f[0])):
                return True
        return False


_MODE = None


class plan_mode:
    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        global _MODE
        self._prev, _MODE = _MODE, self.mode
        return self.mode

    def __exit__(self, *exc):
        global _MODE
        _MODE = self._prev


def current_mode():
    return _MODE


def clear_plan_cache():
    _PLAN_CACHE.clear()


class Relation:
    """'element e belongs to target ids[e]' with its CSR (rowptr, stable perm); e.g. edge -> receiver."""

    def __init__(self, ids, n_targets, rowptr=None, perm=None, is_sorted=None):
        self.ids = ids                    # int32 [E] contiguous, batch indexing
        self.n_targets = int(n_targets)
        self.n_elements = int(ids.shape[0])
        self._rowptr = rowptr
        self._perm = perm
        self._is_sorted = is_sorted
        self._permuted = {}               # id(other relation) -> other.ids[perm]
        self._inv_count = None
        self.long_rows = self.n_long = None
        self.long_cap = 0

    def _build(self):
        if self._rowptr is not None:
            return
        if self._is_sorted is None:
            mode = _MODE
            capture = mode is not None and mode.kind == "capture"
            flags = F.raw_check_sorted(self.ids, self.n_targets, flags=mode.next_flags() if capture else None,
                                       clamp=capture)
            if capture:
                if mode.cursor >= len(mode.hints):
                    raise RuntimeError("CUDA-graph capture met a relation that the warm-up pass did not record")
                self._is_sorted = bool(mode.hints[mode.cursor])
                mode.cursor += 1
                mode.flags.append((self._is_sorted, flags))
            else:
                host = flags.cpu()            # one sync per batch and relation
                if int(host[1]):
                    raise IndexError("index out of range [0, %d) in ragged index tensor" % self.n_targets)
                self._is_sorted = not bool(int(host[0]))
                if mode is not None and mode.kind == "record":
                    mode.hints.append(self._is_sorted)
        self._rowptr, self._perm = F.raw_plan_build(self.ids, self.n_targets, self._is_sorted)
        self._find_long_rows()

    LONG_THRESH = 1024      # segments longer than this are reduced by a whole CTA (hub nodes)

    def max_degree(self):
        """Longest segment (host int).  One device read the first time; cached on the relation, and relations are cached
        per index tensor (get_plan), so a layer called again on the same graph does not synchronise."""
        if getattr(self, "_max_degree", None) is None:
            rp = self.rowptr
            self._max_degree = int((rp[1:] - rp[:-1]).max()) if self.n_targets > 0 and self.n_elements > 0 else 0
        return self._max_degree

    def _find_long_rows(self):
        """Hub rows would serialise one lane group for thousands of edges; list them once per batch."""
        self.long_rows = self.n_long = None
        self.long_cap = 0
        if self.n_targets <= 1024 or self.n_elements <= self.LONG_THRESH:
            return                      # few-row outputs already use the block-per-row mapping
        stats = F.raw_row_stats(self._rowptr, self.n_targets)
        mode = _MODE
        if mode is not None and mode.kind == "capture":
            if mode.cursor >= len(mode.hints):
                raise RuntimeError("CUDA-graph capture met a relation that the warm-up pass did not record")
            has_long = bool(mode.hints[mode.cursor])
            mode.cursor += 1
        else:
            has_long = int(stats.cpu()[0]) > self.LONG_THRESH
            if mode is not None and mode.kind == "record":
                mode.hints.append(has_long)
        if has_long:
            self.long_cap = self.n_elements // self.LONG_THRESH + 1
            self.long_rows, self.n_long = F.raw_long_rows(self._rowptr, self.n_targets, self.LONG_THRESH,
                                                          self.long_cap)

    @property
    def is_sorted(self):
        self._build()
        return self._is_sorted

    @property
    def rowptr(self):
        self._build()
        return self._rowptr

    @property
    def perm(self):
        """int32 [E] stable argsort of ids, or None when ids are already ascending."""
        self._build()
        return self._perm

    def permuted_ids(self, other):
        """other.ids in this relation's CSR order (so the fused kernel reads one index array)."""
        if self.perm is None:
            return other.ids
        key = id(other)
        if key not in self._permuted:
            self._permuted[key] = (F.raw_permute_i32(other.ids, self.perm), other)
        return self._permuted[key][0]

    def inv_count(self):
        """1/len(segment) (0 for empty) as float32 [n_targets, 1]: backward of segment_mean."""
        rp = self.rowptr
        cnt = (rp[1:] - rp[:-1]).to(torch.float32)
        return torch.where(cnt > 0, 1.0 / cnt, torch.zeros_like(cnt)).unsqueeze(1)


class IndexPlan:
    """Batch-indexed view of a ragged index tensor [E, K] that refers to rows of a ragged node tensor."""

    def __init__(self, index_values, index_partition, node_partition):
        if index_values.dim() != 2:
            raise ValueError("index tensor values must have shape [E, K]")
        if not index_values.is_cuda:
            raise KgcnnError("kgcnn_b200 needs CUDA tensors (no CPU fallback); got index tensor on %s"
                             % index_values.device)
        self.E, self.K = int(index_values.shape[0]), int(index_values.shape[1])
        self.N = int(node_partition.nvals)
        self.index_partition = index_partition
        self.node_partition = node_partition
        idx64 = index_values.to(torch.int64).contiguous()
        # column-major [K, E] int32, batch indexing (a1)
        self.idx32 = F.raw_shift_index(idx64, node_partition.row_splits, index_partition.row_splits)
        self._relations = {}

    def relation(self, col):
        col = int(col)
        if col < 0:
            col += self.K
        if col not in self._relations:
            self._relations[col] = Relation(self.idx32[col], self.N)
        return self._relations[col]


_MAX_PLANS_PER_TENSOR = 4   # distinct (node partition, index partition) wrappers kept per index tensor
_PLAN_CACHE = {}   # id(index values tensor) -> (weakref to it, {key: (IndexPlan, partitions...)})


def _cache_slot(vals):
    """Per-tensor dict keyed by object identity (tensors define == elementwise, so no WeakKeyDictionary)."""
    key = id(vals)
    slot = _PLAN_CACHE.get(key)
    if slot is not None and slot[0]() is vals:
        return slot[1]
    per_tensor = {}
    _PLAN_CACHE[key] = (weakref.ref(vals, lambda _, k=key: _PLAN_CACHE.pop(k, None)), per_tensor)
    return per_tensor


def get_plan(index_ragged, node_ragged):
    """Cached IndexPlan for (index tensor, node partition)."""
    vals = index_ragged.values
    node_part = node_ragged.partition
    key = (id(node_part), vals._version, id(index_ragged.partition))
    per_tensor = _cache_slot(vals)
    entry = per_tensor.get(key)
    if entry is None:
        for k in [k for k in per_tensor if k[1] != vals._version]:
            del per_tensor[k]                # an in-place update of the index tensor invalidates its plans
        while len(per_tensor) >= _MAX_PLANS_PER_TENSOR:
            del per_tensor[next(iter(per_tensor))]   # callers that rebuild their partitions per call: oldest out
        # keep the partitions alive next to the plan so id() keys cannot be recycled
        entry = (IndexPlan(vals, index_ragged.partition, node_part), node_part, index_ragged.partition)
        per_tensor[key] = entry
    return entry[0]


def unify_partitions(raggeds):
    """Make ragged tensors whose row_splits are equal share ONE RowPartition object, so that plans and
    int32 CSR views are built once per batch (nodes, coordinates and atomic numbers arrive as separate
    RaggedTensors with equal splits).  One device comparison + sync per distinct pair and batch."""
    canon = []
    out = []
    no_sync = _MODE is not None and _MODE.kind == "capture"
    for r in raggeds:
        if not hasattr(r, "partition"):
            out.append(r)
            continue
        hit = None
        for c in canon:
            if c is r.partition:
                hit = c
                break
            if not no_sync and c.nrows == r.partition.nrows and c.row_splits.device == r.partition.row_splits.device \
                    and int(c.row_splits.shape[0]) == int(r.partition.row_splits.shape[0]) \
                    and (c._nvals is None or r.partition._nvals is None or c._nvals == r.partition._nvals) \
                    and bool(torch.equal(c.row_splits, r.partition.row_splits)):
                hit = c
                break
        if hit is None:
            canon.append(r.partition)
            out.append(r)
        else:
            out.append(r if hit is r.partition else type(r)(r.values, hit))
    return out
