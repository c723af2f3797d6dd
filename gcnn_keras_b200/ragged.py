"""RaggedTensor(values, row_splits) value type with the members the kgcnn layers touch.

The reference passes ``tf.RaggedTensor`` of ``ragged_rank == 1`` between layers and only ever uses
``values``, ``row_splits``, ``row_lengths()``, ``value_rowids()``, ``from_row_splits/lengths``,
``with_values`` and ``to_tensor`` (kgcnn/layers/gather.py:76-77,90; kgcnn/layers/pooling.py:53-55,78,215;
kgcnn/layers/casting.py:80-83; kgcnn/layers/base.py:101-103,153-164).  This class keeps exactly that
contract on top of torch tensors; the row partition is shared between tensors derived from one another so
that int32 CSR views are computed once per batch instead of once per layer call.
"""
import numpy as np
import torch


class RowPartition:
    """row_splits (int64 [B+1]) plus lazily derived views, shared by all tensors of one ragged axis."""

    def __init__(self, row_splits, nvals=None, max_len=None):
        self.max_len = max_len          # optional host-side hint: longest row (avoids a sync in to_tensor)
        if row_splits.dtype != torch.int64:
            row_splits = row_splits.to(torch.int64)
        self.row_splits = row_splits.contiguous()
        self.nrows = int(self.row_splits.shape[0]) - 1
        self._nvals = nvals
        self._row_lengths = None
        self._rowids64 = None
        self._rowids32 = None
        self._splits32 = None
        self._relation = None

    @property
    def device(self):
        return self.row_splits.device

    @property
    def nvals(self):
        if self._nvals is None:
            self._nvals = int(self.row_splits[-1].item())
        return self._nvals

    def row_lengths(self):
        if self._row_lengths is None:
            self._row_lengths = self.row_splits[1:] - self.row_splits[:-1]
        return self._row_lengths

    def splits32(self):
        if self._splits32 is None:
            if self.row_splits.is_cuda:
                from . import functional as F
                self._splits32 = F.raw_cast_i32(self.row_splits)
            else:
                self._splits32 = self.row_splits.to(torch.int32)
        return self._splits32

    def value_rowids32(self):
        if self._rowids32 is None:
            if self.row_splits.is_cuda:
                from . import functional as F
                self._rowids32 = F.raw_value_rowids(self.row_splits, self.nvals)
            else:
                self._rowids32 = torch.repeat_interleave(
                    torch.arange(self.nrows, dtype=torch.int32), self.row_lengths())
        return self._rowids32

    def value_rowids(self):
        if self._rowids64 is None:
            self._rowids64 = self.value_rowids32().to(torch.int64)
        return self._rowids64

    def relation(self):
        """Relation 'value -> row' (sorted by construction): used by PoolingNodes."""
        if self._relation is None:
            from .plan import Relation
            self._relation = Relation(ids=self.value_rowids32(), n_targets=self.nrows,
                                      rowptr=self.splits32(), perm=None, is_sorted=True)
        return self._relation


class RaggedTensor:
    """Ragged tensor of ragged_rank 1: ``values`` [nvals, ...] partitioned into rows by ``row_splits``."""

    ragged_rank = 1

    def __init__(self, values, partition):
        self.values = values
        self.partition = partition
        # how these values were derived, when a layer can use it to fuse: ("gathered", source ragged, relation) or
        # ("distance", coordinates ragged, rel_i, rel_j, eps).  Not propagated by with_values().
        self.provenance = None

    # ---- constructors (tf.RaggedTensor.from_*) -------------------------------------------------
    @classmethod
    def from_row_splits(cls, values, row_splits, validate=False):
        if isinstance(row_splits, RowPartition):
            part = row_splits
        else:
            part = RowPartition(torch.as_tensor(row_splits), nvals=int(values.shape[0]))
        if validate:
            rs = part.row_splits.cpu()
            assert int(rs[0]) == 0 and int(rs[-1]) == values.shape[0] and bool((rs[1:] >= rs[:-1]).all()), \
                "invalid row_splits"
        return cls(values, part)

    @classmethod
    def from_row_lengths(cls, values, row_lengths, validate=False):
        row_lengths = torch.as_tensor(row_lengths).to(torch.int64)
        splits = torch.zeros(row_lengths.shape[0] + 1, dtype=torch.int64, device=row_lengths.device)
        torch.cumsum(row_lengths, 0, out=splits[1:])
        return cls.from_row_splits(values, splits, validate=validate)

    @classmethod
    def from_value_rowids(cls, values, value_rowids, nrows=None, validate=False):
        value_rowids = torch.as_tensor(value_rowids).to(torch.int64)
        if nrows is None:
            nrows = int(value_rowids.max().item()) + 1 if value_rowids.numel() else 0
        lengths = torch.bincount(value_rowids, minlength=nrows)
        return cls.from_row_lengths(values, lengths, validate=validate)

    @classmethod
    def from_nested(cls, rows, dtype=None, device=None):
        """Build from a python list of per-graph arrays (kgcnn/data/utils.py:129-157)."""
        arrs = [np.asarray(r) for r in rows]
        lengths = np.array([len(a) for a in arrs], dtype=np.int64)
        inner = next((a.shape[1:] for a in arrs if len(a)), ())
        flat = np.concatenate([a.reshape((-1,) + tuple(inner)) for a in arrs], axis=0) if arrs else np.zeros((0,))
        values = torch.as_tensor(flat, dtype=dtype)
        out = cls.from_row_lengths(values, torch.as_tensor(lengths))
        return out.to(device) if device is not None else out

    # ---- tf.RaggedTensor members ------------------------------------------------------------------
    @property
    def flat_values(self):
        return self.values

    @property
    def row_splits(self):
        return self.partition.row_splits

    def row_lengths(self):
        return self.partition.row_lengths()

    def value_rowids(self):
        return self.partition.value_rowids()

    def nrows(self):
        return self.partition.nrows

    @property
    def shape(self):
        return (self.partition.nrows, None) + tuple(self.values.shape[1:])

    @property
    def dtype(self):
        return self.values.dtype

    @property
    def device(self):
        return self.values.device

    def with_values(self, new_values):
        return RaggedTensor(new_values, self.partition)

    with_flat_values = with_values

    def to(self, device):
        if self.values.device == torch.device(device) and self.partition.device == torch.device(device):
            return self
        return RaggedTensor(self.values.to(device), RowPartition(self.partition.row_splits.to(device),
                                                                 nvals=int(self.values.shape[0])))

    def _max_len(self):
        if self.partition.max_len is not None:
            return int(self.partition.max_len)
        lengths = self.row_lengths()
        return int(lengths.max().item()) if lengths.numel() else 0     # one sync; pass max_len to avoid it

    def to_tensor(self, default_value=0, shape=None):
        """Pad to [B, max_len, ...] (casting.py:78-80): kgcnn_pad on the device.  Host-resident ragged tensors
        (data containers built by from_numpy_list) are padded with tensor indexing; no kernel math is involved."""
        nmax = self._max_len() if shape is None or shape[1] is None else int(shape[1])
        default_value = 0 if default_value is None else default_value
        if self.values.is_cuda:
            from . import functional as F
            return F.pad(self.values, self.row_splits, nmax, default_value)
        out = self.values.new_full((self.partition.nrows, nmax) + tuple(self.values.shape[1:]), default_value)
        rowids = self.partition.value_rowids()
        pos = torch.arange(self.values.shape[0], device=self.values.device) - self.row_splits[rowids]
        keep = pos < nmax
        out[rowids[keep], pos[keep]] = self.values[keep]
        return out

    def to_tensor_and_mask(self, default_value=0, boolean_mask=False, shape=None):
        """(padded, mask) of ChangeTensorType 'mask' (casting.py:80-84), both from ONE launch on the device."""
        nmax = self._max_len() if shape is None or shape[1] is None else int(shape[1])
        default_value = 0 if default_value is None else default_value
        if self.values.is_cuda and not (self.values.requires_grad and torch.is_grad_enabled()):
            from . import functional as F
            return F.raw_pad(self.values, self.row_splits, nmax, default_value, "bool" if boolean_mask else "like")
        padded = self.to_tensor(default_value, shape=shape)
        ones = torch.ones_like(self.values, dtype=torch.bool if boolean_mask else self.values.dtype)
        return padded, self.with_values(ones).to_tensor(shape=shape)

    @classmethod
    def from_padded(cls, padded, partition):
        """Inverse of to_tensor for a known row partition (force.py:204-211): kgcnn_unpad on the device."""
        if padded.is_cuda:
            from . import functional as F
            return cls(F.unpad(padded, partition.row_splits, partition.nvals), partition)
        rowids = partition.value_rowids()
        pos = torch.arange(partition.nvals, device=padded.device) - partition.row_splits[rowids]
        return cls(padded[rowids, pos], partition)

    def __getitem__(self, i):
        a, b = int(self.row_splits[i]), int(self.row_splits[i + 1])
        return self.values[a:b]

    def __len__(self):
        return self.partition.nrows

    def __repr__(self):
        return "RaggedTensor(shape=%s, dtype=%s, device=%s)" % (self.shape, self.dtype, self.device)
