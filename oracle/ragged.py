"""Minimal ragged container for the oracle: values + int64 row_splits (torch CPU)."""
import numpy as np
import torch


class R:
    """ragged_rank==1 tensor: `values` [nvals, ...], `row_splits` int64 [B+1]."""

    def __init__(self, values, row_splits):
        self.values = values
        self.row_splits = torch.as_tensor(row_splits, dtype=torch.int64)

    @classmethod
    def from_row_lengths(cls, values, row_lengths):
        rl = torch.as_tensor(np.asarray(row_lengths), dtype=torch.int64)
        return cls(values, torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(rl, 0)]))

    def row_lengths(self):
        return self.row_splits[1:] - self.row_splits[:-1]

    def value_rowids(self):
        # kgcnn/ops/partition.py:41-45: repeat(range(B), row_lengths)
        return torch.repeat_interleave(torch.arange(len(self.row_splits) - 1), self.row_lengths())

    def with_values(self, v):
        return R(v, self.row_splits)

    def __getitem__(self, i):
        return self.values[int(self.row_splits[i]):int(self.row_splits[i + 1])]

    def nrows(self):
        return len(self.row_splits) - 1
