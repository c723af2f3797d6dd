"""Oracle restatement of kgcnn/ops/{partition,segment,scatter,activ}.py (torch CPU / NumPy).  Test infra only."""
import math

import numpy as np
import torch


# ---- kgcnn/ops/partition.py ---------------------------------------------------------------------------
def change_partition_by_name(in_partition, in_type, out_type):
    """kgcnn/ops/partition.py:5-93 (the conversions the hot path uses)."""
    p = torch.as_tensor(in_partition, dtype=torch.int64)
    lengths_names, splits_names = ["row_length", "row_lengths"], ["row_split", "row_splits"]
    if in_type == out_type:
        return p
    if in_type in lengths_names and out_type in splits_names:                       # :24-26
        return torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(p, 0)])
    if in_type in lengths_names and out_type == "value_rowids":                     # :28-29
        return torch.repeat_interleave(torch.arange(p.shape[0]), p)
    if in_type in splits_names and out_type in lengths_names:                       # :38-40
        return p[1:] - p[:-1]
    if in_type in splits_names and out_type == "value_rowids":                      # :42-45
        part_sum = p[1:] - p[:-1]
        return torch.repeat_interleave(torch.arange(part_sum.shape[0]), part_sum)
    if in_type == "value_rowids" and out_type in lengths_names:                     # :54-55
        return torch.bincount(p)
    if in_type == "value_rowids" and out_type in splits_names:                      # :57-60
        return torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(torch.bincount(p), 0)])
    raise TypeError("Unknown partition scheme, use: 'value_rowids', 'row_splits', 'row_lengths', etc.")


def partition_row_indexing(tensor_index, part_target, part_index, partition_type_target, partition_type_index,
                           from_indexing="sample", to_indexing="batch"):
    """kgcnn/ops/partition.py:97-161.  Integer, bit-exact."""
    tensor_index = torch.as_tensor(tensor_index, dtype=torch.int64)
    if to_indexing == from_indexing:                                                # :135-137
        return tensor_index
    nod_splits = change_partition_by_name(part_target, partition_type_target, "row_splits")   # :140
    edge_ids = change_partition_by_name(part_index, partition_type_index, "value_rowids")     # :143
    shift_index = nod_splits[edge_ids]                                              # :146
    for _ in range(1, tensor_index.dim()):                                          # :150-151
        shift_index = shift_index.unsqueeze(-1)
    if to_indexing == "batch" and from_indexing == "sample":                        # :154-155
        return tensor_index + shift_index
    if to_indexing == "sample" and from_indexing == "batch":
        return tensor_index - shift_index
    raise TypeError("ERROR:kgcnn: Unknown index change, use: 'sample', 'batch', ...")


# ---- kgcnn/ops/segment.py -----------------------------------------------------------------------------
def _nseg(segment_ids):
    return int(segment_ids[-1]) + 1 if segment_ids.numel() else 0


def segment_sum(data, segment_ids):
    """tf.math.segment_sum on SORTED ids: output has max_id+1 rows, gaps are 0."""
    out = torch.zeros((_nseg(segment_ids),) + tuple(data.shape[1:]), dtype=data.dtype)
    return out.index_add(0, segment_ids, data)


def segment_mean(data, segment_ids):
    n = _nseg(segment_ids)
    cnt = torch.bincount(segment_ids, minlength=n).to(data.dtype).clamp_min(1)
    return segment_sum(data, segment_ids) / cnt.reshape((-1,) + (1,) * (data.dim() - 1))


def _segment_extreme(data, segment_ids, reduce):
    n = _nseg(segment_ids)
    out = torch.zeros((n,) + tuple(data.shape[1:]), dtype=data.dtype)
    idx = segment_ids.reshape((-1,) + (1,) * (data.dim() - 1)).expand_as(data)
    # include_self=False: untouched (gap) rows keep 0, like TF; gradient is split evenly among ties, like
    # TF's _SegmentMinOrMaxGrad.
    return out.scatter_reduce(0, idx, data, reduce=reduce, include_self=False)


def segment_softmax(data, segment_ids, normalize=True):
    """kgcnn/ops/segment.py:4-24 on SORTED ids: shift by the segment max (:16-19), exp, divide by the segment sum."""
    if normalize:
        data = data - segment_ops_by_name("max", data, segment_ids)[segment_ids]
    data_exp = torch.exp(data)
    return data_exp / segment_sum(data_exp, segment_ids)[segment_ids]


def segment_ops_by_name(segment_name, data, segment_ids):
    """kgcnn/ops/segment.py:27-52."""
    if segment_name in ["segment_mean", "mean", "reduce_mean"]:
        return segment_mean(data, segment_ids)
    if segment_name in ["segment_sum", "sum", "reduce_sum"]:
        return segment_sum(data, segment_ids)
    if segment_name in ["segment_max", "max", "reduce_max"]:
        return _segment_extreme(data, segment_ids, "amax")
    if segment_name in ["segment_min", "min", "reduce_min"]:
        return _segment_extreme(data, segment_ids, "amin")
    raise TypeError("Unknown segment operation, choose: 'segment_mean', 'segment_sum', ...")


def segment_sum_sequential_f32(data, segment_ids):
    """float32 NumPy segment sum in strict edge order (what TF's CPU SegmentSum does): the 'TF-CPU-like'
    comparand used to report the reference's own rounding error beside ours."""
    data = np.asarray(data, dtype=np.float32)
    ids = np.asarray(segment_ids)
    n = int(ids[-1]) + 1 if len(ids) else 0
    out = np.zeros((n,) + data.shape[1:], dtype=np.float32)
    for e in range(len(ids)):
        out[ids[e]] += data[e]
    return out


# ---- kgcnn/ops/scatter.py -----------------------------------------------------------------------------
def tensor_scatter_nd_ops_by_name(segment_name, tensor, indices, updates):
    """kgcnn/ops/scatter.py:5-26 for 2-D indices (receiver, relation) into a [N, R, m] tensor."""
    n, r = tensor.shape[0], tensor.shape[1]
    flat = indices[:, 0] * r + indices[:, 1]
    t2 = tensor.reshape((n * r,) + tuple(tensor.shape[2:]))
    if segment_name in ["segment_sum", "sum", "reduce_sum", "add"]:
        out = t2.index_add(0, flat, updates)
    elif segment_name in ["segment_max", "max", "reduce_max"]:
        out = t2.scatter_reduce(0, flat.unsqueeze(-1).expand_as(updates), updates, "amax", include_self=True)
    elif segment_name in ["segment_min", "min", "reduce_min"]:
        out = t2.scatter_reduce(0, flat.unsqueeze(-1).expand_as(updates), updates, "amin", include_self=True)
    else:
        raise TypeError("Unknown pooling, choose: 'mean', 'sum', ...")
    return out.reshape(tensor.shape)


# ---- kgcnn/ops/activ.py + keras activations ----------------------------------------------------------------
def softplus(x):
    return torch.logaddexp(x, torch.zeros_like(x))


def shifted_softplus(x):
    """kgcnn/ops/activ.py:5-15: softplus(x) - log(2)."""
    return softplus(x) - math.log(2.0)


def softplus2(x):
    """kgcnn/ops/activ.py:18-29."""
    return torch.relu(x) + torch.log(0.5 * torch.exp(-torch.abs(x)) + 0.5)


def swish(x, beta=1.0):
    """kgcnn/ops/activ.py:83-107 (and keras 'swish')."""
    return x * torch.sigmoid(beta * x)


def leaky_softplus(x, alpha=0.05):
    return softplus(x) * (1 - alpha) + alpha * x


ACTIVATIONS = {
    None: lambda x: x, "linear": lambda x: x,
    "kgcnn>shifted_softplus": shifted_softplus, "shifted_softplus": shifted_softplus,
    "kgcnn>softplus2": softplus2, "softplus2": softplus2,
    "swish": swish, "kgcnn>swish": swish,
    "tanh": torch.tanh, "relu": torch.relu, "sigmoid": torch.sigmoid, "softplus": softplus,
    "kgcnn>leaky_softplus": leaky_softplus,
    "kgcnn>leaky_relu": lambda x: torch.nn.functional.leaky_relu(x, 0.05),
    "leaky_relu": lambda x: torch.nn.functional.leaky_relu(x, 0.05),
}


def activation(name):
    if callable(name):
        return name
    return ACTIVATIONS[name]


def divide_no_nan(a, b):
    """tf.math.divide_no_nan: 0 where b == 0."""
    safe = torch.where(b == 0, torch.ones_like(b), b)
    return torch.where(b == 0, torch.zeros_like(a * b), a / safe)
