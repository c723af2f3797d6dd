"""PyTorch host binding of the C-ABI: raw launchers + autograd Functions.

Every differentiable primitive's backward is written with the OTHER primitives (gather <-> segment-sum are
adjoint; the fused gather*weight->segment-sum has itself on the transposed relation and the two-sided
gather-product as VJPs), so the set is closed under differentiation to any order.  That is what
``EnergyForceModel`` needs: forces are dE/dx computed inside the step and are then trained on
(kgcnn/model/force.py:159-176).

torch is plumbing here (device memory, streams, the autograd tape); all arithmetic on ragged data runs in
libkgcnn_b200.so.  There is no CPU path: CPU tensors raise.
"""
import torch

from . import _lib
from ._lib import KgcnnError, OP_MAX, OP_MEAN, OP_MIN, OP_SUM, check

_OPS = {
    "segment_sum": OP_SUM, "sum": OP_SUM, "reduce_sum": OP_SUM,
    "segment_mean": OP_MEAN, "mean": OP_MEAN, "reduce_mean": OP_MEAN,
    "segment_max": OP_MAX, "max": OP_MAX, "reduce_max": OP_MAX,
    "segment_min": OP_MIN, "min": OP_MIN, "reduce_min": OP_MIN,
}


def op_code(name):
    """segment_ops_by_name dispatch (kgcnn/ops/segment.py:39-51): unknown names raise TypeError."""
    try:
        return _OPS[name]
    except KeyError:
        raise TypeError("Unknown segment operation, choose: 'segment_mean', 'segment_sum', ...")


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise KgcnnError("kgcnn_b200 kernels need CUDA tensors (there is no CPU fallback); got %s" % t.device)


def _f32c(t):
    if t.dtype != torch.float32:
        raise TypeError("kgcnn_b200 kernels compute in float32, got %s" % t.dtype)
    return t.contiguous()


# ------------------------------------------------------------------------------------------------
# raw launchers (no autograd)
# ------------------------------------------------------------------------------------------------
def raw_shift_index(idx64, node_splits, edge_splits, want_i64=False):
    """a1: batch-indexed indices.  Returns int32 [K, E] (column-major) and optionally int64 [E, K]."""
    _need_cuda(idx64, node_splits, edge_splits)
    E, K = int(idx64.shape[0]), int(idx64.shape[1])
    B = int(edge_splits.shape[0]) - 1
    if int(node_splits.shape[0]) - 1 != B:
        raise ValueError("index tensor and node tensor have different batch sizes: %d vs %d"
                         % (B, int(node_splits.shape[0]) - 1))
    out32 = torch.empty((K, E), dtype=torch.int32, device=idx64.device)
    out64 = torch.empty((E, K), dtype=torch.int64, device=idx64.device) if want_i64 else None
    check(_lib.lib().kgcnn_shift_index(_ptr(idx64), E, K, _ptr(node_splits), _ptr(edge_splits), B, _ptr(out64),
                                       _ptr(out32), _stream()), "kgcnn_shift_index")
    return (out32, out64) if want_i64 else out32


def raw_value_rowids(row_splits, nvals):
    _need_cuda(row_splits)
    ids = torch.empty((nvals,), dtype=torch.int32, device=row_splits.device)
    check(_lib.lib().kgcnn_value_rowids(_ptr(row_splits), int(row_splits.shape[0]) - 1, nvals, _ptr(ids), _stream()),
          "kgcnn_value_rowids")
    return ids


def raw_cast_i32(t64):
    _need_cuda(t64)
    out = torch.empty(t64.shape, dtype=torch.int32, device=t64.device)
    check(_lib.lib().kgcnn_cast_i64_i32(_ptr(t64), t64.numel(), _ptr(out), _stream()), "kgcnn_cast_i64_i32")
    return out


def _elem_words(dtype):
    if dtype in (torch.float32, torch.int32):
        return 1
    if dtype == torch.int64:
        return 2
    raise TypeError("ragged <-> padded conversion moves float32 / int32 / int64 values, got %s" % dtype)


def raw_pad(values, row_splits, max_len, default_value=0, mask=None):
    """a20: (padded [B, max_len, ...], mask | None).  mask: None, 'like' (ones_like of the values, same dtype and
    shape as padded) or 'bool' (one flag per element).  One launch for both outputs."""
    import struct
    _need_cuda(values, row_splits)
    words = _elem_words(values.dtype)
    values = values.contiguous()
    B = int(row_splits.shape[0]) - 1
    tail = tuple(values.shape[1:])
    n_tail = 1
    for t in tail:
        n_tail *= int(t)
    out = torch.empty((B, int(max_len)) + tail, dtype=values.dtype, device=values.device)
    m = None
    if mask == "like":
        m = torch.empty_like(out)
    elif mask == "bool":
        if words != 1:
            raise TypeError("boolean masks are defined for 4-byte element types")
        m = torch.empty(out.shape, dtype=torch.bool, device=values.device)
    elif mask is not None:
        raise ValueError("mask must be None, 'like' or 'bool'")
    if values.dtype == torch.float32:
        w0, w1 = struct.unpack("<I", struct.pack("<f", float(default_value)))[0], 0
    else:
        bits = int(default_value) & 0xFFFFFFFFFFFFFFFF
        w0, w1 = bits & 0xFFFFFFFF, bits >> 32
    if n_tail == 0:
        return out, m
    check(_lib.lib().kgcnn_pad(_ptr(values), _ptr(row_splits), B, int(max_len), n_tail * words, w0, w1, words,
                               _ptr(out), _ptr(m), 1 if mask == "bool" else 0, _stream()), "kgcnn_pad")
    return out, m


def raw_unpad(padded, row_splits, n_values):
    """values[n] = padded[graph_of(n), n - row_splits[graph_of(n)]] (force.py:204-211 without the mask tensor:
    the row partition already says which slots are live)."""
    _need_cuda(padded, row_splits)
    words = _elem_words(padded.dtype)
    padded = padded.contiguous()
    B, max_len = int(padded.shape[0]), int(padded.shape[1])
    tail = tuple(padded.shape[2:])
    n_tail = 1
    for t in tail:
        n_tail *= int(t)
    out = torch.empty((int(n_values),) + tail, dtype=padded.dtype, device=padded.device)
    if n_tail and n_values:
        check(_lib.lib().kgcnn_unpad(_ptr(padded), _ptr(row_splits), B, max_len, n_tail * words, int(n_values),
                                     _ptr(out), _stream()), "kgcnn_unpad")
    return out


class _Pad(torch.autograd.Function):
    """ragged values -> padded tensor; the VJP is the unpad kernel (padding slots carry no gradient)."""

    @staticmethod
    def forward(ctx, values, row_splits, max_len, default_value):
        ctx.row_splits, ctx.n = row_splits, int(values.shape[0])
        return raw_pad(values, row_splits, max_len, default_value)[0]

    @staticmethod
    def backward(ctx, g):
        return _Unpad.apply(g.contiguous(), ctx.row_splits, ctx.n), None, None, None


class _Unpad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, padded, row_splits, n_values):
        ctx.row_splits, ctx.max_len = row_splits, int(padded.shape[1])
        return raw_unpad(padded, row_splits, n_values)

    @staticmethod
    def backward(ctx, g):
        return _Pad.apply(g.contiguous(), ctx.row_splits, ctx.max_len, 0), None, None


def pad(values, row_splits, max_len, default_value=0):
    """RaggedTensor.to_tensor on the device (differentiable for floating values)."""
    if values.is_floating_point() and values.requires_grad and torch.is_grad_enabled():
        return _Pad.apply(values, row_splits, int(max_len), default_value)
    return raw_pad(values, row_splits, max_len, default_value)[0]


def unpad(padded, row_splits, n_values):
    if padded.is_floating_point() and padded.requires_grad and torch.is_grad_enabled():
        return _Unpad.apply(padded, row_splits, int(n_values))
    return raw_unpad(padded, row_splits, n_values)


def raw_check_sorted(key, n_targets, flags=None, clamp=False):
    """flags int32[2] = (not ascending, out of range).  `flags`: a zeroed device slot to write into (the capture
    mode's guard buffer); `clamp`: also rewrite out-of-range ids in place (see include/kgcnn_b200.h)."""
    _need_cuda(key)
    if flags is None:
        flags = torch.zeros((2,), dtype=torch.int32, device=key.device)
    check(_lib.lib().kgcnn_check_sorted(_ptr(key), 1, int(key.shape[0]), n_targets, _ptr(flags),
                                        1 if (clamp and n_targets > 0) else 0, _stream()), "kgcnn_check_sorted")
    return flags


def raw_plan_build(key, n_targets, assume_sorted):
    """CSR row pointer [N+1] and stable permutation (None when sorted)."""
    _need_cuda(key)
    E = int(key.shape[0])
    rowptr = torch.empty((n_targets + 1,), dtype=torch.int32, device=key.device)
    if assume_sorted:
        check(_lib.lib().kgcnn_plan_build(_ptr(key), 1, E, n_targets, 1, _ptr(rowptr), None, None, 0, _stream()),
              "kgcnn_plan_build")
        return rowptr, None
    perm = torch.empty((E,), dtype=torch.int32, device=key.device)
    nbytes = int(_lib.lib().kgcnn_plan_workspace_bytes(E, n_targets))
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=key.device)
    check(_lib.lib().kgcnn_plan_build(_ptr(key), 1, E, n_targets, 0, _ptr(rowptr), _ptr(perm), _ptr(ws), nbytes,
                                      _stream()), "kgcnn_plan_build")
    return rowptr, perm


def raw_row_stats(rowptr, n_rows):
    stats = torch.zeros((1,), dtype=torch.int32, device=rowptr.device)
    check(_lib.lib().kgcnn_row_stats(_ptr(rowptr), n_rows, _ptr(stats), _stream()), "kgcnn_row_stats")
    return stats


def raw_long_rows(rowptr, n_rows, thresh, cap):
    rows = torch.zeros((cap,), dtype=torch.int32, device=rowptr.device)
    count = torch.zeros((1,), dtype=torch.int32, device=rowptr.device)
    check(_lib.lib().kgcnn_long_rows_build(_ptr(rowptr), n_rows, thresh, cap, _ptr(rows), _ptr(count), _stream()),
          "kgcnn_long_rows_build")
    return rows, count


def raw_permute_i32(src, perm):
    out = torch.empty_like(perm)
    check(_lib.lib().kgcnn_permute_i32(_ptr(src), 1, _ptr(perm), int(perm.shape[0]), _ptr(out), _stream()),
          "kgcnn_permute_i32")
    return out


def raw_gather_rows(x2d, ids, out=None, col_offset=0):
    """y[e] = x[ids[e]]; with `out` [E, W] writes into out[:, col_offset:col_offset+F]."""
    _need_cuda(x2d, ids)
    N, Fdim = int(x2d.shape[0]), int(x2d.shape[1])
    E = int(ids.shape[0])
    if x2d.dtype == torch.int64:
        y = torch.empty((E, Fdim), dtype=torch.int64, device=x2d.device) if out is None else out
        ystride = int(y.shape[1])
        yp = y.data_ptr() + col_offset * 8
        check(_lib.lib().kgcnn_gather_rows_i64(_ptr(x2d.contiguous()), N, Fdim, _ptr(ids), 1, E, yp, ystride,
                                               _stream()), "kgcnn_gather_rows_i64")
        return y
    x2d = _f32c(x2d)
    y = torch.empty((E, Fdim), dtype=torch.float32, device=x2d.device) if out is None else out
    ystride = int(y.shape[1])
    yp = y.data_ptr() + col_offset * 4
    check(_lib.lib().kgcnn_gather_rows(_ptr(x2d), N, Fdim, _ptr(ids), 1, E, yp, ystride, _stream()),
          "kgcnn_gather_rows")
    return y


def raw_gms(x2d, w, rel_out, rel_src, op=OP_SUM):
    """y[i] = OP_{e in seg_out(i)} x[src(e)] * w[e]; rel_src None -> x is the element tensor itself."""
    _need_cuda(x2d, w)
    x2d = _f32c(x2d)
    Fdim = int(x2d.shape[1])
    wcols = 0
    if w is not None:
        w = _f32c(w)
        wcols = int(w.shape[1]) if w.dim() == 2 else 1
    n_out = rel_out.n_targets
    if rel_out.n_elements == 0 or x2d.shape[0] == 0:      # every segment is empty -> 0 (no launch needed)
        return torch.zeros((n_out, Fdim), dtype=torch.float32, device=x2d.device)
    y = torch.empty((n_out, Fdim), dtype=torch.float32, device=x2d.device)
    perm = rel_out.perm
    if rel_src is not None:
        gidx = rel_out.permuted_ids(rel_src)
        gperm = 1 if perm is not None else 0
    else:
        gidx, gperm = None, 0
    check(_lib.lib().kgcnn_gather_mul_segment_reduce(_ptr(x2d), Fdim, _ptr(w), wcols, _ptr(rel_out.rowptr),
                                                     _ptr(perm), _ptr(gidx), 1, gperm, n_out, rel_out.n_elements, op,
                                                     _ptr(y), _ptr(rel_out.long_rows), _ptr(rel_out.n_long),
                                                     rel_out.long_cap, rel_out.LONG_THRESH if rel_out.long_cap else 0,
                                                     _stream()), "kgcnn_gather_mul_segment_reduce")
    return y


def raw_g2m(a2d, ids_a, b2d, ids_b, n_elements, reduce):
    _need_cuda(a2d, b2d)
    a2d, b2d = _f32c(a2d), _f32c(b2d)
    Fdim = int(a2d.shape[1])
    z = torch.empty((n_elements, 1 if reduce else Fdim), dtype=torch.float32, device=a2d.device)
    check(_lib.lib().kgcnn_gather2_mul(_ptr(a2d), _ptr(ids_a), 1, _ptr(b2d), _ptr(ids_b), 1, Fdim, n_elements,
                                       1 if reduce else 0, _ptr(z), _stream()), "kgcnn_gather2_mul")
    return z


# ------------------------------------------------------------------------------------------------
# autograd Functions
# ------------------------------------------------------------------------------------------------
def _flat2d(t):
    return t.reshape(t.shape[0], -1) if t.dim() != 2 else t


class _Gather(torch.autograd.Function):
    """y[e] = x[rel.ids[e]]   (kgcnn/layers/gather.py:228); backward = segment-sum over rel."""

    @staticmethod
    def forward(ctx, x, rel):
        ctx.rel = rel
        ctx.tail = tuple(x.shape[1:])
        y = raw_gather_rows(_flat2d(x), rel.ids)
        return y.reshape((rel.n_elements,) + ctx.tail)

    @staticmethod
    def backward(ctx, g):
        return _SegmentSum.apply(g.contiguous(), ctx.rel), None


class _SegmentSum(torch.autograd.Function):
    """y[i] = sum_{e: rel.ids[e]==i} v[e]   (sorted segment_sum + scatter_nd pad, pooling.py:63-76)."""

    @staticmethod
    def forward(ctx, v, rel):
        ctx.rel = rel
        tail = tuple(v.shape[1:])
        y = raw_gms(_flat2d(v), None, rel, None, OP_SUM)
        return y.reshape((rel.n_targets,) + tail)

    @staticmethod
    def backward(ctx, g):
        return _Gather.apply(g.contiguous(), ctx.rel), None


class _GMS(torch.autograd.Function):
    """y[i] = sum_{e: out.ids[e]==i} x[src.ids[e]] * w[e]; w is [E,F] or [E,1]."""

    @staticmethod
    def forward(ctx, x, w, rel_out, rel_src):
        ctx.rel_out, ctx.rel_src = rel_out, rel_src
        ctx.save_for_backward(x, w)
        return raw_gms(x, w, rel_out, rel_src, OP_SUM)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        g = g.contiguous()
        gx = gw = None
        if ctx.needs_input_grad[0]:
            gx = _GMS.apply(g, w, ctx.rel_src, ctx.rel_out)
        if ctx.needs_input_grad[1]:
            reduce = w.shape[1] == 1 and x.shape[1] != 1
            gw = _G2M.apply(x, g, ctx.rel_src, ctx.rel_out, reduce)
        return gx, gw, None, None


class _G2M(torch.autograd.Function):
    """z[e] = a[ra.ids[e]] * b[rb.ids[e]] (optionally summed over features -> [E,1]); rb None = 'row e'."""

    @staticmethod
    def forward(ctx, a, b, rel_a, rel_b, reduce):
        ctx.rel_a, ctx.rel_b, ctx.reduce = rel_a, rel_b, reduce
        ctx.save_for_backward(a, b)
        return raw_g2m(a, rel_a.ids, b, None if rel_b is None else rel_b.ids, rel_a.n_elements, reduce)

    @staticmethod
    def backward(ctx, gz):
        a, b = ctx.saved_tensors
        gz = gz.contiguous()
        ga = gb = None
        if ctx.needs_input_grad[0]:
            ga = _GMS.apply(b, gz, ctx.rel_a, ctx.rel_b) if ctx.rel_b is not None \
                else _GMS_edge.apply(b, gz, ctx.rel_a)
        if ctx.needs_input_grad[1]:
            if ctx.rel_b is not None:
                gb = _GMS.apply(a, gz, ctx.rel_b, ctx.rel_a)
            elif gz.shape[1] == a.shape[1]:
                gb = _G2M.apply(a, gz, ctx.rel_a, None, False)
            else:
                gb = _Gather.apply(a, ctx.rel_a) * gz
        return ga, gb, None, None, None


class _GMS_edge(torch.autograd.Function):
    """y[i] = sum_{e: rel.ids[e]==i} v[e] * w[e]  (PoolingWeightedLocalEdges, pooling.py:152-162)."""

    @staticmethod
    def forward(ctx, v, w, rel):
        ctx.rel = rel
        ctx.save_for_backward(v, w)
        return raw_gms(v, w, rel, None, OP_SUM)

    @staticmethod
    def backward(ctx, g):
        v, w = ctx.saved_tensors
        g = g.contiguous()
        gv = gw = None
        if ctx.needs_input_grad[0]:
            gv = _G2M.apply(g, w, ctx.rel, None, False) if w.shape[1] == v.shape[1] \
                else _Gather.apply(g, ctx.rel) * w
        if ctx.needs_input_grad[1]:
            gw = _G2M.apply(g, v, ctx.rel, None, w.shape[1] == 1 and v.shape[1] != 1)
        return gv, gw, None


class _GatherConcat(torch.autograd.Function):
    """y[e] = concat_k x[rel_k.ids[e]]  (kgcnn/layers/gather.py:83-90: tf.gather [E,K,F] + K slices + tf.concat):
    K launches of the strided-output row gather into ONE [E, K F] buffer, no concat pass.  Backward: the segment sum
    of each column block along its relation."""

    @staticmethod
    def forward(ctx, x, *rels):
        ctx.rels = rels
        x2 = _f32c(_flat2d(x))
        ctx.tail = tuple(x.shape[1:])
        Fd = int(x2.shape[1])
        out = torch.empty((rels[0].n_elements, Fd * len(rels)), dtype=torch.float32, device=x2.device)
        for k, rel in enumerate(rels):
            raw_gather_rows(x2, rel.ids, out=out, col_offset=k * Fd)
        return out

    @staticmethod
    def backward(ctx, g):
        Fd = g.shape[1] // len(ctx.rels)
        total = None
        for k, rel in enumerate(ctx.rels):
            part = _SegmentSum.apply(g[:, k * Fd:(k + 1) * Fd].contiguous(), rel)
            total = part if total is None else total + part
        return (total.reshape((total.shape[0],) + ctx.tail),) + (None,) * len(ctx.rels)


def gather_concat(x, rels):
    """Concatenated row gathers along several relations of the same element set (GatherEmbedding, concat_axis=2)."""
    if len(rels) == 1:
        return gather(x, rels[0])
    if not torch.is_floating_point(x) or x.dim() != 2:
        return torch.cat([gather(x, r) for r in rels], dim=1)
    return _GatherConcat.apply(x, *rels)


def gather(x, rel):
    """Row gather along a relation (any trailing feature shape)."""
    if x.dtype == torch.int64 or not torch.is_floating_point(x):
        y = raw_gather_rows(_flat2d(x.to(torch.int64)), rel.ids)
        return y.reshape((rel.n_elements,) + tuple(x.shape[1:]))
    return _Gather.apply(x, rel)


def segment_reduce(v, rel, op="sum"):
    """Deterministic CSR segment reduce; empty segments give 0 for every op (F4 in SURVEY.md)."""
    code = op if isinstance(op, int) else op_code(op)
    if code == OP_SUM:
        return _SegmentSum.apply(v, rel)
    if code == OP_MEAN:
        s = _SegmentSum.apply(v, rel)
        inv = rel.inv_count()
        return s * inv.reshape((-1,) + (1,) * (v.dim() - 1))
    # max / min: value from the kernel (no tape), gradient split equally among ties like TF's
    # SegmentMax/SegmentMin gradient.
    tail = tuple(v.shape[1:])
    with torch.no_grad():
        y = raw_gms(_flat2d(v), None, rel, None, code).reshape((rel.n_targets,) + tail)
    if not (torch.is_grad_enabled() and v.requires_grad):
        return y
    yg = gather(y, rel)
    mask = (v.detach() == yg).to(v.dtype)
    nties = _SegmentSum.apply(mask, rel).clamp_min(1.0)
    weights = mask / gather(nties, rel)
    # value path carries no gradient; gradient path: y = segment_sum(v * weights) has dy/dv = weights
    return y + (_SegmentSum.apply(v * weights, rel) - _SegmentSum.apply(v.detach() * weights, rel))


class _SegmentSoftmax(torch.autograd.Function):
    """alpha = segment_softmax(a) over rel (kgcnn/ops/segment.py:4-24), one fused launch.  The backward
    da = alpha * (g - gather(segment_sum(alpha * g))) is written with the closed primitives on the SAVED OUTPUT,
    so it can be differentiated again (attention layers inside a force model)."""

    @staticmethod
    def forward(ctx, a, rel, normalize):
        _need_cuda(a)
        a2 = _f32c(_flat2d(a))
        out = torch.zeros_like(a2)
        perm = rel.perm
        check(_lib.lib().kgcnn_segment_softmax(_ptr(a2), _ptr(rel.rowptr), _ptr(perm), rel.n_targets, rel.n_elements,
                                               int(a2.shape[1]), int(bool(normalize)), _ptr(out), _stream()),
              "kgcnn_segment_softmax")
        out = out.reshape(a.shape)
        ctx.rel = rel
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, g):
        (alpha,) = ctx.saved_tensors
        ag = alpha * g
        t = _SegmentSum.apply(ag.contiguous(), ctx.rel)
        return ag - alpha * _Gather.apply(t, ctx.rel), None, None


def segment_softmax(a, rel, normalize=True):
    """Softmax of a[E, ...] within the segments of `rel`, returned in the original element order."""
    if a.shape[0] != rel.n_elements:
        raise ValueError("segment_softmax: %d rows for a relation of %d elements" % (a.shape[0], rel.n_elements))
    return _SegmentSoftmax.apply(a, rel, normalize)


def gather_mul_segment_sum(x, w, rel_out, rel_src):
    """Fused  y_i = sum_{e->i} x[s_e] * w_e  (SchNet cfconv, PaiNN scalar channel, GCN)."""
    tail = tuple(x.shape[1:])
    x2 = _flat2d(x)
    if w is None:
        w2 = None
        y = _GMS_noW.apply(x2, rel_out, rel_src)
    else:
        w2 = _flat2d(w) if w.dim() > 1 else w.unsqueeze(1)
        y = _GMS.apply(x2, w2, rel_out, rel_src)
    return y.reshape((rel_out.n_targets,) + tail)


class _GMS_noW(torch.autograd.Function):
    """y[i] = sum_{e: out.ids[e]==i} x[src.ids[e]]  (plain gather + segment-sum; the roofline op)."""

    @staticmethod
    def forward(ctx, x, rel_out, rel_src):
        ctx.rel_out, ctx.rel_src = rel_out, rel_src
        return raw_gms(x, None, rel_out, rel_src, OP_SUM)

    @staticmethod
    def backward(ctx, g):
        return _GMS_noW.apply(g.contiguous(), ctx.rel_src, ctx.rel_out), None, None


def gather2_mul(a, b, rel_a, rel_b, reduce=False):
    return _G2M.apply(_flat2d(a), _flat2d(b), rel_a, rel_b, reduce)


# ------------------------------------------------------------------------------------------------
# bias + activation family (a19) and deterministic column sum
# ------------------------------------------------------------------------------------------------
def act_code(activation):
    """Keras/kgcnn activation identifier -> kernel enum ('kgcnn>shifted_softplus', 'swish', None, ...)."""
    if activation is None:
        return _lib.ACT_CODES["linear"]
    if isinstance(activation, dict):
        activation = activation.get("class_name", activation.get("config", {}).get("name"))
    name = str(activation)
    if ">" in name:
        name = name.split(">")[-1]
    if name not in _lib.ACT_CODES:
        raise ValueError("Unknown activation function: %s" % activation)
    return _lib.ACT_CODES[name]


def raw_bias_act(pre, bias, g1, g2, act, order):
    _need_cuda(pre, bias, g1, g2)
    pre = _f32c(pre)
    U = int(pre.shape[-1])
    M = pre.numel() // U if U else 0
    out = torch.empty_like(pre)
    g1 = None if g1 is None else _f32c(g1)
    g2 = None if g2 is None else _f32c(g2)
    bias = None if bias is None else _f32c(bias)
    check(_lib.lib().kgcnn_bias_act(_ptr(pre), _ptr(bias), _ptr(g1), _ptr(g2), act, order, M, U, _ptr(out),
                                    _stream()), "kgcnn_bias_act")
    return out


BIAS_ACT_PAIR = True       # False: the two VJPs of an activation node as two launches (diagnostics, A/B timing)


def raw_bias_act_pair(pre, bias, gy, g1, act, order):
    """(gy * g1 * f^(order+1)(pre + bias), gy * f^(order)(pre + bias)) in one launch."""
    _need_cuda(pre, bias, gy, g1)
    pre, gy, g1 = _f32c(pre), _f32c(gy), _f32c(g1)
    U = int(pre.shape[-1])
    M = pre.numel() // U if U else 0
    d_pre, d_g1 = torch.empty_like(pre), torch.empty_like(pre)
    bias = None if bias is None else _f32c(bias)
    check(_lib.lib().kgcnn_bias_act_pair(_ptr(pre), _ptr(bias), _ptr(gy), _ptr(g1), act, order, M, U, _ptr(d_pre),
                                         _ptr(d_g1), _stream()), "kgcnn_bias_act_pair")
    return d_pre, d_g1


def raw_colsum(a2d):
    _need_cuda(a2d)
    a2d = _f32c(a2d)
    M, U = int(a2d.shape[0]), int(a2d.shape[1])
    out = torch.empty((U,), dtype=torch.float32, device=a2d.device)
    nbytes = int(_lib.lib().kgcnn_colsum_workspace_bytes(M, U))
    ws = torch.empty((max(nbytes, 4),), dtype=torch.uint8, device=a2d.device)
    check(_lib.lib().kgcnn_colsum(_ptr(a2d), M, U, _ptr(out), _ptr(ws), nbytes, _stream()), "kgcnn_colsum")
    return out


class _ColSum(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a):
        ctx.rows = int(a.shape[0])
        return raw_colsum(a)

    @staticmethod
    def backward(ctx, g):
        return g.unsqueeze(0).expand(ctx.rows, g.shape[0])


def colsum(a):
    return _ColSum.apply(a.reshape(-1, a.shape[-1]))


class _BiasAct(torch.autograd.Function):
    """out = g1 * g2 * f^(order)(pre + bias); its VJPs are the same kernel at order+1 / with swapped multipliers."""

    @staticmethod
    def forward(ctx, pre, bias, g1, g2, act, order):
        ctx.act, ctx.order = act, order
        ctx.save_for_backward(pre, bias, g1, g2)
        return raw_bias_act(pre, bias, g1, g2, act, order)

    @staticmethod
    def backward(ctx, gy):
        pre, bias, g1, g2 = ctx.saved_tensors
        gy = gy.contiguous()
        act, k = ctx.act, ctx.order
        g_pre = g_bias = g_g1 = g_g2 = None
        want_bias = bias is not None and _want(ctx, 1, bias)
        if (BIAS_ACT_PAIR and not torch.is_grad_enabled() and k <= 2 and ctx.needs_input_grad[0] and not want_bias
                and g1 is not None and g2 is None and ctx.needs_input_grad[2]):
            # last sweep (nothing differentiates through it again): both VJPs in one pass over (pre, gy, g1)
            g_pre, g_g1 = raw_bias_act_pair(pre, bias, gy, g1, act, k)
            return g_pre, None, g_g1, None, None, None
        if ctx.needs_input_grad[0] or want_bias:
            if k >= 3:
                raise NotImplementedError("activation derivatives above order 3 are not instantiated")
            if g1 is not None and g2 is not None:
                t = _BiasAct.apply(pre, bias, gy * g1, g2, act, k + 1)
            else:
                t = _BiasAct.apply(pre, bias, gy, g1 if g1 is not None else g2, act, k + 1)
            if ctx.needs_input_grad[0]:
                g_pre = t
            if want_bias:
                g_bias = colsum(t)
        if g1 is not None and ctx.needs_input_grad[2]:
            g_g1 = _BiasAct.apply(pre, bias, gy, g2, act, k)
        if g2 is not None and ctx.needs_input_grad[3]:
            g_g2 = _BiasAct.apply(pre, bias, gy, g1, act, k)
        return g_pre, g_bias, g_g1, g_g2, None, None


def bias_act(pre, bias, activation):
    """activation(pre + bias) with any-order derivatives through our kernels."""
    code = activation if isinstance(activation, int) else act_code(activation)
    if code == _lib.ACT_CODES["linear"]:
        return pre if bias is None else pre + bias
    return _BiasAct.apply(pre, bias, None, None, code, 0)


class WeightGradSink:
    """Deferred, grouped weight gradients of one backward sweep.

    The (dW, db) = (x^T g, colsum g) products of the Dense layers are leaves of the backward graph: nothing but
    the optimizer reads them.  While a sink is active (and no graph is being recorded for a further derivative),
    ``_Linear`` / ``_MatMul`` hand their operands over instead of launching one small GEMM + reduce + column sum
    each; ``flush()`` submits the whole list to ``kgcnn_dense_tn_grouped`` (two launches), which ACCUMULATES into
    the registered flat gradient buffer.  Operands are kept alive until the flush; ``max_bytes`` bounds that."""

    def __init__(self, params, flat_grad, offsets, max_bytes=6 << 30, max_items=256):
        self.flat = flat_grad
        self.slots = {}
        for p, off in zip(params, offsets):
            self.slots[p.data_ptr()] = (off, tuple(p.shape), p)
        self.items = []          # (slot_offset, x, g, K, U, bias_offset | None)
        self.bytes = 0
        self.max_bytes, self.max_items = max_bytes, max_items
        self.covered = set()

    def lookup(self, w):
        s = self.slots.get(w.data_ptr())
        return s if s is not None and s[1] == tuple(w.shape) else None

    def add(self, w, x, g, bias=None):
        """dW(w) += x^T g (and db(bias) += colsum g).  Returns False if `w` is not a registered variable."""
        if x.dim() != 2 or g.dim() != 2 or x.dtype != torch.float32 or not x.is_cuda:
            return False
        sw = self.lookup(w)
        sb = self.lookup(bias) if bias is not None else None
        if sw is None or (bias is not None and sb is None):
            return False
        x, g = _f32c(x.detach()), _f32c(g.detach())
        self.items.append((sw[0], x, g, int(x.shape[1]), int(g.shape[1]), None if sb is None else sb[0]))
        self.covered.add(sw[0])
        if sb is not None:
            self.covered.add(sb[0])
        self.bytes += (x.numel() + g.numel()) * 4
        if self.bytes > self.max_bytes or len(self.items) >= self.max_items:
            self.flush()
        return True

    def flush(self):
        import ctypes
        import numpy as np
        if not self.items:
            return
        items = sorted(self.items, key=lambda it: (it[0], it[5] is None))     # same dW adjacent, bias carrier first
        n = len(items)
        base = self.flat.data_ptr()
        X = np.array([it[1].data_ptr() for it in items], dtype=np.uint64)
        G = np.array([it[2].data_ptr() for it in items], dtype=np.uint64)
        M = np.array([it[1].shape[0] for it in items], dtype=np.int64)
        K = np.array([it[3] for it in items], dtype=np.int32)
        U = np.array([it[4] for it in items], dtype=np.int32)
        dW = np.array([base + 4 * it[0] for it in items], dtype=np.uint64)
        db = np.array([0 if it[5] is None else base + 4 * it[5] for it in items], dtype=np.uint64)
        hp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
        nbytes = int(_lib.lib().kgcnn_dense_tn_grouped_workspace_bytes(n, hp(M), hp(K), hp(U)))
        ws = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device=self.flat.device)
        check(_lib.lib().kgcnn_dense_tn_grouped(n, hp(X), hp(G), hp(M), hp(K), hp(U), hp(dW), hp(db), _ptr(ws),
                                                nbytes, _stream()), "kgcnn_dense_tn_grouped")
        self.items = []
        self.bytes = 0


_INPUT_ONLY = None


class input_gradients_only:
    """Context for a derivative with respect to DATA tensors only (the inner dE/dx of EnergyForceModel):
    ``ctx.needs_input_grad`` of a Python autograd Function is fixed at forward time, so without this the backward
    of every Dense would also compute its (unused) weight and bias gradients in that sweep.  Inside the context
    the gradients of LEAF operands other than `wanted` are skipped (returned as None)."""

    def __init__(self, *wanted):
        self.ptrs = frozenset(t.data_ptr() for t in wanted)

    def __enter__(self):
        global _INPUT_ONLY
        self._prev, _INPUT_ONLY = _INPUT_ONLY, self.ptrs

    def __exit__(self, *exc):
        global _INPUT_ONLY
        _INPUT_ONLY = self._prev


def _want(ctx, i, t):
    """Does this backward sweep need the gradient of input i (tensor t)?"""
    if not ctx.needs_input_grad[i]:
        return False
    return not (_INPUT_ONLY is not None and t is not None and t.is_leaf and t.data_ptr() not in _INPUT_ONLY)


_SINK = None


class weight_grad_sink:
    """Context: route Dense weight gradients of non-differentiated backward sweeps into `sink`."""

    def __init__(self, sink):
        self.sink = sink

    def __enter__(self):
        global _SINK
        self._prev, _SINK = _SINK, self.sink
        return self.sink

    def __exit__(self, *exc):
        global _SINK
        if exc[0] is None:
            self.sink.flush()
        else:
            self.sink.items, self.sink.bytes = [], 0
        _SINK = self._prev


def _sink():
    """The active sink, unless this backward sweep is itself being recorded (create_graph=True)."""
    return _SINK if (_SINK is not None and not torch.is_grad_enabled()) else None


def _dense_operands(*tensors):
    """Dense contractions exist on the tcgen05 path only: CUDA float32 matrices, anything else raises."""
    _need_cuda(*tensors)
    for t in tensors:
        if t is not None and t.dtype != torch.float32:
            raise TypeError("kgcnn_b200 Dense kernels compute in float32, got %s" % t.dtype)


class _MatMul(torch.autograd.Function):
    """Bilinear products closed under differentiation.  mode 0: a @ b, 1: a @ b^T, 2: a^T @ b, all three on the
    tcgen05 tensor cores (three-piece bf16 split, fp32 TMEM accumulation; dense_tc.cu)."""

    @staticmethod
    def forward(ctx, a, b, mode):
        ctx.mode = mode
        ctx.save_for_backward(a, b)
        _dense_operands(a, b)
        if a.dim() != 2 or b.dim() != 2:
            raise ValueError("matmul operands must be matrices, got %s and %s" % (tuple(a.shape), tuple(b.shape)))
        return raw_dense_tc(mode, a, b)              # tcgen05 tensor cores, bf16x3 split

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        ga = gb = None
        mode = ctx.mode
        sink = _sink()
        if mode == 0:       # y = a b
            if _want(ctx, 0, a):
                ga = _MatMul.apply(g, b, 1)          # g b^T
            if _want(ctx, 1, b) and not (sink is not None and sink.add(b, a, g)):
                gb = _MatMul.apply(a, g, 2)          # a^T g
        elif mode == 1:     # y = a b^T
            if _want(ctx, 0, a):
                ga = _MatMul.apply(g, b, 0)          # g b
            if _want(ctx, 1, b) and not (sink is not None and sink.add(b, g, a)):
                gb = _MatMul.apply(g, a, 2)          # g^T a
        else:               # y = a^T b
            if _want(ctx, 0, a):
                ga = _MatMul.apply(b, g, 1)          # b g^T
            if _want(ctx, 1, b):
                gb = _MatMul.apply(a, g, 0)          # a g
        return ga, gb, None


class _BiasAdd(torch.autograd.Function):
    """pre + bias with the bias gradient through the deterministic column-sum kernel."""

    @staticmethod
    def forward(ctx, pre, bias):
        ctx.bias_leaf = bias.is_leaf
        return pre + bias

    @staticmethod
    def backward(ctx, g):
        skip = _INPUT_ONLY is not None and ctx.bias_leaf
        return g, (colsum(g) if ctx.needs_input_grad[1] and not skip else None)


def matmul(a, b):
    return _MatMul.apply(a, b, 0)


DENSE_FUSE_ACT = True      # False: Dense = GEMM launch + bias/activation launch (diagnostics, A/B timing)


class _Linear(torch.autograd.Function):
    """pre = x @ W + b.  Backward: gx = g W^T (tcgen05 NT); (gW, gb) deferred to the grouped TN launch or _TNCol."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        ctx.bias_ref = b.detach() if b is not None else None      # identity only (the sink's variable lookup)
        ctx.bias_leaf = b is not None and b.is_leaf
        _dense_operands(x, w, b)
        return raw_dense_tc(0, x, w, b)              # tcgen05 tensor cores, bf16x3 split, bias in the epilogue

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        return _linear_backward(ctx, x, w, g.contiguous())


def _linear_backward(ctx, x, w, g):
    """(gx, gW, gb) of pre = x W + b for the cotangent g of pre (shared by _Linear and _LinearAct)."""
    gx = gw = gb = None
    if _want(ctx, 0, x):
        gx = _MatMul.apply(g, w, 1)
    if _want(ctx, 1, w) or (ctx.has_bias and _want(ctx, 2, ctx.bias_ref if ctx.bias_leaf else None)):
        sink = _sink()
        if sink is not None and ctx.needs_input_grad[1] and sink.add(w, x, g, ctx.bias_ref):
            pass                                  # deferred: accumulated by the grouped kernel at flush
        elif ctx.has_bias:
            gw, gb = _TNCol.apply(x, g)
        else:
            gw = _MatMul.apply(x, g, 2)
    return gx, gw, gb


class _LinearAct(torch.autograd.Function):
    """(pre, y) = (x W + b, f(x W + b)) in ONE launch (activation in the tcgen05 epilogue, csrc/dense_tc.cu).
    Backward is the composition the unfused path records: g = g_pre + g_y * f'(pre) through the bias/activation
    kernel family (differentiable again: `pre` is this node's own output), then the Linear backward."""

    @staticmethod
    def forward(ctx, x, w, b, act):
        ctx.has_bias = b is not None
        ctx.bias_ref = b.detach() if b is not None else None
        ctx.bias_leaf = b is not None and b.is_leaf
        ctx.act = act
        ctx.set_materialize_grads(False)
        pre, y = raw_dense_act(x, w, b, act)
        ctx.save_for_backward(x, w, pre)
        return pre, y

    @staticmethod
    def backward(ctx, g_pre, g_y):
        x, w, pre = ctx.saved_tensors
        g = None
        if g_y is not None:
            g = _BiasAct.apply(pre, None, g_y.contiguous(), None, ctx.act, 1)
        if g_pre is not None:
            g = g_pre if g is None else g + g_pre
        if g is None:
            return None, None, None, None
        return _linear_backward(ctx, x, w, g.contiguous()) + (None,)


class _TNCol(torch.autograd.Function):
    """(dW, db) = (x^T g, colsum(g)); its VJP is a Linear again: closed under differentiation."""

    @staticmethod
    def forward(ctx, x, g):
        ctx.save_for_backward(x, g)
        _dense_operands(x, g)
        return raw_dense_tc(2, x, g), raw_colsum(g)

    @staticmethod
    def backward(ctx, ggw, ggb):
        x, g = ctx.saved_tensors
        gx = gg = None
        if ctx.needs_input_grad[0]:
            gx = _MatMul.apply(g, ggw.contiguous(), 1)              # g ggW^T
        if ctx.needs_input_grad[1]:
            gg = _Linear.apply(x, ggw.contiguous(), ggb)            # x ggW + ggb
        return gx, gg


def dense(x, kernel, bias, activation):
    """DenseEmbedding on flat values (kgcnn/layers/modules.py:85): activation(x @ kernel + bias).
    x @ W (+ b), g @ W^T and x^T g run on the tcgen05 tensor cores (bf16x3 split); the activation and all its derivatives
    in the fused bias/activation kernel family."""
    code = activation if isinstance(activation, int) else act_code(activation)
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    _dense_operands(x2, kernel, bias)
    if code != _lib.ACT_CODES["linear"] and DENSE_FUSE_ACT:
        _, out = _LinearAct.apply(x2, kernel, bias, code)
    else:
        pre = _Linear.apply(x2, kernel, bias)
        out = pre if code == _lib.ACT_CODES["linear"] else _BiasAct.apply(pre, None, None, None, code, 0)
    return out.reshape(tuple(lead) + (kernel.shape[1],))


# ------------------------------------------------------------------------------------------------
# fused continuous-filter convolution (a13): filter network inside the gather-scale-scatter
# ------------------------------------------------------------------------------------------------
CFCONV_FUSED = True     # False: Dense -> Dense -> fused gather-multiply-pool (filter [E, units] through HBM)


class CfconvSpec:
    """Static part of one fused cfconv call (kgcnn/layers/conv/schnet_conv.py:73-79 behind GaussBasisLayer,
    kgcnn/layers/geom.py:567-570): receiver / sender relations, Gaussian basis, activation of the first Dense."""

    def __init__(self, recv, send, bins, gamma, offset, step, act):
        self.recv, self.send = recv, send
        self.R, self.gamma, self.offset, self.step, self.act = int(bins), float(gamma), float(offset), float(step), int(act)

    @staticmethod
    def supported(units, bins, act, w1, w2):
        return (CFCONV_FUSED and units == 128 and 1 <= bins <= 32
                and act in (_lib.ACT_CODES["shifted_softplus"], _lib.ACT_CODES["softplus2"])
                and tuple(w1.shape) == (bins, 128) and tuple(w2.shape) == (128, 128)
                and w1.is_cuda and w1.dtype == torch.float32 and w2.dtype == torch.float32)


_STATIC_WEIGHT_RANGE = None          # [lo, hi) device addresses of the flat buffer whose owner vouches for its writers


class static_weights:
    """Opt-in of the code that OWNS the model variables: while this context is active, tensors that live inside `flat`
    (FlatAdam's parameter buffer) are written by nothing but the owner's own update kernel, which is never the launch
    right before a layer kernel -- so the fused cfconv may stage them before its launch-dependency wait
    (csrc/common.cuh: otherwise no global access before pdl_wait()).  Without it (the default) the kernel waits first:
    a foreign optimizer, an EMA / constraint kernel or a load_state_dict copy may have written the weights last.
    `flat=None` is a no-op (the optimizer does not exist yet)."""

    def __init__(self, flat):
        self.range = None if flat is None else (flat.data_ptr(), flat.data_ptr() + flat.numel() * flat.element_size())

    def __enter__(self):
        global _STATIC_WEIGHT_RANGE
        self._prev, _STATIC_WEIGHT_RANGE = _STATIC_WEIGHT_RANGE, (self.range or _STATIC_WEIGHT_RANGE)
        return self

    def __exit__(self, *exc):
        global _STATIC_WEIGHT_RANGE
        _STATIC_WEIGHT_RANGE = self._prev


def static_weight_flag(tensors):
    """1 only if an owner has opted in (static_weights) and every tensor handed to the kernel -- AFTER the contiguous /
    fp32 conversion, so a fresh copy never qualifies -- lies inside the owner's buffer."""
    if _STATIC_WEIGHT_RANGE is None:
        return 0
    lo, hi = _STATIC_WEIGHT_RANGE
    return int(all(t is None or (t.numel() > 0 and lo <= t.data_ptr() and
                                 t.data_ptr() + t.numel() * t.element_size() <= hi) for t in tensors))


def raw_cfconv(mode, spec, d, w1, b1, w2, b2, rel_out, rel_other, a, b=None, c=None, tang=None, want_wg=False):
    """One launch of kgcnn_cfconv_fused (+ its fix-up): (out [n_out,128], edge_out [E] | None, wg | None).
    want_wg (modes 1, 3): also the operand pairs of the filter network's weight gradients, as a dict
    {"w2": [(x, g, with_bias), ...], "w1": [...]} of [E, *] tensors (rows in the CSR order of this call)."""
    import ctypes
    _need_cuda(d, w1, w2, a)
    E = rel_out.n_elements
    dev = a.device
    d = _f32c(d.reshape(-1))
    if int(d.shape[0]) != E:
        raise ValueError("cfconv: %d distances for %d edges" % (int(d.shape[0]), E))
    a = _f32c(a)
    b = None if b is None else _f32c(b)
    c = None if c is None else _f32c(c)
    tang = None if tang is None else _f32c(tang.reshape(-1))
    w1, w2 = _f32c(w1), _f32c(w2)
    b1 = None if b1 is None else _f32c(b1)
    b2 = None if b2 is None else _f32c(b2)
    # staged before the kernel waits for its launch dependency only when the variables' owner has vouched for their
    # writers (static_weights); decided on the tensors the kernel really reads, after any copy
    static_w = static_weight_flag((w1, b1, w2, b2))
    out = torch.empty((rel_out.n_targets, 128), dtype=torch.float32, device=dev)
    edge_out = torch.empty((E,), dtype=torch.float32, device=dev) if mode in (1, 3) else None
    nbytes = int(_lib.lib().kgcnn_cfconv_fused_workspace_bytes(E))
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    wg = wg_arr = None
    if want_wg:
        new = lambda cols: torch.empty((E, cols), dtype=torch.float32, device=dev)
        t = [new(spec.R), new(128), new(128), new(128)]
        if mode == 3:
            t += [new(spec.R), new(128), new(128), new(128)]
        wg_arr = (ctypes.c_void_p * 8)(*([x.data_ptr() for x in t] + [None] * (8 - len(t))))
        wg = {"w2": [(t[1], t[2], True)], "w1": [(t[0], t[3], True)]}
        if mode == 3:
            wg["w2"].append((t[5], t[6], False))
            wg["w1"].append((t[4], t[7], False))
    check(_lib.lib().kgcnn_cfconv_fused(
        mode, _ptr(d), E, spec.R, spec.gamma, spec.offset, spec.step, _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), 128,
        spec.act, static_w, _ptr(rel_out.rowptr), _ptr(rel_out.permuted_ids(rel_out)),
        _ptr(rel_out.permuted_ids(rel_other)),
        _ptr(rel_out.perm), rel_out.n_targets, _ptr(a), _ptr(b), _ptr(c), _ptr(tang), _ptr(out), _ptr(edge_out),
        wg_arr, _ptr(ws), nbytes, _stream()), "kgcnn_cfconv_fused")
    return out, edge_out, wg


def raw_gauss_basis(d, spec, order, scale=None):
    d = _f32c(d.reshape(-1))
    E = int(d.shape[0])
    out = torch.empty((E, spec.R), dtype=torch.float32, device=d.device)
    scale = None if scale is None else _f32c(scale.reshape(-1))
    check(_lib.lib().kgcnn_gauss_basis(_ptr(d), E, spec.R, spec.gamma, spec.offset, spec.step, order, _ptr(scale),
                                       _ptr(out), _stream()), "kgcnn_gauss_basis")
    return out


class _GaussBasis(torch.autograd.Function):
    """rbf^(order)_k(d) = (d/dd)^order exp(-gamma (d - offset - k step)^2) on distances that do not come from the fused
    geometry kernel (models fed with edge distances, kgcnn/literature/Schnet.py make_distance=False;
    kgcnn/layers/geom.py:567-570).  The VJP of order n is order n + 1 contracted with the cotangent: closed up to the
    second derivative the kernel provides."""

    @staticmethod
    def forward(ctx, d, basis, order):
        ctx.basis, ctx.order = basis, order
        ctx.save_for_backward(d)
        return raw_gauss_basis(d, basis, order)

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        if ctx.order >= 2:
            raise RuntimeError("Gaussian basis: derivatives beyond the third are not provided by kgcnn_gauss_basis")
        nxt = _GaussBasis.apply(d, ctx.basis, ctx.order + 1)
        return (g * nxt).sum(dim=1).reshape(d.shape), None, None


class GaussBasisParams:
    def __init__(self, bins, gamma, offset, step):
        self.R, self.gamma, self.offset, self.step = int(bins), float(gamma), float(offset), float(step)


def gauss_basis(d, bins, gamma, offset, step):
    """[E] or [E,1] distances -> [E, bins] Gaussian expansion (one launch; differentiable twice)."""
    _need_cuda(d)
    return _GaussBasis.apply(_f32c(d), GaussBasisParams(bins, gamma, offset, step), 0)


def _cfconv_weight_grads(wg, w1, b1, w2, b2):
    """Hand the operand pairs a WG launch produced to the grouped weight-gradient launch (active sink) or contract
    them here.  Returns (gW1, gb1, gW2, gb2); entries that went to the sink are None."""
    sink = _sink()
    direct = {}
    for w, bias, key in ((w2, b2, "w2"), (w1, b1, "w1")):
        for x, g, with_bias in wg[key]:
            bb = bias if with_bias else None
            if sink is not None and sink.add(w, x, g, bb):
                continue
            direct[id(w)] = direct.get(id(w), 0) + raw_dense_tc(2, x, g)
            if bb is not None:
                direct[id(bb)] = direct.get(id(bb), 0) + raw_colsum(g)
    return tuple(direct.get(id(t)) if t is not None else None for t in (w1, b1, w2, b2))


def _cfconv_unfused(x, d, w1, b1, w2, b2, spec):
    """The same layer op by op (basis, Dense, Dense, fused gather-multiply-pool) with the closed primitives."""
    k = torch.arange(spec.R, dtype=d.dtype, device=d.device)
    t = d.reshape(-1, 1) - spec.offset - spec.step * k
    rbf = torch.exp(-spec.gamma * t * t)
    filt = dense(dense(rbf, w1, b1, spec.act), w2, b2, _lib.ACT_CODES["linear"])
    return gather_mul_segment_sum(x, filt, spec.recv, spec.send)


class _CfConv(torch.autograd.Function):
    """y[i] = sum_{e->i} x[s_e] * phi(d_e),  phi = Dense2(ssp(Dense1(gauss(d))))  -- one fused launch, nothing of size
    [E, units] in HBM (csrc/cfconv.cu mode 0).  Backward = _CfConvVjp (differentiable once more)."""

    @staticmethod
    def forward(ctx, x, d, w1, b1, w2, b2, spec):
        ctx.spec = spec
        ctx.save_for_backward(x, d, w1, b1, w2, b2)
        return raw_cfconv(0, spec, d, w1, b1, w2, b2, spec.recv, spec.send, x)[0]

    @staticmethod
    def backward(ctx, gy):
        x, d, w1, b1, w2, b2 = ctx.saved_tensors
        spec = ctx.spec
        gy = gy.contiguous()
        wanted_w = [t is not None and _want(ctx, i, t) for i, t in ((2, w1), (3, b1), (4, w2), (5, b2))]
        if any(wanted_w) and torch.is_grad_enabled():
            # weight gradients that are themselves differentiated again (not a training step): record them through
            # the unfused chain of closed primitives
            ins = [x, d, w1, b1, w2, b2]
            want = [_want(ctx, 0, x), _want(ctx, 1, d)] + wanted_w
            with torch.enable_grad():
                y2 = _cfconv_unfused(x, d, w1, b1, w2, b2, spec)
                gs = torch.autograd.grad(y2, [t for t, w in zip(ins, want) if w], gy, create_graph=True,
                                         allow_unused=True)
            it = iter(gs)
            return tuple(next(it) if w else None for w in want) + (None,)
        if any(wanted_w):      # weight-gradient sweep: the VJP kernel also emits the operands of dW1, db1, dW2, db2
            gx, gd, wg = raw_cfconv(1, spec, d, w1, b1, w2, b2, spec.send, spec.recv, gy, b=x, want_wg=True)
            gw = _cfconv_weight_grads(wg, w1, b1, w2, b2)
            return (gx, gd.reshape(d.shape)) + gw + (None,)
        gx = gd = None
        if _want(ctx, 0, x) or _want(ctx, 1, d):
            gx, gd = _CfConvVjp.apply(gy, x, d, w1, b1, w2, b2, spec)
            gd = gd.reshape(d.shape)
        return (gx, gd, None, None, None, None, None)


class _CfConvVjp(torch.autograd.Function):
    """(gx[s], gd[e]) = (sum_{e: s_e = s} gy[r_e] * phi(d_e),  <gy[r_e] * x[s_e], phi'(d_e)>)  (csrc/cfconv.cu mode 1).
    Its own VJP (the weight-gradient sweep of force training) is the JVP of the layer (mode 2) for gy, mode 3 for x
    and d, and the reverse sweep through the filter network for the weights."""

    @staticmethod
    def forward(ctx, gy, x, d, w1, b1, w2, b2, spec):
        ctx.spec = spec
        ctx.save_for_backward(gy, x, d, w1, b1, w2, b2)
        gx, gd, _ = raw_cfconv(1, spec, d, w1, b1, w2, b2, spec.send, spec.recv, gy, b=x)
        return gx, gd

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, ggx, ggd):
        gy, x, d, w1, b1, w2, b2 = ctx.saved_tensors
        spec = ctx.spec
        ggx = torch.zeros_like(x) if ggx is None else ggx.contiguous()
        ggd = torch.zeros((spec.recv.n_elements,), dtype=torch.float32, device=x.device) if ggd is None \
            else ggd.contiguous().reshape(-1)
        ggy = g2x = g2d = None
        if ctx.needs_input_grad[0]:
            ggy = raw_cfconv(2, spec, d, w1, b1, w2, b2, spec.recv, spec.send, ggx, b=x, tang=ggd)[0]
        want_w = any(t is not None and _want(ctx, i, t) for i, t in ((3, w1), (4, b1), (5, w2), (6, b2)))
        gw = (None,) * 4
        if want_w or _want(ctx, 1, x) or _want(ctx, 2, d):
            g2x, g2d, wg = raw_cfconv(3, spec, d, w1, b1, w2, b2, spec.send, spec.recv, gy, b=x, c=ggx, tang=ggd,
                                      want_wg=want_w)
            g2d = g2d.reshape(d.shape)
            if want_w:
                gw = _cfconv_weight_grads(wg, w1, b1, w2, b2)
        return (ggy, g2x, g2d) + gw + (None,)


def cfconv_fused(x, d, w1, b1, w2, b2, spec):
    """SchNet cfconv from the distances: x [N,128], d [E] or [E,1] -> [N,128]."""
    return _CfConv.apply(x, d, w1, b1, w2, b2, spec)


# ------------------------------------------------------------------------------------------------
# fused ACSF G2 / G4 (a15, a16): G, J^T v, J u
# ------------------------------------------------------------------------------------------------
COORDINATE_HESSIAN = True     # False inside a training step: coordinates are not trained, d2/dx2 is never needed


class no_coordinate_hessian:
    """Inside this context the fused geometry kernels do not provide second derivatives with respect to the
    coordinates (only what force TRAINING needs: gradients of the force with respect to the weights)."""

    def __enter__(self):
        global COORDINATE_HESSIAN
        self._prev, COORDINATE_HESSIAN = COORDINATE_HESSIAN, False

    def __exit__(self, *exc):
        global COORDINATE_HESSIAN
        COORDINATE_HESSIAN = self._prev


def _no_hessian(what):
    if COORDINATE_HESSIAN:
        raise NotImplementedError(
            "second derivatives with respect to the coordinates through the fused %s kernels are not implemented; "
            "wrap weight-gradient computations in functional.no_coordinate_hessian() (ForceTrainStep does) or use "
            "the unfused layer path (layers.conv.acsf_conv.FUSED = False)" % what)
    return None


class ACSFSpec:
    """Per-batch constants of one ACSF layer call (index relations, element maps, parameter table)."""

    def __init__(self, kind, rel_centre, table, R, m, eps, rel_atom=None, tset_atom=None, rel_send=None,
                 z_atom=None, rev_pair=None, zmax=31, rel_j=None, rel_k=None, multiplicity=0.0, uniform_rc=False,
                 weight=None, mu_table=None, factor=None):
        self.uniform_rc = uniform_rc
        self.factor = factor       # G4: (factor_table, factor_index) device tensors from g4_factor_tables, or None
        self.weight, self.mu_table = weight, mu_table      # wACSF: per-pair/triple factor, Gaussian shifts
        self.kind, self.rel_centre, self.table, self.R, self.m, self.eps = kind, rel_centre, table, R, m, eps
        self.rel_atom, self.tset_atom, self.rel_send = rel_atom, tset_atom, rel_send
        self.z_atom, self.rev_pair, self.zmax, self.rel_j, self.rel_k = z_atom, rev_pair, zmax, rel_j, rel_k
        self.multiplicity = multiplicity


G4_FACTORISE = True        # False: always the unfactorised parameter loop (diagnostics)


def g4_factor_tables(table, mu=None):
    """Host side, once per layer: distinct Gaussian (eta, mu) and angular (zeta, lambda) terms of every (T, R) row
    of a G4 parameter table [..., m, 4] = (eta, zeta, lambda, Rc) and the slot pair of each of its m parameter sets
    (include/kgcnn_b200.h, kgcnn_acsf_g4).  Returns (factor_table float32 [rows, 32, 2], factor_index int32
    [rows, m]) as NumPy arrays, or None when a row has more than 16 distinct terms of either kind or the cutoff
    radius is not uniform (the kernel then runs its general loop)."""
    import numpy as np
    table = np.asarray(table, dtype=np.float32)
    m = table.shape[-2]
    t = table.reshape(-1, m, 4)
    if not np.all(t[..., 3] == t[0, 0, 3]):
        return None
    mus = np.zeros(t.shape[:2], np.float32) if mu is None else np.asarray(mu, dtype=np.float32).reshape(-1, m)
    fac = np.zeros((t.shape[0], 32, 2), np.float32)
    idx = np.zeros((t.shape[0], m), np.int32)
    for row in range(t.shape[0]):
        gauss, ang = {}, {}
        for q in range(m):
            ge = gauss.setdefault((float(t[row, q, 0]), float(mus[row, q])), len(gauss))
            an = ang.setdefault((float(t[row, q, 1]), float(t[row, q, 2])), len(ang))
            idx[row, q] = ge | ((16 + an) << 8)
        if len(gauss) > 16 or len(ang) > 16:
            return None
        for (a, b), slot in gauss.items():
            fac[row, slot] = (a, b)
        for (a, b), slot in ang.items():
            fac[row, 16 + slot] = (a, b)
    return fac, idx


def _g4_factor_tables(spec):
    if not (G4_FACTORISE and spec.uniform_rc and spec.factor is not None):
        return None, None
    return spec.factor


def raw_acsf(xyz, spec, mode, cot=None):
    _need_cuda(xyz, cot)
    xyz = _f32c(xyz)
    N = int(xyz.shape[0])
    RM = spec.R * spec.m
    rc = spec.rel_centre
    out = torch.empty((N, 3 if mode == 1 else RM), dtype=torch.float32, device=xyz.device)
    cot = None if cot is None else _f32c(cot)
    if spec.kind == 2:
        M = rc.n_elements
        edge_out = torch.empty((M, 3), dtype=torch.float32, device=xyz.device) if mode == 1 else None
        send = rc.permuted_ids(spec.rel_send)
        check(_lib.lib().kgcnn_acsf_g2(_ptr(xyz), _ptr(spec.rel_atom), _ptr(spec.tset_atom), _ptr(rc.rowptr),
                                       _ptr(rc.perm), _ptr(send), 1 if rc.perm is not None else 0, _ptr(spec.table),
                                       spec.R, spec.m, spec.eps, N, _ptr(spec.weight), mode, _ptr(cot), _ptr(out),
                                       _ptr(edge_out), _stream()), "kgcnn_acsf_g2")
        return out, edge_out
    A = rc.n_elements
    tri = torch.empty((2, A, 3), dtype=torch.float32, device=xyz.device) if mode == 1 else None
    fac, fidx = _g4_factor_tables(spec)
    check(_lib.lib().kgcnn_acsf_g4(_ptr(xyz), _ptr(spec.z_atom), _ptr(spec.tset_atom), _ptr(spec.rev_pair), spec.zmax,
                                   _ptr(rc.rowptr), _ptr(rc.perm), _ptr(spec.rel_j.ids), _ptr(spec.rel_k.ids),
                                   _ptr(spec.table), spec.R, spec.m, spec.eps, float(spec.multiplicity or 0.0),
                                   1 if spec.uniform_rc else 0, N, _ptr(spec.mu_table), _ptr(spec.weight), _ptr(fac),
                                   _ptr(fidx), A, mode, _ptr(cot), _ptr(out), _ptr(tri), _stream()), "kgcnn_acsf_g4")
    return out, tri


def _acsf_vjp_raw(xyz, dG, spec):
    centre, nb = raw_acsf(xyz, spec, 1, dG)
    if spec.kind == 2:
        return centre + raw_gms(nb, None, spec.rel_send, None)
    return centre + raw_gms(nb[0], None, spec.rel_j, None) + raw_gms(nb[1], None, spec.rel_k, None)


class _ACSF(torch.autograd.Function):
    """G(xyz) for one ACSF layer (parameters constant)."""

    @staticmethod
    def forward(ctx, xyz, spec):
        ctx.spec = spec
        ctx.save_for_backward(xyz)
        return raw_acsf(xyz, spec, 0)[0]

    @staticmethod
    def backward(ctx, dG):
        (xyz,) = ctx.saved_tensors
        return _ACSFVjp.apply(xyz, dG.contiguous(), ctx.spec), None


class _ACSFVjp(torch.autograd.Function):
    """J(xyz)^T dG; linear in dG, so its VJP with respect to dG is J h (the JVP kernel)."""

    @staticmethod
    def forward(ctx, xyz, dG, spec):
        ctx.spec = spec
        ctx.save_for_backward(xyz)
        return _acsf_vjp_raw(xyz, dG, spec)

    @staticmethod
    def backward(ctx, h):
        (xyz,) = ctx.saved_tensors
        g_xyz = _no_hessian("ACSF") if ctx.needs_input_grad[0] else None
        g_dG = _ACSFJvp.apply(xyz, h.contiguous(), ctx.spec) if ctx.needs_input_grad[1] else None
        return g_xyz, g_dG, None


class _ACSFJvp(torch.autograd.Function):
    """J(xyz) u; its VJP with respect to u is J^T g."""

    @staticmethod
    def forward(ctx, xyz, u, spec):
        ctx.spec = spec
        ctx.save_for_backward(xyz)
        return raw_acsf(xyz, spec, 2, u)[0]

    @staticmethod
    def backward(ctx, g):
        (xyz,) = ctx.saved_tensors
        g_xyz = _no_hessian("ACSF") if ctx.needs_input_grad[0] else None
        g_u = _ACSFVjp.apply(xyz, g.contiguous(), ctx.spec) if ctx.needs_input_grad[1] else None
        return g_xyz, g_u, None


def acsf(xyz, spec):
    return _ACSF.apply(xyz, spec)


# ------------------------------------------------------------------------------------------------
# Dense contractions on tcgen05 tensor cores (bf16x3 split)
# ------------------------------------------------------------------------------------------------
def raw_dense_tc(op, a, b, bias=None):
    """op 0: a[M,K] @ b[K,U] (+bias); op 1: a[M,U] @ b[K,U]^T; op 2: a[M,K]^T @ b[M,U]."""
    _need_cuda(a, b, bias)
    a, b = _f32c(a), _f32c(b)
    M = int(a.shape[0])
    if op == 0:
        K, U = int(b.shape[0]), int(b.shape[1])
        out = torch.empty((M, U), dtype=torch.float32, device=a.device)
    elif op == 1:
        K, U = int(b.shape[0]), int(b.shape[1])
        out = torch.empty((M, K), dtype=torch.float32, device=a.device)
    else:
        K, U = int(a.shape[1]), int(b.shape[1])
        out = torch.empty((K, U), dtype=torch.float32, device=a.device)
    nbytes = int(_lib.lib().kgcnn_dense_tc_workspace_bytes(op, M, K, U))
    ws = torch.empty((max(nbytes, 4),), dtype=torch.uint8, device=a.device) if nbytes else None
    bias = None if bias is None else _f32c(bias)
    check(_lib.lib().kgcnn_dense_tc(op, _ptr(a), _ptr(b), M, K, U, _ptr(bias), _ptr(out), _ptr(ws), nbytes, _stream()),
          "kgcnn_dense_tc")
    return out


def raw_dense_act(x, w, bias, act):
    """(pre, f(pre)) with pre = x[M,K] @ w[K,U] + bias: the activation runs in the GEMM epilogue (one launch)."""
    _need_cuda(x, w, bias)
    x, w = _f32c(x), _f32c(w)
    M, K, U = int(x.shape[0]), int(w.shape[0]), int(w.shape[1])
    pre = torch.empty((M, U), dtype=torch.float32, device=x.device)
    y = torch.empty((M, U), dtype=torch.float32, device=x.device)
    bias = None if bias is None else _f32c(bias)
    check(_lib.lib().kgcnn_dense_act(_ptr(x), _ptr(w), M, K, U, _ptr(bias), int(act), _ptr(pre), _ptr(y), _stream()),
          "kgcnn_dense_act")
    return pre, y


# ------------------------------------------------------------------------------------------------
# relational Dense (f1): rows sorted by relation, one grouped launch per product
# ------------------------------------------------------------------------------------------------
def raw_dense_rel(op, a, b, offs, n_rel, bias=None):
    """Rows of `a` (and of `b` for op 2) are sorted by relation, offs = int32 [R + 1] device row offsets.
    op 0: a[M,K] @ b[R,K,U][rel] (+bias) -> [M,U]; op 1: a[M,U] @ b[R,K,U][rel]^T -> [M,K];
    op 2: per relation a[M,K]^T @ b[M,U] -> [R,K,U]."""
    _need_cuda(a, b, offs, bias)
    a, b = _f32c(a), _f32c(b)
    M, R = int(a.shape[0]), int(n_rel)
    if offs.dtype != torch.int32 or int(offs.shape[0]) != R + 1:
        raise ValueError("offs must be int32 [num_relations + 1]")
    if op == 2:
        K, U = int(a.shape[1]), int(b.shape[1])
        out = torch.empty((R, K, U), dtype=torch.float32, device=a.device)
    else:
        K, U = int(b.shape[1]), int(b.shape[2])
        out = torch.empty((M, U if op == 0 else K), dtype=torch.float32, device=a.device)
    nbytes = int(_lib.lib().kgcnn_dense_rel_workspace_bytes(op, R, K, U))
    ws = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device=a.device) if nbytes else None
    bias = None if bias is None else _f32c(bias)
    check(_lib.lib().kgcnn_dense_rel(op, _ptr(a), _ptr(b), _ptr(offs), R, M, K, U, _ptr(bias), _ptr(out), _ptr(ws),
                                     nbytes, _stream()), "kgcnn_dense_rel")
    return out


class _RelMatMul(torch.autograd.Function):
    """The three relational products, closed under differentiation like _MatMul (mode = op of raw_dense_rel)."""

    @staticmethod
    def forward(ctx, a, b, offs, n_rel, mode):
        ctx.mode, ctx.n_rel = mode, n_rel
        ctx.save_for_backward(a, b, offs)
        _dense_operands(a, b)
        return raw_dense_rel(mode, a, b, offs, n_rel)

    @staticmethod
    def backward(ctx, g):
        a, b, offs = ctx.saved_tensors
        R, mode = ctx.n_rel, ctx.mode
        g = g.contiguous()
        ga = gb = None
        if mode == 0:       # y = a kernel[rel]
            if _want(ctx, 0, a):
                ga = _RelMatMul.apply(g, b, offs, R, 1)
            if _want(ctx, 1, b):
                gb = _RelMatMul.apply(a, g, offs, R, 2)
        elif mode == 1:     # y = a kernel[rel]^T
            if _want(ctx, 0, a):
                ga = _RelMatMul.apply(g, b, offs, R, 0)
            if _want(ctx, 1, b):
                gb = _RelMatMul.apply(g, a, offs, R, 2)
        else:               # y[rel] = a^T b
            if _want(ctx, 0, a):
                ga = _RelMatMul.apply(b, g, offs, R, 1)
            if _want(ctx, 1, b):
                gb = _RelMatMul.apply(a, g, offs, R, 0)
        return ga, gb, None, None, None


def dense_rel(x_sorted, kernels, offs):
    """x_sorted[M,K] @ kernels[R,K,U][relation of the row]; rows sorted by relation (see RelationalDense)."""
    return _RelMatMul.apply(x_sorted, kernels, offs, int(kernels.shape[0]), 0)


class _PermuteRows(torch.autograd.Function):
    """out[i] = x[perm[i]] for a PERMUTATION perm (inv = its inverse): the gradient is the inverse gather, so no
    scatter is needed in either direction."""

    @staticmethod
    def forward(ctx, x, perm, inv):
        ctx.save_for_backward(perm, inv)
        return raw_gather_rows(_f32c(x), perm)

    @staticmethod
    def backward(ctx, g):
        perm, inv = ctx.saved_tensors
        return _PermuteRows.apply(g.contiguous(), inv, perm), None, None


def permute_rows(x, perm, inv):
    return _PermuteRows.apply(x, perm, inv)


# ------------------------------------------------------------------------------------------------
# fused PaiNN message (a14): one multilinear form, three kernels, closed under differentiation
# ------------------------------------------------------------------------------------------------
_PM_ORDER = ("gs", "gv", "phi", "v", "w", "r")
_PM_TERMS = {"gs": 1, "gv": 6, "phi": 7, "v": 2, "w": 7, "r": 4}        # terms of M that contain the operand
_PM_SLOTS = {"G": ("gs", "gv"), "J": ("phi", "v"), "E": ("w", "r")}     # kernel -> the two derivatives it writes


def raw_painn_msg(slot, mask, want_a, want_b, ops, recv, send, F_dim, acc_a=None):
    """(dM/dXa, dM/dXb) for the operand pair of `slot` (csrc/painn.cu); None for an output that is not wanted.
    acc_a (slot 'E' only): the first output is acc_a + dM/dW (sum of two substitutions without an add pass)."""
    t = {k: (None if x is None else _f32c(x)) for k, x in ops.items()}
    _need_cuda(*[x for x in t.values() if x is not None])
    dev = recv.ids.device
    N, E = recv.n_targets, recv.n_elements
    L = _lib.lib()

    def new(rows, cols, want):
        return torch.empty((rows, cols), dtype=torch.float32, device=dev) if want else None

    if slot == "G":
        a, b = new(N, F_dim, want_a), new(N, 3 * F_dim, want_b)
        m = mask & ((1 if want_a else 0) | (6 if want_b else 0))
        check(L.kgcnn_painn_msg_nodes(_ptr(t["phi"]), _ptr(t["v"]), _ptr(t["w"]), _ptr(t["r"]), _ptr(recv.rowptr),
                                      _ptr(recv.perm), _ptr(recv.permuted_ids(send)), 1 if recv.perm is not None else 0,
                                      N, E, F_dim, m, _ptr(a), _ptr(b), _stream()), "kgcnn_painn_msg_nodes")
        if b is not None:
            b = b.view(N, 3, F_dim)
    elif slot == "J":
        a, b = new(N, 3 * F_dim, want_a), new(N, 3 * F_dim, want_b)
        check(L.kgcnn_painn_msg_senders(_ptr(t["gs"]), _ptr(t["gv"]), _ptr(t["phi"]), _ptr(t["v"]), _ptr(t["w"]),
                                        _ptr(t["r"]), _ptr(send.rowptr), _ptr(send.perm),
                                        _ptr(send.permuted_ids(recv)), 1 if send.perm is not None else 0,
                                        N, E, F_dim, mask, _ptr(a), _ptr(b), _stream()), "kgcnn_painn_msg_senders")
        if b is not None:
            b = b.view(N, 3, F_dim)
    else:
        a, b = new(E, 3 * F_dim, want_a), new(E, 3, want_b)        # gr of an inactive term 3 is written as zeros
        check(L.kgcnn_painn_msg_edges(_ptr(t["gs"]), _ptr(t["gv"]), _ptr(t["phi"]), _ptr(t["v"]), _ptr(t["w"]),
                                      _ptr(t["r"]), _ptr(recv.ids), _ptr(send.ids), E, F_dim, mask,
                                      _ptr(None if acc_a is None or a is None else _f32c(acc_a)), _ptr(a), _ptr(b),
                                      _stream()), "kgcnn_painn_msg_edges")
    return a, b


class _PainnMsg(torch.autograd.Function):
    """dM/d(operand pair of `slot`) of the PaiNN message form M(Gs,Gv,Phi,V,W,R) (csrc/painn.cu).  slot 'G' with
    (Gs,Gv) absent is the layer's forward (ds, dv) of painn_conv.py:104-114.  Because M is linear in every operand,
    the derivative of any output with respect to any other operand is again such a partial derivative with the
    cotangent substituted for the differentiated operand: the class is closed under differentiation."""

    @staticmethod
    def forward(ctx, slot, mask, want_a, want_b, gs, gv, phi, v, w, r, recv, send, F_dim, acc_a=None):
        ctx.slot, ctx.mask, ctx.recv, ctx.send, ctx.F_dim = slot, mask, recv, send, F_dim
        ctx.save_for_backward(gs, gv, phi, v, w, r)
        ops = dict(zip(_PM_ORDER, (gs, gv, phi, v, w, r)))
        a, b = raw_painn_msg(slot, mask, want_a, want_b, ops, recv, send, F_dim, acc_a)
        return a, b

    @staticmethod
    def backward(ctx, ca, cb):
        ops = dict(zip(_PM_ORDER, ctx.saved_tensors))
        needs = dict(zip(_PM_ORDER, ctx.needs_input_grad[4:10]))
        if not COORDINATE_HESSIAN:
            needs["r"] = False          # r_hat depends on the coordinates only
        grads = dict.fromkeys(_PM_ORDER)
        xa, xb = _PM_SLOTS[ctx.slot]
        subs = []                       # (active terms, operands with the cotangent substituted, replaced names)
        if ctx.slot == "G":             # Gs and Gv share no term: substitute both at once
            m, o = 0, dict(ops)
            if ca is not None:
                m |= ctx.mask & 1
                o["gs"] = ca.contiguous()
            if cb is not None:
                m |= ctx.mask & 6
                o["gv"] = cb.contiguous()
            if m:
                subs.append((m, o, (xa, xb)))
        else:
            for x, c in ((xa, ca), (xb, cb)):
                m = ctx.mask & _PM_TERMS[x]
                if c is not None and m:
                    o = dict(ops)
                    o[x] = c.contiguous()
                    subs.append((m, o, (x,)))
        for m, o, replaced in subs:
            for slot, (ya, yb) in _PM_SLOTS.items():
                wa = ya not in replaced and needs[ya] and bool(_PM_TERMS[ya] & m)
                wb = yb not in replaced and needs[yb] and bool(_PM_TERMS[yb] & m)
                if not (wa or wb):
                    continue
                # the [E, 3F] filter gradient of a second substitution is summed inside the kernel
                acc = grads["w"] if (slot == "E" and wa and grads["w"] is not None) else None
                ra, rb = _PainnMsg.apply(slot, m, wa, wb, o["gs"], o["gv"], o["phi"], o["v"], o["w"], o["r"],
                                         ctx.recv, ctx.send, ctx.F_dim, acc)
                for y, res, wanted in ((ya, ra, wa), (yb, rb, wb)):
                    if wanted:
                        res = res.reshape(ops[y].shape)
                        if y == "w" and acc is not None:
                            grads[y] = res
                        else:
                            grads[y] = res if grads[y] is None else grads[y] + res
        # the accumulate operand enters its output with derivative 1
        g_acc = ca if (ctx.slot == "E" and ctx.needs_input_grad[-1]) else None
        return (None, None, None, None) + tuple(grads[k] for k in _PM_ORDER) + (None, None, None, g_acc)


def painn_message(phi, v, w, r, recv, send):
    """(ds [N,F], dv [N,3,F]) of PAiNNconv (painn_conv.py:104-114) for phi [N,3F], v [N,3,F], w [E,3F], r [E,3]."""
    F_dim = int(phi.shape[1]) // 3
    return _PainnMsg.apply("G", 7, True, True, None, None, phi, v, w, r, recv, send, F_dim, None)


# ------------------------------------------------------------------------------------------------
# fused edge geometry (a9-a12): d, u, radial basis, cosine cutoff straight from the coordinates
# ------------------------------------------------------------------------------------------------
class GeometrySpec:
    """What one fused geometry call computes.  want: subset of ("d", "u", "rbf", "fc") in that order."""

    def __init__(self, rel_i, rel_j, want, basis=0, R=0, basis_params=(0.0, 0.0, 0.0), frequencies=None, eps=0.0,
                 no_nan=True, cos_cutoff=0.0):
        self.rel_i, self.rel_j = rel_i, rel_j
        self.want = tuple(w for w in ("d", "u", "rbf", "fc") if w in want)
        self.basis, self.R = int(basis), int(R)
        self.basis_params = tuple(float(b) for b in basis_params)
        self.frequencies = frequencies
        self.eps, self.no_nan, self.cos_cutoff = float(eps), bool(no_nan), float(cos_cutoff)

    def shapes(self, E):
        return {"d": (E, 1), "u": (E, 3), "rbf": (E, self.R), "fc": (E, 1)}


def raw_edge_geometry(xyz, spec, mode, cots=None, tangent=None):
    """mode 0 / 2: dict name -> tensor for spec.want; mode 1: g_v [E,3]."""
    import ctypes
    _need_cuda(xyz)
    xyz = _f32c(xyz)
    E = spec.rel_i.n_elements
    dev = xyz.device
    bp = (ctypes.c_float * 3)(*spec.basis_params)
    outs = {k: None for k in ("d", "u", "rbf", "fc")}
    gv = None
    if mode in (1, 3):
        gv = torch.empty((E, 3), dtype=torch.float32, device=dev)
        cots = {k: (None if v is None else _f32c(v)) for k, v in cots.items()}
    else:
        shp = spec.shapes(E)
        for k in spec.want:
            outs[k] = torch.empty(shp[k], dtype=torch.float32, device=dev)
        cots = {}
    tangent = None if tangent is None else _f32c(tangent)
    check(_lib.lib().kgcnn_edge_geometry(
        _ptr(xyz), _ptr(spec.rel_i.ids), _ptr(spec.rel_j.ids), E, mode, spec.basis, spec.R,
        ctypes.cast(bp, ctypes.c_void_p), _ptr(spec.frequencies), spec.eps, 1 if spec.no_nan else 0, spec.cos_cutoff,
        _ptr(cots.get("d")), _ptr(cots.get("u")), _ptr(cots.get("rbf")), _ptr(cots.get("fc")), _ptr(tangent),
        _ptr(outs["d"]), _ptr(outs["u"]), _ptr(outs["rbf"]), _ptr(outs["fc"]), _ptr(gv), _stream()),
        "kgcnn_edge_geometry")
    return gv if mode in (1, 3) else tuple(outs[k] for k in spec.want)


def _geom_composite(xyz, spec):
    """The same quantities with differentiable tensor ops (any derivative order); used for the coordinate Hessian
    terms the fused kernels do not provide, and as the comparand in the tests."""
    import math
    xi, xj = xyz[spec.rel_i.ids.long()], xyz[spec.rel_j.ids.long()]
    v = xi - xj
    d = torch.sqrt(torch.relu(torch.sum(v * v, dim=1, keepdim=True)) + spec.eps)
    out = {}
    if "d" in spec.want:
        out["d"] = d
    if "u" in spec.want:
        inv = torch.where(d == 0, torch.zeros_like(d), 1.0 / torch.where(d == 0, torch.ones_like(d), d)) \
            if spec.no_nan else 1.0 / d
        out["u"] = v * inv
    if "rbf" in spec.want:
        if spec.basis == 1:
            gamma, offset, step = spec.basis_params
            mu = torch.arange(spec.R, dtype=xyz.dtype, device=xyz.device) * step
            out["rbf"] = torch.exp(-gamma * torch.square(d - offset - mu))
        else:
            inv_c, p = spec.basis_params[0], int(spec.basis_params[1])
            x = d * inv_c
            a, b, c = -(p + 1) * (p + 2) / 2, p * (p + 2), -p * (p + 1) / 2
            env = torch.where(x < 1, 1.0 / x + a * x ** (p - 1) + b * x ** p + c * x ** (p + 1), torch.zeros_like(x))
            out["rbf"] = env * torch.sin(spec.frequencies.to(xyz.dtype) * x)
    if "fc" in spec.want:
        rc = spec.cos_cutoff
        out["fc"] = (torch.cos(torch.clamp(d, -rc, rc) * math.pi / rc) + 1) * 0.5
    return tuple(out[k] for k in spec.want)


GEOMETRY_HESSIAN_KERNEL = True    # False: the coordinate-Hessian term through the op-by-op composite (tests, A/B)


def _geom_hessian_term(xyz, spec, cots, hvec):
    """d/dxyz <J(xyz)^T cots, hvec>: the coordinate-Hessian term (only when the caller really differentiates the
    force with respect to the coordinates; force TRAINING never does).  One launch of the geometry kernel in mode 3
    + the two deterministic segment sums; Bessel bases with trainable frequencies included (frequencies constant)."""
    if GEOMETRY_HESSIAN_KERNEL:
        gv = raw_edge_geometry(xyz, spec, 3, cots=dict(zip(spec.want, cots)), tangent=hvec)
        return raw_gms(gv, None, spec.rel_i, None) - raw_gms(gv, None, spec.rel_j, None)
    with torch.enable_grad():
        x = xyz.detach().requires_grad_(True)
        outs = _geom_composite(x, spec)
        pairs = [(o, c) for o, c in zip(outs, cots) if c is not None]
        (vjp,) = torch.autograd.grad([o for o, _ in pairs], x, [c for _, c in pairs], create_graph=True)
        (g,) = torch.autograd.grad(vjp, x, hvec, allow_unused=True)
    return torch.zeros_like(xyz) if g is None else g


def _geom_vjp_raw(xyz, cots, spec):
    gv = raw_edge_geometry(xyz, spec, 1, cots=dict(zip(spec.want, cots)))
    return raw_gms(gv, None, spec.rel_i, None) - raw_gms(gv, None, spec.rel_j, None)


def raw_bessel_frequency_grad(xyz, spec, cot_rbf, tangent=None):
    """d(scalar)/d(frequencies) [R] of the Bessel basis: through the basis values (tangent None) or through their
    coordinate VJP contracted with `tangent` [N,3] (include/kgcnn_b200.h, kgcnn_bessel_frequency_grad)."""
    import ctypes
    _need_cuda(xyz, cot_rbf)
    xyz, cot_rbf = _f32c(xyz), _f32c(cot_rbf)
    tangent = None if tangent is None else _f32c(tangent)
    out = torch.empty((spec.R,), dtype=torch.float32, device=xyz.device)
    nbytes = int(_lib.lib().kgcnn_bessel_frequency_grad_workspace_bytes(spec.R))
    ws = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device=xyz.device)
    bp = (ctypes.c_float * 3)(*spec.basis_params)
    check(_lib.lib().kgcnn_bessel_frequency_grad(
        _ptr(xyz), _ptr(spec.rel_i.ids), _ptr(spec.rel_j.ids), spec.rel_i.n_elements, spec.R,
        ctypes.cast(bp, ctypes.c_void_p), _ptr(spec.frequencies), spec.eps, _ptr(cot_rbf), _ptr(tangent), _ptr(out),
        _ptr(ws), nbytes, _stream()), "kgcnn_bessel_frequency_grad")
    return out


def _rbf_cotangent(spec, cots):
    return dict(zip(spec.want, cots)).get("rbf")


class _EdgeGeom(torch.autograd.Function):
    """freq: the trainable Bessel frequencies (a leaf Parameter whose storage spec.frequencies views) or None."""

    @staticmethod
    def forward(ctx, xyz, spec, freq=None):
        ctx.spec = spec
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(xyz, freq)
        return raw_edge_geometry(xyz, spec, 0)

    @staticmethod
    def backward(ctx, *cots):
        xyz, freq = ctx.saved_tensors
        cots = [None if c is None else c.contiguous() for c in cots]
        g_xyz = g_f = None
        # inside the weight sweep of a training step (no_coordinate_hessian) nobody reads d/d(coordinates)
        if ctx.needs_input_grad[0] and (COORDINATE_HESSIAN or _INPUT_ONLY is not None or torch.is_grad_enabled()):
            g_xyz = _EdgeGeomVjp.apply(xyz, ctx.spec, freq, *cots)
        c_rbf = _rbf_cotangent(ctx.spec, cots)
        if freq is not None and c_rbf is not None and _want(ctx, 2, freq):
            g_f = raw_bessel_frequency_grad(xyz.detach(), ctx.spec, c_rbf.detach())
        return g_xyz, None, g_f


class _EdgeGeomVjp(torch.autograd.Function):
    """J^T (cotangents): linear in the cotangents, so its VJP with respect to them is the JVP kernel."""

    @staticmethod
    def forward(ctx, xyz, spec, freq, *cots):
        ctx.spec = spec
        ctx.present = [c is not None for c in cots]
        ctx.save_for_backward(xyz, freq, *[c for c in cots if c is not None])
        return _geom_vjp_raw(xyz, cots, spec)

    @staticmethod
    def backward(ctx, hvec):
        xyz, freq, saved = ctx.saved_tensors[0], ctx.saved_tensors[1], list(ctx.saved_tensors[2:])
        cots = [saved.pop(0) if p else None for p in ctx.present]
        g_xyz = g_f = None
        if hvec is not None and ctx.needs_input_grad[0] and COORDINATE_HESSIAN:
            g_xyz = _geom_hessian_term(xyz, ctx.spec, [None if c is None else c.detach() for c in cots],
                                       hvec.detach())
        c_rbf = _rbf_cotangent(ctx.spec, cots)
        if hvec is not None and freq is not None and c_rbf is not None and _want(ctx, 2, freq):
            # the force depends on the frequencies through d rbf / d d: mixed second derivative, summed over the edges
            g_f = raw_bessel_frequency_grad(xyz.detach(), ctx.spec, c_rbf.detach(), tangent=hvec.detach())
        g_cots = [None] * len(ctx.present)
        if hvec is not None and any(ctx.needs_input_grad[3:]):
            tang = _EdgeGeomJvp.apply(xyz, ctx.spec, hvec.contiguous())
            g_cots = [t if (p and n) else None for t, p, n in zip(tang, ctx.present, ctx.needs_input_grad[3:])]
        return (g_xyz, None, g_f) + tuple(g_cots)


class _EdgeGeomJvp(torch.autograd.Function):
    """J h; its VJP with respect to h is J^T again."""

    @staticmethod
    def forward(ctx, xyz, spec, hvec):
        ctx.spec = spec
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(xyz)
        return raw_edge_geometry(xyz, spec, 2, tangent=hvec)

    @staticmethod
    def backward(ctx, *cots):
        (xyz,) = ctx.saved_tensors
        g_xyz = _no_hessian("edge geometry (third order)") if ctx.needs_input_grad[0] else None
        g_h = None
        if ctx.needs_input_grad[2]:
            g_h = _EdgeGeomVjp.apply(xyz, ctx.spec, None, *[None if c is None else c.contiguous() for c in cots])
        return g_xyz, None, g_h


def edge_geometry(xyz, spec, trainable_frequencies=None):
    """Tuple of tensors for spec.want, differentiable to the order force training needs.  trainable_frequencies: the
    Bessel layer's frequency Parameter (spec.frequencies must be its detached view); it then receives its gradient
    through the energy AND through the forces."""
    return _EdgeGeom.apply(xyz, spec, trainable_frequencies)
