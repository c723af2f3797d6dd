"""Oracle restatement of the kgcnn L2 layers on the hot path (torch CPU, dtype follows the inputs).
Each function cites the reference lines it follows.  Test infrastructure only."""
import math

import numpy as np
import torch

from . import ops
from .ragged import R

KERAS_EPSILON = 1e-7  # ks.backend.epsilon()


def _batch_index(index, nodes):
    # the call every gather / pooling layer makes: gather.py:78-82, pooling.py:57-61
    return ops.partition_row_indexing(index.values, nodes.row_splits, index.row_lengths(),
                                      partition_type_target="row_splits", partition_type_index="row_length",
                                      to_indexing="batch", from_indexing="sample")


# ---- kgcnn/layers/gather.py ---------------------------------------------------------------------------
def gather_embedding(nodes, index, concat_axis=2, split_axis=None):
    """GatherEmbedding._disjoint_implementation, gather.py:69-99."""
    if split_axis is not None and concat_axis is not None:
        raise ValueError("Can not both split and concatenate new index axis. At least one must be `None`.")
    disjoint = _batch_index(index, nodes)                      # :78-82
    out = nodes.values[disjoint]                               # :83  [E, K, F...]
    if concat_axis == 2:                                       # :85-90
        out = torch.cat([out[:, i] for i in range(disjoint.shape[1])], dim=1)
        return index.with_values(out)
    if split_axis == 2:                                        # :92-97
        return [index.with_values(out[:, i]) for i in range(disjoint.shape[1])]
    return index.with_values(out)


def gather_embedding_selection(nodes, index, selection_index):
    """GatherEmbeddingSelection.call, gather.py:218-230."""
    if isinstance(selection_index, int):
        selection_index = [selection_index]
    indexlist = _batch_index(index, nodes)                     # :223-227
    return [index.with_values(nodes.values[indexlist[:, i]]) for i in selection_index]   # :228-229


def gather_nodes_ingoing(nodes, index):
    return gather_embedding_selection(nodes, index, [0])[0]    # gather.py:249-282


def gather_nodes_outgoing(nodes, index):
    return gather_embedding_selection(nodes, index, [1])[0]    # gather.py:286-319


def gather_state(state, target):
    """GatherState.call, gather.py:357-370."""
    return target.with_values(torch.repeat_interleave(state, target.row_lengths(), dim=0))


# ---- kgcnn/layers/pooling.py --------------------------------------------------------------------------
def _pad_rows(out, n):
    # tf.scatter_nd(range(len(out)), out, [N, ...]) : pooling.py:74-76
    if out.shape[0] == n:
        return out
    pad = torch.zeros((n - out.shape[0],) + tuple(out.shape[1:]), dtype=out.dtype)
    return torch.cat([out, pad], dim=0)


def pooling_local_edges(nodes, edges, index, pooling_method="mean", pooling_index=0, is_sorted=False,
                        has_unconnected=True):
    """PoolingLocalEdges.call, pooling.py:37-79."""
    shiftind = _batch_index(index, nodes)                      # :57-61
    nodind = shiftind[:, pooling_index]                        # :63
    dens = edges.values
    if not is_sorted:                                          # :65-68
        node_order = torch.argsort(nodind, stable=True)
        nodind = nodind[node_order]
        dens = dens[node_order]
    out = ops.segment_ops_by_name(pooling_method, dens, nodind)   # :71
    if has_unconnected:                                        # :74-76
        out = _pad_rows(out, nodes.values.shape[0])
    return nodes.with_values(out)                              # :78


def pooling_weighted_local_edges(nodes, edges, index, weights, pooling_method="mean", normalize_by_weights=False,
                                 pooling_index=0, is_sorted=False, has_unconnected=True):
    """PoolingWeightedLocalEdges.call, pooling.py:125-172."""
    shiftind = _batch_index(index, nodes)
    wval = weights.values
    dens = edges.values * wval                                 # :152
    nodind = shiftind[:, pooling_index]
    if not is_sorted:                                          # :155-159
        node_order = torch.argsort(nodind, stable=True)
        nodind, dens, wval = nodind[node_order], dens[node_order], wval[node_order]
    get = ops.segment_ops_by_name(pooling_method, dens, nodind)   # :162
    if normalize_by_weights:                                   # :164-165
        get = ops.divide_no_nan(get, ops.segment_sum(wval, nodind))
    if has_unconnected:
        get = _pad_rows(get, nodes.values.shape[0])
    return nodes.with_values(get)


def pooling_local_edges_attention(nodes, edges, attention, index, pooling_index=0, is_sorted=False,
                                  has_unconnected=True):
    """PoolingLocalEdgesAttention.call, pooling.py:505-540."""
    shiftind = _batch_index(index, nodes)
    nodind = shiftind[:, pooling_index]
    dens, ats = edges.values, attention.values
    if not is_sorted:                                          # :522-527
        node_order = torch.argsort(nodind, stable=True)
        nodind, dens, ats = nodind[node_order], dens[node_order], ats[node_order]
    ats = ops.segment_softmax(ats, nodind)                     # :529
    get = ops.segment_sum(dens * ats, nodind)                  # :530-531
    if has_unconnected:
        get = _pad_rows(get, nodes.values.shape[0])
    return nodes.with_values(get)


def pooling_local_edges_lstm(nodes, edges, index, kernel, recurrent_kernel, bias=None, pooling_index=0,
                             is_sorted=False, go_backwards=False):
    """PoolingLocalEdgesLSTM.call, pooling.py:398-446: stable sort by receiver (:419-423), ragged [N, (deg), F]
    (:427), ks.layers.LSTM -> last hidden state (:428), zero rows for nodes without edges (:430-434).
    The cell is Keras' LSTMCell (third party, TensorFlow 2.12 keras/layers/rnn/lstm.py): z = x W + h U + b,
    gate order (i, f, c, o), c' = sigmoid(f) c + sigmoid(i) tanh(g), h' = sigmoid(o) tanh(c').  PARITY UNPINNED:
    the reference holds no numeric test for this layer.  Plain loop per node: small cases only."""
    shiftind = _batch_index(index, nodes)
    nodind = shiftind[:, pooling_index]
    dens = edges.values
    if not is_sorted:
        node_order = torch.argsort(nodind, stable=True)
        nodind, dens = nodind[node_order], dens[node_order]
    n, u = nodes.values.shape[0], recurrent_kernel.shape[0]
    rows = []
    for i in range(n):
        seq = dens[nodind == i]
        if go_backwards:
            seq = torch.flip(seq, dims=[0])
        h = torch.zeros(u, dtype=dens.dtype)
        c = torch.zeros(u, dtype=dens.dtype)
        for x in seq:
            z = x @ kernel + h @ recurrent_kernel
            if bias is not None:
                z = z + bias
            zi, zf, zg, zo = z[:u], z[u:2 * u], z[2 * u:3 * u], z[3 * u:]
            c = torch.sigmoid(zf) * c + torch.sigmoid(zi) * torch.tanh(zg)
            h = torch.sigmoid(zo) * torch.tanh(c)
        rows.append(h)
    return nodes.with_values(torch.stack(rows) if rows else torch.zeros((0, u), dtype=dens.dtype))


def pooling_embedding_attention(nodes, attention):
    """PoolingEmbeddingAttention.call (= PoolingNodesAttention), pooling.py:583-589: dense [B, F]."""
    batchi = nodes.value_rowids()
    ats = ops.segment_softmax(attention.values, batchi)
    return ops.segment_sum(nodes.values * ats, batchi)


def pooling_embedding(inputs, pooling_method="mean"):
    """PoolingEmbedding.call (= PoolingNodes), pooling.py:203-219.  Returns a dense [B, F] tensor.
    Trailing empty graphs are dropped by TF's sorted segment op; the product pads them with 0."""
    return ops.segment_ops_by_name(pooling_method, inputs.values, inputs.value_rowids())


def pooling_weighted_embedding(inputs, weights, pooling_method="mean"):
    """PoolingWeightedEmbedding.call, pooling.py:257-278."""
    return ops.segment_ops_by_name(pooling_method, inputs.values * weights.values, inputs.value_rowids())


def relational_pooling_local_edges(nodes, edges, index, relation, num_relations, pooling_method="sum",
                                   pooling_index=0):
    """RelationalPoolingLocalEdges.call, pooling.py:630-668."""
    shiftind = _batch_index(index, nodes)                      # :649-653
    indices = shiftind[:, pooling_index]                       # :655
    relations = relation.values.to(indices.dtype)              # :656
    scatter_indices = torch.stack([indices, relations], dim=-1)   # :657
    out_tensor = torch.zeros((nodes.values.shape[0], num_relations) + tuple(edges.values.shape[1:]),
                             dtype=edges.values.dtype)         # :658-660
    out = ops.tensor_scatter_nd_ops_by_name(pooling_method, out_tensor, scatter_indices, edges.values)   # :663
    return nodes.with_values(out)


# ---- kgcnn/layers/geom.py -----------------------------------------------------------------------------
def node_position(position, index, selection_index=(0, 1)):
    """NodePosition.call, geom.py:54-67."""
    return gather_embedding_selection(position, index, list(selection_index))


def euclidean_norm(x, axis=-1, keepdims=False, invert_norm=False, add_eps=False, no_nan=True, square_norm=False):
    """EuclideanNorm._compute_euclidean_norm, geom.py:165-193 (axis counts on the values tensor)."""
    out = torch.relu(torch.sum(torch.square(x), dim=axis, keepdim=keepdims))   # :181
    if add_eps:
        out = out + KERAS_EPSILON                              # :184-185
    if not square_norm:
        out = torch.sqrt(out)                                  # :186-187
    if invert_norm:
        out = ops.divide_no_nan(torch.ones_like(out), out) if no_nan else 1 / out   # :188-192
    return out


def node_distance_euclidean(pos_a, pos_b, add_eps=False, no_nan=True):
    """NodeDistanceEuclidean.call, geom.py:306-321 -> [E, 1]."""
    diff = pos_a.values - pos_b.values
    return pos_a.with_values(euclidean_norm(diff, axis=1, keepdims=True, add_eps=add_eps, no_nan=no_nan))


def edge_direction_normalized(pos_a, pos_b, add_eps=False, no_nan=True):
    """EdgeDirectionNormalized.call, geom.py:357-371."""
    diff = pos_a.values - pos_b.values
    norm = euclidean_norm(diff, axis=1, keepdims=True, invert_norm=True, add_eps=add_eps, no_nan=no_nan)
    return pos_a.with_values(diff * norm)


def scalar_product(a, b, axis=1):
    """ScalarProduct, geom.py:218-281: sum(a*b, axis) on the values tensor."""
    return a.with_values(torch.sum(a.values * b.values, dim=axis))


def gauss_basis(distance, bins=20, dist=4.0, sigma=0.4, offset=0.0):
    """GaussBasisLayer._compute_gauss_basis, geom.py:553-571."""
    d = distance.values
    gamma = 1 / sigma / sigma / 2                              # :549
    gbs = torch.arange(0, bins, 1, dtype=d.dtype) / float(bins) * dist   # :567
    out = d - offset                                           # :568
    out = torch.square(out - gbs) * (gamma * (-1.0))           # :569
    return distance.with_values(torch.exp(out))                # :570


def bessel_basis(distance, num_radial, cutoff, envelope_exponent=5, frequencies=None):
    """BesselBasisLayer.expand_bessel_basis + envelope, geom.py:772-785.  inv_cutoff and the initial
    frequencies are float32 constants in the reference (:758, :766-767) and are rounded accordingly."""
    d = distance.values
    inv_cutoff = torch.tensor(np.float32(1 / cutoff), dtype=d.dtype)
    if frequencies is None:
        frequencies = torch.tensor(np.pi * np.arange(1, num_radial + 1, dtype=np.float32), dtype=d.dtype)
    d_scaled = d * inv_cutoff                                  # :782
    p = envelope_exponent + 1                                  # :774
    a = -(p + 1) * (p + 2) / 2
    b = p * (p + 2)
    c = -p * (p + 1) / 2
    env_val = 1.0 / d_scaled + a * d_scaled ** (p - 1) + b * d_scaled ** p + c * d_scaled ** (p + 1)   # :778
    d_cutoff = torch.where(d_scaled < 1, env_val, torch.zeros_like(d_scaled))   # :779
    return distance.with_values(d_cutoff * torch.sin(frequencies * d_scaled))   # :784


def cos_cutoff_envelope(distance, cutoff):
    """CosCutOffEnvelope._compute_cutoff_envelope, geom.py:831-837 (cutoff None -> 1e8, :829)."""
    cutoff = float(abs(cutoff)) if cutoff is not None else 1e8
    fc = torch.clamp(distance.values, -cutoff, cutoff)
    return distance.with_values((torch.cos(fc * np.pi / cutoff) + 1) * 0.5)


def cos_cutoff(distance, cutoff):
    """CosCutOff._compute_cutoff, geom.py:883-888."""
    return distance.with_values(cos_cutoff_envelope(distance, cutoff).values * distance.values)


# ---- kgcnn/layers/modules.py --------------------------------------------------------------------------
def dense(x, kernel, bias=None, activation=None):
    """DenseEmbedding.call, modules.py:75-86: activation(x @ kernel + bias) on the flat values."""
    v = x.values if isinstance(x, R) else x
    out = v @ kernel
    if bias is not None:
        out = out + bias
    out = ops.activation(activation)(out)
    return x.with_values(out) if isinstance(x, R) else out


def ragged_to_tensor(r, default_value=0, max_len=None):
    """tf.RaggedTensor.to_tensor(shape, default_value) as ChangeTensorType uses it (kgcnn/layers/casting.py:78-80):
    [B, max_len, ...], row b holds its values left-aligned, the rest is `default_value`; rows longer than a given
    max_len are truncated.  NumPy-style loop over rows (integer/copy work, bit-exact comparand)."""
    lengths = r.row_lengths()
    nmax = int(lengths.max()) if max_len is None and len(lengths) else int(max_len or 0)
    out = r.values.new_full((r.nrows(), nmax) + tuple(r.values.shape[1:]), default_value)
    for b in range(r.nrows()):
        n = min(int(lengths[b]), nmax)
        out[b, :n] = r[b][:n]
    return out


def change_tensor_type_mask(r, default_value=0, boolean_mask=False, max_len=None):
    """ChangeTensorType(ragged -> 'mask') (kgcnn/layers/casting.py:80-84): (padded, mask) with
    mask = ones_like(flat_values) (or bool ones) padded with zeros."""
    padded = ragged_to_tensor(r, default_value, max_len)
    ones = torch.ones_like(r.values, dtype=torch.bool if boolean_mask else r.values.dtype)
    return padded, ragged_to_tensor(r.with_values(ones), 0, max_len)


def padded_to_ragged(x_pad, x_mask):
    """EnergyForceModel._cast_coordinates_pad_to_ragged (kgcnn/model/force.py:204-211): boolean-mask gather of the
    padded tensor by mask[:, :, 0], row lengths = number of live slots."""
    from .ragged import R
    live = x_mask[:, :, 0].to(torch.bool)
    return R.from_row_lengths(x_pad[live], live.sum(dim=-1).numpy())
