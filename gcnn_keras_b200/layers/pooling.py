"""Pooling layers of kgcnn/layers/pooling.py backed by the deterministic CSR segment-reduce kernel.

PoolingLocalEdges in the reference = partition_row_indexing + stable argsort + 2 permutation gathers +
sorted segment op + scatter_nd zero padding (pooling.py:57-78); here it is ONE kernel over the cached
receiver CSR.  Empty segments give 0 for every method (incl. max/min), as in the reference.
"""
import torch

from .. import functional as F
from ..plan import Relation, get_plan
from .base import GraphBaseLayer, register_keras_serializable


@register_keras_serializable(package="kgcnn", name="PoolingLocalEdges")
class PoolingLocalEdges(GraphBaseLayer):
    """kgcnn/layers/pooling.py:11-88.  call([nodes, edges, tensor_index]) -> ragged [B, [N], F]."""

    def __init__(self, pooling_method="mean", pooling_index=0, **kwargs):
        super().__init__(**kwargs)
        self.pooling_method = pooling_method
        self.pooling_index = pooling_index

    def call(self, inputs, **kwargs):
        nodes, edges, index = self.assert_ragged_input_rank(inputs)
        code = F.op_code(self.pooling_method)                     # TypeError on unknown names (segment.py:51)
        rel = get_plan(index, nodes).relation(self.pooling_index)
        return nodes.with_values(F.segment_reduce(edges.values, rel, code))

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method, "pooling_index": self.pooling_index})
        return config


PoolingLocalMessages = PoolingLocalEdges


@register_keras_serializable(package="kgcnn", name="PoolingWeightedLocalEdges")
class PoolingWeightedLocalEdges(GraphBaseLayer):
    """kgcnn/layers/pooling.py:92-182.  call([nodes, edges, tensor_index, weights])."""

    def __init__(self, pooling_method="mean", normalize_by_weights=False, pooling_index=0, **kwargs):
        super().__init__(**kwargs)
        self.pooling_method = pooling_method
        self.normalize_by_weights = normalize_by_weights
        self.pooling_index = pooling_index

    def call(self, inputs, **kwargs):
        nodes, edges, index, weights = self.assert_ragged_input_rank(inputs)
        code = F.op_code(self.pooling_method)
        rel = get_plan(index, nodes).relation(self.pooling_index)
        w = weights.values
        if code in (F.OP_SUM, F.OP_MEAN) and w.dim() == 2 and edges.values.dim() == 2 \
                and w.shape[1] in (1, edges.values.shape[1]):
            # edge * weight fused into the reduction (pooling.py:152 + :162)
            out = F._GMS_edge.apply(edges.values, w, rel)
            if code == F.OP_MEAN:
                out = out * rel.inv_count()
        else:
            out = F.segment_reduce(edges.values * w, rel, code)
        if self.normalize_by_weights:                             # pooling.py:164-165
            wsum = F.segment_reduce(w, rel, F.OP_SUM)
            safe = torch.where(wsum == 0, torch.ones_like(wsum), wsum)
            out = torch.where(wsum == 0, torch.zeros_like(out), out / safe)
        return nodes.with_values(out)

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method, "normalize_by_weights": self.normalize_by_weights,
                       "pooling_index": self.pooling_index})
        return config


PoolingWeightedLocalMessages = PoolingWeightedLocalEdges


@register_keras_serializable(package="kgcnn", name="PoolingEmbedding")
class PoolingEmbedding(GraphBaseLayer):
    """kgcnn/layers/pooling.py:186-229: per-graph pooling -> dense [B, F] tensor."""

    def __init__(self, pooling_method="mean", **kwargs):
        super().__init__(**kwargs)
        self.pooling_method = pooling_method

    def call(self, inputs, **kwargs):
        inputs = self.assert_ragged_input_rank(inputs)
        code = F.op_code(self.pooling_method)
        return F.segment_reduce(inputs.values, inputs.partition.relation(), code)

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method})
        return config


PoolingNodes = PoolingEmbedding
PoolingGlobalEdges = PoolingEmbedding


@register_keras_serializable(package="kgcnn", name="PoolingWeightedEmbedding")
class PoolingWeightedEmbedding(GraphBaseLayer):
    """kgcnn/layers/pooling.py:233-284."""

    def __init__(self, pooling_method="mean", **kwargs):
        super().__init__(**kwargs)
        self.pooling_method = pooling_method

    def call(self, inputs, **kwargs):
        nodes, weights = self.assert_ragged_input_rank(inputs)
        code = F.op_code(self.pooling_method)
        return F.segment_reduce(nodes.values * weights.values, nodes.partition.relation(), code)

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method})
        return config


PoolingWeightedNodes = PoolingWeightedEmbedding
PoolingWeightedGlobalEdges = PoolingWeightedEmbedding


@register_keras_serializable(package="kgcnn", name="PoolingLocalEdgesAttention")
class PoolingLocalEdgesAttention(GraphBaseLayer):
    """kgcnn/layers/pooling.py:464-546.  call([nodes, edges, attention, tensor_index]):
    n_i = sum_j softmax_j(a_ij) e_ij.  Two launches: fused segment softmax (original edge order) and the fused
    edge*alpha -> CSR segment sum; the reference runs argsort, 3 permutation gathers, segment_max, 2 gathers, exp,
    2 segment_sums and a scatter."""

    def __init__(self, pooling_index=0, **kwargs):
        super().__init__(**kwargs)
        self.pooling_index = pooling_index

    def call(self, inputs, **kwargs):
        nodes, edges, attention, index = self.assert_ragged_input_rank(inputs)
        rel = get_plan(index, nodes).relation(self.pooling_index)
        alpha = F.segment_softmax(attention.values, rel)
        e = edges.values
        if e.dim() == 2 and alpha.dim() == 2 and alpha.shape[1] in (1, e.shape[1]):
            out = F._GMS_edge.apply(e, alpha, rel)
        else:
            out = F.segment_reduce(e * alpha, rel, F.OP_SUM)
        return nodes.with_values(out)

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_index": self.pooling_index})
        return config


PoolingLocalMessagesAttention = PoolingLocalEdgesAttention


def _snake(name):
    """'GlorotUniform' / 'glorot_uniform' / 'Zeros' -> 'glorot_uniform' / 'zeros' (Keras accepts both spellings)."""
    if not isinstance(name, str):
        return name
    out = "".join(("_" + ch.lower()) if ch.isupper() else ch for ch in name)
    return out.lstrip("_")


@register_keras_serializable(package="kgcnn", name="PoolingLocalEdgesLSTM")
class PoolingLocalEdgesLSTM(GraphBaseLayer):
    """kgcnn/layers/pooling.py:288-460.  call([nodes, edges, tensor_index]) -> the LAST hidden state of an LSTM run
    over the edges of every node in (stable) receiver order; nodes without edges keep the zero initial state.

    The reference sorts the edges, builds a ragged [N, (deg), F] tensor and hands it to ``ks.layers.LSTM``
    (pooling.py:417-437), whose cell is  z = x W + h U + b,  (i, f, g, o) = split(z),
    c' = sigma(f) c + sigma(i) tanh(g),  h' = sigma(o) tanh(c')  (Keras LSTMCell, gate order i, f, c, o), with the
    state frozen once a sequence has ended.  Here the input projection of ALL edges is one tcgen05 Dense, the rows are
    brought into CSR order by the permutation kernel and padded to [N, max_deg, 4 units] by ``kgcnn_pad``; each time
    step is then one recurrent Dense [N, units] x [units, 4 units] plus the gate arithmetic on the rows still active.
    Supported: the configuration the reference's models use (return_sequences / return_state / stateful False, no
    dropout); anything else raises at construction."""

    def __init__(self, units, pooling_method="LSTM", pooling_index=0, activation="tanh",
                 recurrent_activation="sigmoid", use_bias=True, kernel_initializer="glorot_uniform",
                 recurrent_initializer="orthogonal", bias_initializer="zeros", unit_forget_bias=True,
                 kernel_regularizer=None, recurrent_regularizer=None, bias_regularizer=None,
                 activity_regularizer=None, kernel_constraint=None, recurrent_constraint=None, bias_constraint=None,
                 dropout=0.0, recurrent_dropout=0.0, return_sequences=False, return_state=False, go_backwards=False,
                 stateful=False, time_major=False, unroll=False, **kwargs):
        super().__init__(**kwargs)
        if return_sequences or return_state or stateful:
            raise ValueError("PoolingLocalEdgesLSTM: return_sequences / return_state / stateful are not supported "
                             "(the pooled output is the last hidden state per node)")
        if dropout or recurrent_dropout:
            raise ValueError("PoolingLocalEdgesLSTM: dropout inside the recurrence is not supported")
        # the variables are always glorot-uniform / orthogonal / zeros and unconstrained: a config that asks for
        # anything else is refused instead of being echoed by get_config and ignored
        for key, given, default in (("kernel_initializer", kernel_initializer, "glorot_uniform"),
                                    ("recurrent_initializer", recurrent_initializer, "orthogonal"),
                                    ("bias_initializer", bias_initializer, "zeros")):
            name = given.get("class_name") if isinstance(given, dict) else given
            if _snake(name) != default:
                raise ValueError("PoolingLocalEdgesLSTM: %s=%r is not supported (only %r)" % (key, given, default))
        for key, given in (("kernel_constraint", kernel_constraint), ("recurrent_constraint", recurrent_constraint),
                           ("bias_constraint", bias_constraint)):
            if given is not None:
                raise ValueError("PoolingLocalEdgesLSTM: %s is not supported" % key)
        self.units = int(units)
        self.pooling_method = pooling_method
        self.pooling_index = pooling_index
        self.use_bias, self.unit_forget_bias, self.go_backwards = use_bias, unit_forget_bias, go_backwards
        self._cfg = {"activation": activation, "recurrent_activation": recurrent_activation, "use_bias": use_bias,
                     "kernel_initializer": kernel_initializer, "recurrent_initializer": recurrent_initializer,
                     "bias_initializer": bias_initializer, "unit_forget_bias": unit_forget_bias,
                     "kernel_regularizer": kernel_regularizer, "recurrent_regularizer": recurrent_regularizer,
                     "bias_regularizer": bias_regularizer, "activity_regularizer": activity_regularizer,
                     "kernel_constraint": kernel_constraint, "recurrent_constraint": recurrent_constraint,
                     "bias_constraint": bias_constraint, "dropout": dropout, "recurrent_dropout": recurrent_dropout,
                     "implementation": 2, "return_sequences": return_sequences, "return_state": return_state,
                     "go_backwards": go_backwards, "stateful": stateful, "time_major": time_major, "unroll": unroll}
        self._act, self._rec_act = self._activation(activation), self._activation(recurrent_activation)
        self.kernel = self.recurrent_kernel = self.bias = None
        if self.pooling_method not in ["LSTM", "lstm"]:
            print("Warning: Pooling method does not match with layer, expected 'LSTM' but got", self.pooling_method)

    @staticmethod
    def _activation(name):
        table = {"tanh": torch.tanh, "sigmoid": torch.sigmoid, "relu": torch.relu, "linear": lambda x: x,
                 None: lambda x: x, "hard_sigmoid": lambda x: torch.clamp(0.2 * x + 0.5, 0.0, 1.0)}
        if name not in table:
            raise ValueError("PoolingLocalEdgesLSTM: unsupported activation %r" % (name,))
        return table[name]

    def build(self, input_shape):
        super().build(input_shape)
        from .modules import _glorot_uniform_
        in_dim, u = int(input_shape[-1]), self.units
        self.kernel = torch.nn.Parameter(_glorot_uniform_(torch.empty(in_dim, 4 * u)))
        self.recurrent_kernel = torch.nn.Parameter(torch.cat(
            [torch.nn.init.orthogonal_(torch.empty(u, u)) for _ in range(4)], dim=1))
        if self.use_bias:
            b = torch.zeros(4 * u)
            if self.unit_forget_bias:
                b[u:2 * u] = 1.0
            self.bias = torch.nn.Parameter(b)

    def forward(self, inputs, **kwargs):
        if not self.built:
            v = inputs[1].values
            self.build(tuple(v.shape))
            self.to(v.device)
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        nodes, edges, index = self.assert_ragged_input_rank(inputs)
        rel = get_plan(index, nodes).relation(self.pooling_index)
        u, n = self.units, rel.n_targets
        dev = edges.values.device
        h = torch.zeros((n, u), dtype=torch.float32, device=dev)
        if rel.n_elements == 0 or n == 0:
            return nodes.with_values(h)
        zx = F.dense(edges.values, self.kernel, self.bias, "linear")            # [E, 4u]: x W + b for every edge
        perm = rel.perm
        if perm is not None:                                                   # stable receiver order (pooling.py:419-423)
            inv = torch.empty_like(perm)
            inv[perm.long()] = torch.arange(perm.shape[0], dtype=perm.dtype, device=dev)
            zx = F.permute_rows(zx, perm, inv)
        rowptr = rel.rowptr.to(torch.int64)
        deg = rowptr[1:] - rowptr[:-1]
        max_len = rel.max_degree()                                             # host int cached with the plan
        zp = F.pad(zx, rowptr, max_len)                                        # [N, max_len, 4u]
        c = torch.zeros_like(h)
        for step in range(max_len):
            active = (deg > step).unsqueeze(1)
            if self.go_backwards:                                              # position deg - 1 - step of every row
                pos = (deg - 1 - step).clamp_min(0).view(n, 1, 1).expand(n, 1, 4 * u)
                z_in = torch.gather(zp, 1, pos).squeeze(1)
            else:
                z_in = zp[:, step]
            z = z_in + F.matmul(h, self.recurrent_kernel)
            zi, zf, zg, zo = z[:, :u], z[:, u:2 * u], z[:, 2 * u:3 * u], z[:, 3 * u:]
            c_new = self._rec_act(zf) * c + self._rec_act(zi) * self._act(zg)
            h_new = self._rec_act(zo) * self._act(c_new)
            c = torch.where(active, c_new, c)
            h = torch.where(active, h_new, h)
        return nodes.with_values(h)

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method, "pooling_index": self.pooling_index, "units": self.units})
        config.update(self._cfg)
        return config


PoolingLocalMessagesLSTM = PoolingLocalEdgesLSTM


@register_keras_serializable(package="kgcnn", name="PoolingEmbeddingAttention")
class PoolingEmbeddingAttention(GraphBaseLayer):
    """kgcnn/layers/pooling.py:550-599.  call([nodes, attention]) -> dense [B, F]: s = sum_i softmax_i(a_i) n_i."""

    def call(self, inputs, **kwargs):
        nodes, attention = self.assert_ragged_input_rank(inputs)
        rel = nodes.partition.relation()
        alpha = F.segment_softmax(attention.values, rel)
        x = nodes.values
        if x.dim() == 2 and alpha.dim() == 2 and alpha.shape[1] in (1, x.shape[1]):
            return F._GMS_edge.apply(x, alpha, rel)
        return F.segment_reduce(x * alpha, rel, F.OP_SUM)


PoolingNodesAttention = PoolingEmbeddingAttention


@register_keras_serializable(package="kgcnn", name="RelationalPoolingLocalEdges")
class RelationalPoolingLocalEdges(GraphBaseLayer):
    """kgcnn/layers/pooling.py:603-675.  The reference scatters with tensor_scatter_nd_add on
    (receiver, relation) (atomic on GPU TF); here the composite key receiver*R + relation gets its own CSR
    and the same deterministic segment kernel runs.  call([nodes, edges, tensor_index, edge_relation])."""

    def __init__(self, num_relations, pooling_method="sum", pooling_index=0, **kwargs):
        super().__init__(**kwargs)
        self.num_relations = num_relations
        self.pooling_method = pooling_method
        self.pooling_index = pooling_index

    def call(self, inputs, **kwargs):
        nodes, edges, index, relation = self.assert_ragged_input_rank(inputs)
        if self.pooling_method not in ["segment_sum", "sum", "reduce_sum", "add", "segment_max", "max",
                                       "reduce_max", "segment_min", "min", "reduce_min"]:
            raise TypeError("Unknown pooling, choose: 'mean', 'sum', ...")   # scatter.py:25
        code = F.OP_SUM if self.pooling_method == "add" else F.op_code(self.pooling_method)
        plan = get_plan(index, nodes)
        recv = plan.relation(self.pooling_index)
        R = int(self.num_relations)
        key = (recv.ids.to(torch.int64) * R + relation.values.to(torch.int64)).to(torch.int32)
        rel = Relation(key, plan.N * R)
        out = F.segment_reduce(edges.values, rel, code)
        if code == F.OP_MAX:      # tensor_scatter_nd_max into zeros (scatter.py:21): max(0, max_e)
            out = torch.clamp_min(out, 0.0)
        elif code == F.OP_MIN:
            out = torch.clamp_max(out, 0.0)
        return nodes.with_values(out.reshape((plan.N, R) + tuple(edges.values.shape[1:])))

    def get_config(self):
        config = super().get_config()
        config.update({"pooling_method": self.pooling_method, "pooling_index": self.pooling_index,
                       "num_relations": self.num_relations})
        return config
