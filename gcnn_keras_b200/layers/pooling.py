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
