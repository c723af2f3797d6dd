"""GCN layer of kgcnn/layers/conv/gcn_conv.py:10-103.

Reference: Dense -> GatherNodesOutgoing -> PoolingWeightedLocalEdges -> Activation, i.e. the [E, F] edge tensor is
materialised, multiplied by the edge weight, permuted and reduced (>= 5 HBM passes over [E, F], gcn_conv.py:85-90,
pooling.py:152-172).  Here gather * weight -> segment-sum is ONE kernel; no [E, F] tensor exists.
"""
import torch

from ... import functional as F
from ...plan import get_plan
from ..base import GraphBaseLayer, register_keras_serializable
from ..modules import Activation, Dense


@register_keras_serializable(package="kgcnn", name="GCN")
class GCN(GraphBaseLayer):
    """call([nodes, edges(weights, (batch,[M],1)), edge_index]) -> nodes."""

    def __init__(self, units, pooling_method="sum", normalize_by_weights=False, activation="kgcnn>leaky_relu",
                 use_bias=True, kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None,
                 kernel_constraint=None, bias_constraint=None, kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.normalize_by_weights = normalize_by_weights
        self.pooling_method = pooling_method
        self.units = units
        self.activation = activation
        self.use_bias = use_bias
        self.lay_dense = Dense(units=self.units, activation="linear", use_bias=use_bias,
                               kernel_initializer=kernel_initializer, bias_initializer=bias_initializer)
        self.lay_act = Activation(activation)

    def call(self, inputs, **kwargs):
        node, edges, edge_index = inputs
        code = F.op_code(self.pooling_method)
        if code not in (F.OP_SUM, F.OP_MEAN):
            raise NotImplementedError("GCN with max/min pooling is outside the fused hot path")
        no = self.lay_dense(node, **kwargs)                                   # gcn_conv.py:86
        plan = get_plan(edge_index, node)
        recv, send = plan.relation(0), plan.relation(1)
        w = edges.values
        nu = F.gather_mul_segment_sum(no.values, w, recv, send)               # :87-88 fused
        if code == F.OP_MEAN:
            nu = nu * recv.inv_count()
        if self.normalize_by_weights:                                         # pooling.py:164-165
            wsum = F.segment_reduce(w, recv, F.OP_SUM)
            safe = torch.where(wsum == 0, torch.ones_like(wsum), wsum)
            nu = torch.where(wsum == 0, torch.zeros_like(nu), nu / safe)
        return self.lay_act(node.with_values(nu), **kwargs)                   # :89

    def get_config(self):
        config = super().get_config()
        config.update({"normalize_by_weights": self.normalize_by_weights, "pooling_method": self.pooling_method,
                       "units": self.units, "activation": self.activation, "use_bias": self.use_bias})
        return config
