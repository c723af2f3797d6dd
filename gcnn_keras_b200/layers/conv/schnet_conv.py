"""SchNetCFconv / SchNetInteraction of kgcnn/layers/conv/schnet_conv.py.

The reference cfconv is Dense -> Dense -> GatherNodesOutgoing -> LazyMultiply -> PoolingLocalEdges
(schnet_conv.py:73-79), i.e. four HBM passes over [E, units] plus a sort; here gather * filter -> segment-sum
is ONE kernel over the cached receiver CSR, and its VJPs are the same kernel on the sender CSC plus the
two-sided gather-product (closed under differentiation for force training).
"""
import torch

from ... import functional as F
from ...plan import get_plan
from ..base import GraphBaseLayer, register_keras_serializable
from ..gather import GatherNodesOutgoing
from ..modules import Dense, LazyAdd, LazyMultiply
from ..pooling import PoolingLocalEdges

_DENSE_KEYS = ["kernel_regularizer", "activity_regularizer", "bias_regularizer", "kernel_constraint",
               "bias_constraint", "kernel_initializer", "bias_initializer"]


@register_keras_serializable(package="kgcnn", name="SchNetCFconv")
class SchNetCFconv(GraphBaseLayer):
    """kgcnn/layers/conv/schnet_conv.py:9-89.  call([nodes, edges, edge_index]) -> nodes."""

    def __init__(self, units, cfconv_pool="segment_sum", use_bias=True, activation="kgcnn>shifted_softplus",
                 kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None, kernel_constraint=None,
                 bias_constraint=None, kernel_initializer="glorot_uniform", bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.cfconv_pool = cfconv_pool
        self.units = units
        self.use_bias = use_bias
        self.activation = activation
        kernel_args = {"kernel_regularizer": kernel_regularizer, "activity_regularizer": activity_regularizer,
                       "bias_regularizer": bias_regularizer, "kernel_constraint": kernel_constraint,
                       "bias_constraint": bias_constraint, "kernel_initializer": kernel_initializer,
                       "bias_initializer": bias_initializer}
        self._kernel_args = kernel_args
        self.lay_dense1 = Dense(units=self.units, activation=activation, use_bias=self.use_bias, **kernel_args)
        self.lay_dense2 = Dense(units=self.units, activation="linear", use_bias=self.use_bias, **kernel_args)
        self.lay_sum = PoolingLocalEdges(pooling_method=cfconv_pool)
        self.gather_n = GatherNodesOutgoing()
        self.lay_mult = LazyMultiply()

    def call(self, inputs, **kwargs):
        node, edge, indexlist = inputs
        code = F.op_code(self.cfconv_pool)
        fused = self._fused_source(node, edge) if code in (F.OP_SUM, F.OP_MEAN) else None
        if fused is not None:        # :74-78 in ONE kernel: the filter [E, units] never reaches HBM (csrc/cfconv.cu)
            dist, gamma, offset, step, bins = fused
            plan = get_plan(indexlist, node)
            recv, send = plan.relation(self.lay_sum.pooling_index), plan.relation(self.gather_n.selection_index[0])
            spec = F.CfconvSpec(recv, send, bins, gamma, offset, step, self.lay_dense1._act_code)
            out = F.cfconv_fused(node.values, dist.values, self.lay_dense1.kernel, self.lay_dense1.bias,
                                 self.lay_dense2.kernel, self.lay_dense2.bias, spec)
            if code == F.OP_MEAN:
                out = out * recv.inv_count()
            return node.with_values(out)
        x = self.lay_dense1(edge, **kwargs)                                   # schnet_conv.py:74
        x = self.lay_dense2(x, **kwargs)                                      # :75
        if code in (F.OP_SUM, F.OP_MEAN):
            plan = get_plan(indexlist, node)
            recv, send = plan.relation(self.lay_sum.pooling_index), plan.relation(self.gather_n.selection_index[0])
            out = F.gather_mul_segment_sum(node.values, x.values, recv, send)    # :76-78 fused
            if code == F.OP_MEAN:
                out = out * recv.inv_count()
            return node.with_values(out)
        node2exp = self.gather_n([node, indexlist], **kwargs)                 # max / min: unfused chain
        x = self.lay_mult([node2exp, x], **kwargs)
        return self.lay_sum([node, x, indexlist], **kwargs)

    def _fused_source(self, node, edge):
        """(distance ragged, gamma, offset, step, bins) when `edge` is the untouched output of GaussBasisLayer and the
        shapes are the ones csrc/cfconv.cu is instantiated for (units 128, <= 32 bins, shifted softplus)."""
        p = getattr(edge, "provenance", None)
        if p is None or p[0] != "gauss" or self.lay_dense1._act_code is None:
            return None
        dist, gamma, offset, step, bins = p[1:]
        v, dv = node.values, dist.values
        if not (v.is_cuda and v.dtype == torch.float32 and v.dim() == 2 and int(v.shape[1]) == self.units
                and dv.dtype == torch.float32 and dv.numel() == edge.values.shape[0]):
            return None
        for lay, in_dim in ((self.lay_dense1, bins), (self.lay_dense2, self.units)):
            if not lay.built:
                lay.build((None, in_dim))
                lay.to(v.device)
        if not F.CfconvSpec.supported(self.units, bins, self.lay_dense1._act_code, self.lay_dense1.kernel,
                                      self.lay_dense2.kernel):
            return None
        return dist, gamma, offset, step, bins

    def get_config(self):
        config = super().get_config()
        config.update({"cfconv_pool": self.cfconv_pool, "units": self.units, "activation": self.activation,
                       "use_bias": self.use_bias})
        config.update(self._kernel_args)
        return config


@register_keras_serializable(package="kgcnn", name="SchNetInteraction")
class SchNetInteraction(GraphBaseLayer):
    """kgcnn/layers/conv/schnet_conv.py:93-174.  call([nodes, edges, tensor_index]) -> nodes."""

    def __init__(self, units=128, cfconv_pool="sum", use_bias=True, activation="kgcnn>shifted_softplus",
                 kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None, kernel_constraint=None,
                 bias_constraint=None, kernel_initializer="glorot_uniform", bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.cfconv_pool = cfconv_pool
        self.use_bias = use_bias
        self.units = units
        self.activation = activation
        kernel_args = {"kernel_regularizer": kernel_regularizer, "activity_regularizer": activity_regularizer,
                       "bias_regularizer": bias_regularizer, "kernel_constraint": kernel_constraint,
                       "bias_constraint": bias_constraint, "kernel_initializer": kernel_initializer,
                       "bias_initializer": bias_initializer}
        self._kernel_args = kernel_args
        conv_args = {"units": self.units, "use_bias": use_bias, "activation": activation, "cfconv_pool": cfconv_pool}
        self.lay_cfconv = SchNetCFconv(**conv_args, **kernel_args)
        self.lay_dense1 = Dense(units=self.units, activation="linear", use_bias=False, **kernel_args)
        self.lay_dense2 = Dense(units=self.units, activation=activation, use_bias=self.use_bias, **kernel_args)
        self.lay_dense3 = Dense(units=self.units, activation="linear", use_bias=self.use_bias, **kernel_args)
        self.lay_add = LazyAdd()

    def call(self, inputs, **kwargs):
        node, edge, indexlist = inputs
        x = self.lay_dense1(node, **kwargs)                                   # schnet_conv.py:160
        x = self.lay_cfconv([x, edge, indexlist], **kwargs)                   # :161
        x = self.lay_dense2(x, **kwargs)                                      # :162
        x = self.lay_dense3(x, **kwargs)                                      # :163
        return self.lay_add([node, x], **kwargs)                              # :164

    def get_config(self):
        config = super().get_config()
        config.update({"cfconv_pool": self.cfconv_pool, "units": self.units, "use_bias": self.use_bias,
                       "activation": self.activation})
        config.update(self._kernel_args)
        return config
