"""SchNetCFconv / SchNetInteraction of kgcnn/layers/conv/schnet_conv.py.

The reference cfconv is Dense -> Dense -> GatherNodesOutgoing -> LazyMultiply -> PoolingLocalEdges
(schnet_conv.py:73-79), i.e. four HBM passes over [E, units] plus a sort; here gather * filter -> segment-sum
is ONE kernel over the cached receiver CSR, and its VJPs are the same kernel on the sender CSC plus the
two-sided gather-product (closed under differentiation for force training).
"""
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
        x = self.lay_dense1(edge, **kwargs)                                   # schnet_conv.py:74
        x = self.lay_dense2(x, **kwargs)                                      # :75
        code = F.op_code(self.cfconv_pool)
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
