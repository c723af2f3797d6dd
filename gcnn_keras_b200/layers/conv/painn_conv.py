"""PAiNNconv / PAiNNUpdate / EquivariantInitialize / SplitEmbedding of kgcnn/layers/conv/painn_conv.py.

Gathers and segment sums run through the sm_100a row-gather / CSR segment kernels over the cached plan (the
reference re-derives the batch indices four times and sorts twice per PAiNNconv call, painn_conv.py:100-114).
"""
import torch

from ... import functional as F
from ...plan import get_plan
from ..base import GraphBaseLayer, register_keras_serializable
from ..geom import compute_euclidean_norm
from ..modules import Dense

KERAS_EPSILON = 1e-7
FUSED = True      # False: compose the message from the gather / segment primitives (test comparand)


@register_keras_serializable(package="kgcnn", name="PAiNNconv")
class PAiNNconv(GraphBaseLayer):
    """kgcnn/layers/conv/painn_conv.py:12-125.
    call([nodes, equivariant, rbf, envelope, r_ij, edge_index]) -> (ds, dv)."""

    def __init__(self, units, conv_pool="sum", use_bias=True, activation="swish", cutoff=None,
                 kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None, kernel_constraint=None,
                 bias_constraint=None, kernel_initializer="glorot_uniform", bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.conv_pool = conv_pool
        self.units = units
        self.use_bias = use_bias
        self.cutoff = cutoff
        self.activation = activation
        kernel_args = {"kernel_regularizer": kernel_regularizer, "activity_regularizer": activity_regularizer,
                       "bias_regularizer": bias_regularizer, "kernel_constraint": kernel_constraint,
                       "bias_constraint": bias_constraint, "kernel_initializer": kernel_initializer,
                       "bias_initializer": bias_initializer}
        self._kernel_args = kernel_args
        self.lay_dense1 = Dense(units=self.units, activation=activation, use_bias=self.use_bias, **kernel_args)
        self.lay_phi = Dense(units=self.units * 3, activation="linear", use_bias=self.use_bias, **kernel_args)
        self.lay_w = Dense(units=self.units * 3, activation="linear", use_bias=self.use_bias, **kernel_args)

    def call(self, inputs, **kwargs):
        node, equivariant, rbf, envelope, r_ij, indexlist = inputs
        code = F.op_code(self.conv_pool)
        plan = get_plan(indexlist, node)
        recv, send = plan.relation(0), plan.relation(1)
        s = self.lay_dense1(node)                                            # painn_conv.py:98
        s = self.lay_phi(s)                                                  # :99
        w = self.lay_w(rbf).values                                           # :101
        if self.cutoff is not None:                                          # :102-103
            w = w * envelope.values
        u = self.units
        v = equivariant.values
        if (FUSED and code in (F.OP_SUM, F.OP_MEAN) and u % 4 == 0 and v.dim() == 3 and v.shape[1] == 3 and r_ij.values.shape[-1] == 3):
            # :100-114 as ONE pass over the edges (csrc/painn.cu): no [E,3F] / [E,3,F] intermediate reaches HBM
            ds, dv = F.painn_message(s.values, v, w, r_ij.values, recv, send)
            if code == F.OP_MEAN:
                inv = recv.inv_count()
                ds, dv = ds * inv, dv * inv.unsqueeze(-1)
            return node.with_values(ds), equivariant.with_values(dv)
        s_j = F.gather(s.values, send)                                       # :100
        sw = s_j * w                                                         # :104
        sw1, sw2, sw3 = sw[:, :u], sw[:, u:2 * u], sw[:, 2 * u:]             # :105
        ds = F.segment_reduce(sw1.contiguous(), recv, code)                  # :106
        vj = F.gather(v, send)                                               # :107  [E,3,F]
        dv = sw2.unsqueeze(-2) * vj + sw3.unsqueeze(-2) * r_ij.values.unsqueeze(-1)   # :108-113
        dv = F.segment_reduce(dv, recv, code)                                # :114
        return node.with_values(ds), equivariant.with_values(dv)

    def get_config(self):
        config = super().get_config()
        config.update({"conv_pool": self.conv_pool, "units": self.units, "cutoff": self.cutoff,
                       "activation": self.activation, "use_bias": self.use_bias})
        config.update(self._kernel_args)
        return config


@register_keras_serializable(package="kgcnn", name="PAiNNUpdate")
class PAiNNUpdate(GraphBaseLayer):
    """kgcnn/layers/conv/painn_conv.py:129-224.  call([nodes, equivariant]) -> (ds, dv)."""

    def __init__(self, units, use_bias=True, activation="swish", kernel_regularizer=None, bias_regularizer=None,
                 activity_regularizer=None, kernel_constraint=None, bias_constraint=None,
                 kernel_initializer="glorot_uniform", bias_initializer="zeros", add_eps=False, **kwargs):
        super().__init__(**kwargs)
        self.units = units
        self.use_bias = use_bias
        self.activation = activation
        self.add_eps = add_eps
        kernel_args = {"kernel_regularizer": kernel_regularizer, "activity_regularizer": activity_regularizer,
                       "bias_regularizer": bias_regularizer, "kernel_constraint": kernel_constraint,
                       "bias_constraint": bias_constraint, "kernel_initializer": kernel_initializer,
                       "bias_initializer": bias_initializer}
        self._kernel_args = kernel_args
        self.lay_dense1 = Dense(units=self.units, activation=activation, use_bias=self.use_bias, **kernel_args)
        self.lay_lin_u = Dense(self.units, activation="linear", use_bias=False, **kernel_args)
        self.lay_lin_v = Dense(self.units, activation="linear", use_bias=False, **kernel_args)
        self.lay_a = Dense(units=self.units * 3, activation="linear", use_bias=self.use_bias, **kernel_args)

    def call(self, inputs, **kwargs):
        node, equivariant = inputs
        v_v = self.lay_lin_v(equivariant, **kwargs).values                   # painn_conv.py:202
        v_u = self.lay_lin_u(equivariant, **kwargs).values                   # :203
        v_prod = torch.sum(v_u * v_v, dim=1)                                 # :204 ScalarProduct(axis=2)
        v_norm = compute_euclidean_norm(v_v, axis=1, add_eps=self.add_eps)   # :205 EuclideanNorm(axis=2)
        a = torch.cat([node.values, v_norm], dim=-1)                         # :206
        a = self.lay_dense1(a, **kwargs)                                     # :207
        a = self.lay_a(a, **kwargs)                                          # :208
        u = self.units
        a_vv, a_sv, a_ss = a[:, :u], a[:, u:2 * u], a[:, 2 * u:]             # :209
        dv = a_vv.unsqueeze(-2) * v_u                                        # :210-211
        ds = v_prod * a_sv + a_ss                                            # :212-213
        return node.with_values(ds), equivariant.with_values(dv)

    def get_config(self):
        config = super().get_config()
        config.update({"units": self.units, "activation": self.activation, "use_bias": self.use_bias})
        config.update(self._kernel_args)
        return config


@register_keras_serializable(package="kgcnn", name="EquivariantInitialize")
class EquivariantInitialize(GraphBaseLayer):
    """kgcnn/layers/conv/painn_conv.py:228-297."""

    def __init__(self, dim=3, method="zeros", value=1.0, stddev=1.0, **kwargs):
        super().__init__(**kwargs)
        self.dim = int(dim)
        self.method = str(method)
        self.value = float(value)
        self.stddev = float(stddev)

    def call(self, inputs, **kwargs):
        inputs = self.assert_ragged_input_rank(inputs)
        v = inputs.values
        if self.method == "zeros":
            out = torch.zeros_like(v).unsqueeze(1).repeat(1, self.dim, 1)
        elif self.method == "eps":
            out = (torch.zeros_like(v) + KERAS_EPSILON).unsqueeze(1).repeat(1, self.dim, 1)
        elif self.method == "ones":
            out = torch.ones_like(v).unsqueeze(1).repeat(1, self.dim, 1)
        elif self.method == "eye":
            out = torch.eye(self.dim, v.shape[1], dtype=v.dtype, device=v.device).unsqueeze(0).repeat(v.shape[0], 1, 1)
        elif self.method == "normal":
            out = (torch.randn(self.dim, v.shape[1], dtype=v.dtype, device=v.device) * self.stddev).unsqueeze(0)
            out = out.repeat(v.shape[0], 1, 1)
        elif self.method == "const":
            out = (torch.ones_like(v) * self.value).unsqueeze(1).repeat(1, self.dim, 1)
        elif self.method == "node":
            out = v.unsqueeze(1).repeat(1, self.dim, 1)
        else:
            raise ValueError("Unknown initialization method %s" % self.method)
        return inputs.with_values(out)

    def get_config(self):
        config = super().get_config()
        config.update({"dim": self.dim, "method": self.method, "value": self.value, "stddev": self.stddev})
        return config


@register_keras_serializable(package="kgcnn", name="SplitEmbedding")
class SplitEmbedding(GraphBaseLayer):
    """kgcnn/layers/conv/painn_conv.py:301-346."""

    def __init__(self, num_or_size_splits, axis=-1, num=None, **kwargs):
        super().__init__(**kwargs)
        self.num_or_size_splits = num_or_size_splits
        self.axis = axis
        self.out_num = num

    def call(self, inputs, **kwargs):
        self.assert_ragged_input_rank(inputs, ragged_rank=1)
        axis = self.axis
        rank = len(inputs.shape)
        axis = axis + rank if axis < 0 else axis
        if axis <= 1:
            raise ValueError("Can not split tensor at axis <= 1.")
        v = inputs.values
        if isinstance(self.num_or_size_splits, int):
            parts = torch.chunk(v, self.num_or_size_splits, dim=axis - 1)
        else:
            parts = torch.split(v, list(self.num_or_size_splits), dim=axis - 1)
        return [inputs.with_values(p) for p in parts]

    def get_config(self):
        config = super().get_config()
        config.update({"num_or_size_splits": self.num_or_size_splits, "axis": self.axis, "num": self.out_num})
        return config
