"""RelationalDense / RelationalMLP of kgcnn/layers/relational.py:204-238 and kgcnn/layers/mlp.py:329-416.

out[n] = activation(x[n] @ kernel[relation[n]] + bias): one weight matrix per relation (atomic number in HDNNP).
The per-atom kernel gather runs through the row-gather kernel so that the kernel gradient is the deterministic
CSR segment-sum over relation (the reference gathers with tf.gather, whose gradient scatters atomically).
"""
import math

import torch

from .. import functional as F
from ..plan import Relation
from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable
from .mlp import _per_layer


@register_keras_serializable(package="kgcnn", name="RelationalDense")
class RelationalDense(GraphBaseLayer):
    def __init__(self, units, num_relations, num_bases=None, num_blocks=None, activation=None, use_bias=True,
                 kernel_initializer="glorot_uniform", bias_initializer="zeros", kernel_regularizer=None,
                 bias_regularizer=None, activity_regularizer=None, kernel_constraint=None, bias_constraint=None,
                 **kwargs):
        super().__init__(**kwargs)
        if num_bases is not None or num_blocks is not None:
            raise NotImplementedError("basis / block decompositions are not used by the named models")
        self.units = int(units)
        self.num_relations = int(num_relations)
        self.use_bias = use_bias
        self.activation = activation
        self._act_code = F.act_code(activation) if not callable(activation) else None
        self.kernel = None
        self.bias = None

    def build(self, input_shape):
        super().build(input_shape)
        in_dim = int(input_shape[-1])
        limit = math.sqrt(6.0 / (in_dim + self.units))        # each relation kernel initialised separately
        self.kernel = torch.nn.Parameter(torch.empty(self.num_relations, in_dim, self.units).uniform_(-limit, limit))
        if self.use_bias:
            self.bias = torch.nn.Parameter(torch.zeros(self.units))

    def forward(self, inputs, **kwargs):
        if not self.built:
            v = inputs[0].values if isinstance(inputs[0], RaggedTensor) else inputs[0]
            self.build(tuple(v.shape))
            self.to(v.device)
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        features, relations = self.assert_ragged_input_rank(inputs, ragged_rank=1)
        x = features.values
        r, fi, fo = self.kernel.shape
        # The reference gathers a full (F_in, units) kernel per atom (relational.py:233: a 1.47 GB temporary at
        # config 4) and batch_dots.  Equivalent without the temporary: ONE Dense against all relation kernels side
        # by side, [N, F_in] @ [F_in, R*units], then pick each atom's own block; the kernel gradient becomes the
        # split-M weight-gradient GEMM instead of a per-atom scatter.
        wcat = self.kernel.permute(1, 0, 2).reshape(fi, r * fo)
        y = F._Linear.apply(x, wcat, None).reshape(-1, r, fo)
        sel = relations.values.reshape(-1).to(torch.int64).clamp(0, r - 1)
        new = torch.gather(y, 1, sel.view(-1, 1, 1).expand(-1, 1, fo)).squeeze(1)                 # :233-234
        if self._act_code is None:
            new = self.activation(new + self.bias if self.use_bias else new)
        else:
            new = F.bias_act(new, self.bias if self.use_bias else None, self._act_code)           # :235-237
        return features.with_values(new)

    def get_config(self):
        config = super().get_config()
        config.update({"units": self.units, "use_bias": self.use_bias, "num_relations": self.num_relations,
                       "activation": self.activation})
        return config


@register_keras_serializable(package="kgcnn", name="RelationalMLP")
class RelationalMLP(GraphBaseLayer):
    """kgcnn/layers/mlp.py:329-416: stack of RelationalDense; call([features, relations])."""

    def __init__(self, units, num_relations, num_bases=None, num_blocks=None, use_bias=True, activation=None,
                 **kwargs):
        for k in list(kwargs):
            if k not in ("name", "trainable", "dtype", "node_indexing", "ragged_validate", "is_sorted",
                         "has_unconnected"):
                kwargs.pop(k)
        super().__init__(**kwargs)
        if isinstance(units, int):
            units = [units]
        depth = len(units)
        self._conf_units = list(units)
        self._conf_num_relations = _per_layer(num_relations, depth)
        self._conf_use_bias = _per_layer(use_bias, depth)
        self._conf_activation = _per_layer(activation, depth)
        self.mlp_dense_layer_list = torch.nn.ModuleList(
            [RelationalDense(units=u, num_relations=r, use_bias=b, activation=a)
             for u, r, b, a in zip(self._conf_units, self._conf_num_relations, self._conf_use_bias,
                                   self._conf_activation)])

    def call(self, inputs, **kwargs):
        x, relations = inputs
        for layer in self.mlp_dense_layer_list:
            x = layer([x, relations])
        return x

    def get_config(self):
        config = super().get_config()
        config.update({"units": self._conf_units, "num_relations": self._conf_num_relations,
                       "use_bias": self._conf_use_bias, "activation": self._conf_activation})
        return config
