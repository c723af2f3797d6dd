"""RelationalDense / RelationalMLP of kgcnn/layers/relational.py:204-238 and kgcnn/layers/mlp.py:329-416.

out[n] = activation(x[n] @ kernel[relation[n]] + bias): one weight matrix per relation (atomic number in HDNNP).
The reference gathers a full (F_in, units) kernel per atom (relational.py:233) and batch_dots.  Here the atoms are
sorted by relation once per batch (the stable CSR plan of the relation tensor), every relation's rows then form
one contiguous Dense problem against its own kernel, and all relations run in ONE grouped tcgen05 launch
(kgcnn_dense_rel: forward, input gradient and per-relation kernel gradient, closed under differentiation).
A RelationalMLP sorts on entry and un-sorts on exit, not per layer.
"""
import math

import torch

from .. import functional as F
from ..plan import Relation
from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable
from .mlp import _per_layer


class RelationGroups:
    """Rows sorted by relation: `offs` int32 [R + 1] (device) row offsets, `perm` / `inv` the stable sort and its
    inverse (None when the relation tensor is already ascending).  A relation outside [0, R) raises IndexError
    (the reference's tf.gather fails on CPU and returns zeros on GPU); inside a captured step it sets the step's
    guard flag instead, and the optimizer update of that replay is skipped (see plan.PlanMode)."""

    def __init__(self, relation_values, n_rel):
        ids = relation_values.reshape(-1)
        if ids.dtype != torch.int32:
            ids = F.raw_cast_i32(ids.to(torch.int64).contiguous()) if ids.is_cuda else ids.to(torch.int32)
        rel = Relation(ids.contiguous(), int(n_rel))
        self.n_rel = int(n_rel)
        self.offs = rel.rowptr
        self.perm = rel.perm
        self.inv = None
        if self.perm is not None:
            self.inv = torch.empty_like(self.perm)
            self.inv[self.perm.long()] = torch.arange(self.perm.shape[0], dtype=torch.int32, device=self.perm.device)

    @classmethod
    def of(cls, relation_values, n_rel):
        """Cached on the relation tensor (identity + version), like the index plans."""
        from ..plan import _cache_slot
        slot = _cache_slot(relation_values)
        key = ("relation_groups", relation_values._version, int(n_rel))
        if key not in slot:
            for k in [k for k in slot if k[0] == "relation_groups" and k != key]:
                del slot[k]
            slot[key] = cls(relation_values, n_rel)
        return slot[key]

    def sort(self, x):
        return x if self.perm is None else F.permute_rows(x, self.perm, self.inv)

    def unsort(self, y):
        return y if self.perm is None else F.permute_rows(y, self.inv, self.perm)


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
        group = kwargs.get("relation_groups")
        if group is None:                                   # a layer used on its own sorts and un-sorts itself
            group = RelationGroups.of(relations.values, self.num_relations)
            return features.with_values(group.unsort(self._call_sorted(group.sort(features.values), group)))
        return features.with_values(self._call_sorted(features.values, group))

    def _call_sorted(self, x, group):
        if group.n_rel != self.num_relations:
            raise ValueError("relation groups built for %d relations, layer has %d" % (group.n_rel, self.num_relations))
        new = F.dense_rel(x, self.kernel, group.offs)                                              # :233-234
        if self._act_code is None:
            return self.activation(new + self.bias if self.use_bias else new)
        return F.bias_act(new, self.bias if self.use_bias else None, self._act_code)               # :235-237

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
        if len(set(self._conf_num_relations)) != 1:         # layers with different relation counts: each sorts itself
            for layer in self.mlp_dense_layer_list:
                x = layer([x, relations])
            return x
        group = RelationGroups.of(relations.values, self._conf_num_relations[0])
        x = x.with_values(group.sort(x.values))
        for layer in self.mlp_dense_layer_list:
            x = layer([x, relations], relation_groups=group)
        return x.with_values(group.unsort(x.values))

    def get_config(self):
        config = super().get_config()
        config.update({"units": self._conf_units, "num_relations": self._conf_num_relations,
                       "use_bias": self._conf_use_bias, "activation": self._conf_activation})
        return config
