"""Ragged-safe Dense / Activation / Lazy* / ExpandDims / OptionalInputEmbedding of kgcnn/layers/modules.py.

All of them act on the flat ``values`` tensor (map_values, kgcnn/layers/base.py:112-171).  Dense = library GEMM
+ our fused bias/activation kernel family; the Lazy* element-wise glue stays as thin tensor ops here and is
absorbed by the fused kernels inside the conv blocks (SchNetCFconv, PAiNNconv, GCN).
"""
import math

import torch

from .. import functional as F
from ..plan import Relation
from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable


def _glorot_uniform_(w):
    fan_in, fan_out = w.shape[-2], w.shape[-1]
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        return w.uniform_(-limit, limit)


def _values(x):
    return x.values if isinstance(x, RaggedTensor) else x


def _rewrap(like, values):
    return like.with_values(values) if isinstance(like, RaggedTensor) else values


@register_keras_serializable(package="kgcnn", name="DenseEmbedding")
class DenseEmbedding(GraphBaseLayer):
    """kgcnn/layers/modules.py:15-90: activation(x W + b) on the last axis of the flat values."""

    def __init__(self, units, activation=None, use_bias=True, kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None,
                 kernel_constraint=None, bias_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self.units = int(units)
        self.activation = activation
        self.use_bias = use_bias
        self._cfg = {"kernel_initializer": kernel_initializer, "bias_initializer": bias_initializer,
                     "kernel_regularizer": kernel_regularizer, "bias_regularizer": bias_regularizer,
                     "activity_regularizer": activity_regularizer, "kernel_constraint": kernel_constraint,
                     "bias_constraint": bias_constraint}
        self._act_code = F.act_code(activation) if not callable(activation) else None
        self.kernel = None
        self.bias = None

    def build(self, input_shape):
        super().build(input_shape)
        in_dim = int(input_shape[-1])
        self.kernel = torch.nn.Parameter(_glorot_uniform_(torch.empty(in_dim, self.units)))
        if self.use_bias:
            self.bias = torch.nn.Parameter(torch.zeros(self.units))

    def forward(self, inputs, **kwargs):
        if not self.built:
            v = _values(inputs)
            self.build(tuple(v.shape))
            self.to(v.device)
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        v = _values(inputs)
        if self._act_code is None:
            out = self.activation(F.dense(v, self.kernel, self.bias, "linear"))
        else:
            out = F.dense(v, self.kernel, self.bias, self._act_code)
        return _rewrap(inputs, out)

    def get_config(self):
        config = super().get_config()
        config.update({"units": self.units, "activation": self.activation, "use_bias": self.use_bias})
        config.update(self._cfg)
        return config


Dense = DenseEmbedding


@register_keras_serializable(package="kgcnn", name="ActivationEmbedding")
class ActivationEmbedding(GraphBaseLayer):
    """kgcnn/layers/modules.py:93-131."""

    def __init__(self, activation, activity_regularizer=None, **kwargs):
        super().__init__(**kwargs)
        self.activation = activation
        self.activity_regularizer = activity_regularizer
        self._act_code = F.act_code(activation) if not callable(activation) else None

    def call(self, inputs, **kwargs):
        v = _values(inputs)
        out = self.activation(v) if self._act_code is None else F.bias_act(v, None, self._act_code)
        return _rewrap(inputs, out)

    def get_config(self):
        config = super().get_config()
        config.update({"activation": self.activation, "activity_regularizer": self.activity_regularizer})
        return config


Activation = ActivationEmbedding


class _Lazy(GraphBaseLayer):
    def _op(self, vals):
        raise NotImplementedError

    def call(self, inputs, **kwargs):
        return self.map_values(self._op, list(inputs))


@register_keras_serializable(package="kgcnn", name="LazyAdd")
class LazyAdd(_Lazy):
    """kgcnn/layers/modules.py:187-222."""

    def _op(self, vals):
        out = vals[0]
        for v in vals[1:]:
            out = out + v
        return out


@register_keras_serializable(package="kgcnn", name="LazySubtract")
class LazySubtract(_Lazy):
    """kgcnn/layers/modules.py:226-262."""

    def _op(self, vals):
        return vals[0] - vals[1]


@register_keras_serializable(package="kgcnn", name="LazyMultiply")
class LazyMultiply(_Lazy):
    """kgcnn/layers/modules.py:305-340."""

    def _op(self, vals):
        out = vals[0]
        for v in vals[1:]:
            out = out * v
        return out


@register_keras_serializable(package="kgcnn", name="LazyAverage")
class LazyAverage(_Lazy):
    def _op(self, vals):
        return sum(vals[1:], vals[0]) / float(len(vals))


@register_keras_serializable(package="kgcnn", name="LazyConcatenate")
class LazyConcatenate(GraphBaseLayer):
    """kgcnn/layers/modules.py:344-364: concatenate along a feature axis of the flat values."""

    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def call(self, inputs, **kwargs):
        axis = self.axis
        if axis > 1:
            axis -= 1
        elif axis in (0, 1):
            raise ValueError("LazyConcatenate can not concatenate ragged tensors along axis %d" % self.axis)
        vals = [_values(x) for x in inputs]
        return _rewrap(inputs[0], torch.cat(vals, dim=axis))

    def get_config(self):
        config = super().get_config()
        config.update({"axis": self.axis})
        return config


@register_keras_serializable(package="kgcnn", name="ExpandDims")
class ExpandDims(GraphBaseLayer):
    """kgcnn/layers/modules.py:368-416."""

    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def call(self, inputs, **kwargs):
        if isinstance(inputs, RaggedTensor):
            axis = self.axis
            if axis > 1:
                axis -= 1
            elif axis in (0, 1):
                raise ValueError("ExpandDims can not insert axis %d into a ragged tensor" % self.axis)
            return inputs.with_values(inputs.values.unsqueeze(axis))
        return inputs.unsqueeze(self.axis)

    def get_config(self):
        config = super().get_config()
        config.update({"axis": self.axis})
        return config


@register_keras_serializable(package="kgcnn", name="ZerosLike")
class ZerosLike(GraphBaseLayer):
    def call(self, inputs, **kwargs):
        return self.map_values(torch.zeros_like, inputs)


@register_keras_serializable(package="kgcnn", name="OptionalInputEmbedding")
class OptionalInputEmbedding(GraphBaseLayer):
    """kgcnn/layers/modules.py:450-534: ks.layers.Embedding on integer node attributes (if use_embedding).
    Forward is our row-gather kernel on the table; the table gradient is the deterministic CSR segment-sum
    over atom type (TF uses an IndexedSlices scatter)."""

    def __init__(self, input_dim, output_dim, use_embedding=False, embeddings_initializer="uniform",
                 embeddings_regularizer=None, activity_regularizer=None, embeddings_constraint=None, mask_zero=False,
                 input_length=None, **kwargs):
        super().__init__(**kwargs)
        self.use_embedding = use_embedding
        self.input_dim, self.output_dim = int(input_dim), int(output_dim)
        self._cfg = {"embeddings_initializer": embeddings_initializer,
                     "embeddings_regularizer": embeddings_regularizer, "activity_regularizer": activity_regularizer,
                     "embeddings_constraint": embeddings_constraint, "mask_zero": mask_zero,
                     "input_length": input_length}
        if use_embedding:
            table = torch.empty(self.input_dim, self.output_dim).uniform_(-0.05, 0.05)   # keras 'uniform'
            self.embeddings = torch.nn.Parameter(table)

    def call(self, inputs, **kwargs):
        if not self.use_embedding:
            return inputs
        v = _values(inputs)
        if self.embeddings.device != v.device:
            self.to(v.device)
        cache = getattr(inputs, "_embedding_relation", None) if isinstance(inputs, RaggedTensor) else None
        if cache is not None and cache[0] == (v._version, self.input_dim):
            rel = cache[1]
        else:
            rel = Relation(v.reshape(-1).to(torch.int32), self.input_dim)   # sorted by atom type once per batch
            if isinstance(inputs, RaggedTensor):
                inputs._embedding_relation = ((v._version, self.input_dim), rel)
        out = F.gather(self.embeddings, rel).reshape(tuple(v.shape) + (self.output_dim,))
        return _rewrap(inputs, out)

    def get_config(self):
        config = super().get_config()
        config.update({"use_embedding": self.use_embedding, "input_dim": self.input_dim,
                       "output_dim": self.output_dim})
        config.update(self._cfg)
        return config
