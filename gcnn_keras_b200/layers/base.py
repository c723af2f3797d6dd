"""GraphBaseLayer: the Keras-style layer contract of kgcnn/layers/base.py:8-173 on torch.nn.Module.

Kept from the reference: constructor flags ``node_indexing, ragged_validate, is_sorted, has_unconnected``
(base.py:18-45), ``get_config()`` keys (base.py:47-64), ``assert_ragged_input_rank`` (base.py:70-110),
``map_values`` (base.py:112-171), lazy ``build(input_shape)`` on first call, ``call(inputs, **kwargs)``.
"""
import torch

from ..ragged import RaggedTensor

_LAYER_COUNTERS = {}
_REGISTRY = {}


def register_keras_serializable(package="kgcnn", name=None):
    """Same registry key scheme as ``ks.utils.register_keras_serializable`` ('kgcnn>Name')."""

    def deco(cls):
        _REGISTRY["%s>%s" % (package, name or cls.__name__)] = cls
        cls._keras_registered_name = "%s>%s" % (package, name or cls.__name__)
        return cls

    return deco


def get_registered(name):
    return _REGISTRY[name]


def _auto_name(cls_name):
    snake = "".join("_" + c.lower() if c.isupper() and i and not cls_name[i - 1].isupper() else c.lower()
                    for i, c in enumerate(cls_name))
    n = _LAYER_COUNTERS.get(snake, 0)
    _LAYER_COUNTERS[snake] = n + 1
    return snake if n == 0 else "%s_%d" % (snake, n)


def _shape_of(x):
    if isinstance(x, (list, tuple)):
        return [_shape_of(i) for i in x]
    return tuple(x.shape)


class GraphBaseLayer(torch.nn.Module):
    def __init__(self, node_indexing="sample", ragged_validate=False, is_sorted=False, has_unconnected=True,
                 name=None, trainable=True, dtype="float32", **kwargs):
        if kwargs:
            raise TypeError("Keyword argument not understood: %s" % sorted(kwargs))
        super().__init__()
        self.node_indexing = node_indexing
        self.ragged_validate = ragged_validate
        self.is_sorted = is_sorted
        self.has_unconnected = has_unconnected
        self.name = name or _auto_name(type(self).__name__)
        self.trainable = trainable
        self._dtype_name = dtype
        self.built = False
        if self.node_indexing != "sample":   # base.py:42-43
            raise ValueError("Indexing for disjoint representation is not supported as of version 1.0")

    # ---- Keras surface --------------------------------------------------------------------------------
    def build(self, input_shape):
        self.built = True

    def call(self, inputs, **kwargs):
        raise NotImplementedError

    def forward(self, inputs, **kwargs):
        if not self.built:
            self.build(_shape_of(inputs))
            self.built = True
        return self.call(inputs, **kwargs)

    def get_config(self):
        return {"name": self.name, "trainable": self.trainable, "dtype": self._dtype_name,
                "node_indexing": self.node_indexing, "ragged_validate": self.ragged_validate,
                "is_sorted": self.is_sorted, "has_unconnected": self.has_unconnected}

    @classmethod
    def from_config(cls, config):
        return cls(**config)

    # ---- helpers ----------------------------------------------------------------------------------------
    def assert_ragged_input_rank(self, inputs, mask=None, ragged_rank=1):
        """base.py:70-110: inputs must be ragged (rank-1); dense [B, N, ...] tensors are cast."""
        if mask is not None:
            raise ValueError("Using `mask` argument in `assert_ragged_input_rank` is not yet supported.")

        def validate_or_cast(x):
            if isinstance(x, RaggedTensor):
                assert x.ragged_rank == ragged_rank, "'%s' must have input with ragged_rank=%s." % (
                    self.name, ragged_rank)
                return x
            if isinstance(x, torch.Tensor):
                if ragged_rank != 1:
                    raise ValueError("Casting to ragged is only supported for ragged_rank=1 at the moment.")
                if x.dim() <= ragged_rank:
                    raise ValueError("Rank of inputs must be > ragged_rank but found '%s <= %s' "
                                     % (x.dim(), ragged_rank))
                lengths = torch.full((x.shape[0],), x.shape[1], dtype=torch.int64, device=x.device)
                return RaggedTensor.from_row_lengths(x.reshape((x.shape[0] * x.shape[1],) + tuple(x.shape[2:])),
                                                     lengths)
            raise ValueError("Unsupported tensor type '%s' in '%s'." % (type(x), self.name))

        if isinstance(inputs, (list, tuple)):
            return [validate_or_cast(i) for i in inputs]
        return validate_or_cast(inputs)

    def map_values(self, fun, inputs, **kwargs):
        """base.py:112-171: call `fun` on the values tensor(s); `axis` kwargs are shifted by one."""
        kwargs_values = dict(kwargs)
        if "axis" in kwargs:
            axis = kwargs["axis"]
            axis_values = None
            if isinstance(axis, int):
                if axis > 1:
                    axis_values = axis - 1
                elif axis < 0:
                    axis_values = axis
            elif isinstance(axis, (list, tuple)) and all(a > 1 for a in axis):
                axis_values = [a - 1 for a in axis]
            if axis_values is None:
                raise ValueError("Layer %s can not apply `axis=%s` on the values of a ragged tensor."
                                 % (self.name, axis))
            kwargs_values["axis"] = axis_values
        if isinstance(inputs, (list, tuple)):
            if all(isinstance(x, RaggedTensor) for x in inputs):
                out = fun([x.values for x in inputs], **kwargs_values)
                if isinstance(out, (list, tuple)):
                    return [inputs[i].with_values(o) for i, o in enumerate(out)]
                return inputs[0].with_values(out)
            return fun(list(inputs), **kwargs)
        if isinstance(inputs, RaggedTensor):
            out = fun(inputs.values, **kwargs_values)
            if isinstance(out, (list, tuple)):
                return [inputs.with_values(o) for o in out]
            return inputs.with_values(out)
        return fun(inputs, **kwargs)

    call_on_values_tensor_of_ragged = map_values
