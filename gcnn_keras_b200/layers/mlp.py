"""MLP / GraphMLP of kgcnn/layers/mlp.py:246-416 (stack of Dense layers with per-layer kwargs)."""
import torch

from .base import GraphBaseLayer, register_keras_serializable
from .modules import Dense


def _per_layer(value, depth):
    if isinstance(value, (list, tuple)):
        if len(value) != depth:
            raise ValueError("Provide matching list of kwargs for MLP, got %s for depth %d" % (value, depth))
        return list(value)
    return [value] * depth


class MLPBase(GraphBaseLayer):
    """kgcnn/layers/mlp.py:13-242 (the Dense part; normalization/dropout are not on the hot path)."""

    def __init__(self, units, use_bias=True, activation=None, activity_regularizer=None, kernel_regularizer=None,
                 bias_regularizer=None, kernel_initializer="glorot_uniform", bias_initializer="zeros",
                 kernel_constraint=None, bias_constraint=None, use_normalization=False, use_dropout=False,
                 rate=None, normalization_technique="BatchNormalization", **kwargs):
        for k in ("axis", "momentum", "epsilon", "center", "scale", "beta_initializer", "gamma_initializer",
                  "moving_mean_initializer", "moving_variance_initializer", "beta_regularizer", "gamma_regularizer",
                  "beta_constraint", "gamma_constraint", "noise_shape", "seed", "mean_shift"):
            kwargs.pop(k, None)
        super().__init__(**kwargs)
        if isinstance(units, int):
            units = [units]
        self._depth = len(units)
        self._conf_units = list(units)
        self._conf_use_bias = _per_layer(use_bias, self._depth)
        self._conf_activation = _per_layer(activation, self._depth)
        if any(_per_layer(use_normalization, self._depth)) or any(_per_layer(use_dropout, self._depth)):
            raise NotImplementedError("normalization / dropout inside MLP are outside the accelerated hot path")
        self.mlp_dense_layer_list = torch.nn.ModuleList(
            [Dense(units=u, use_bias=b, activation=a, kernel_initializer=kernel_initializer,
                   bias_initializer=bias_initializer)
             for u, b, a in zip(self._conf_units, self._conf_use_bias, self._conf_activation)])

    def call(self, inputs, **kwargs):
        x = inputs
        for layer in self.mlp_dense_layer_list:
            x = layer(x)
        return x

    def get_config(self):
        config = super().get_config()
        config.update({"units": self._conf_units, "use_bias": self._conf_use_bias,
                       "activation": self._conf_activation})
        return config


@register_keras_serializable(package="kgcnn", name="MLP")
class MLP(MLPBase):
    """kgcnn/layers/mlp.py:246-322: Dense stack on dense tensors and, through map_values, on ragged embeddings."""


GraphMLP = MLP      # kgcnn/layers/mlp.py:325 (an alias, registered once as 'kgcnn>MLP')
