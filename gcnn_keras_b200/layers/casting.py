"""ChangeTensorType of kgcnn/layers/casting.py:13-102 (ragged <-> padded + mask), used by EnergyForceModel."""
import torch

from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable


@register_keras_serializable(package="kgcnn", name="ChangeTensorType")
class ChangeTensorType(GraphBaseLayer):
    def __init__(self, partition_type="row_length", input_tensor_type="RaggedTensor", output_tensor_type="RaggedTensor",
                 shape=None, default_value=None, boolean_mask=False, **kwargs):
        super().__init__(**kwargs)
        self.partition_type = partition_type
        self.input_tensor_type = str(input_tensor_type)
        self.output_tensor_type = str(output_tensor_type)
        self.shape = shape
        self.default_value = default_value
        self.boolean_mask = boolean_mask
        self._str_type_ragged = ["ragged", "RaggedTensor"]
        self._str_type_tensor = ["Tensor", "tensor"]
        self._str_type_mask = ["padded", "masked", "mask"]

    def call(self, inputs, **kwargs):
        if self.input_tensor_type in self._str_type_ragged:
            if self.output_tensor_type in self._str_type_ragged:
                return inputs
            dv = 0 if self.default_value is None else self.default_value
            if self.output_tensor_type in self._str_type_tensor:
                return inputs.to_tensor(default_value=dv, shape=self.shape)           # casting.py:78-79
            if self.output_tensor_type in self._str_type_mask:         # casting.py:80-84 (padded + mask, one launch)
                return inputs.to_tensor_and_mask(default_value=dv, boolean_mask=self.boolean_mask, shape=self.shape)
        raise NotImplementedError("Unsupported conversion from '%s' to '%s'."
                                  % (self.input_tensor_type, self.output_tensor_type))

    def get_config(self):
        config = super().get_config()
        config.update({"partition_type": self.partition_type, "input_tensor_type": self.input_tensor_type,
                       "output_tensor_type": self.output_tensor_type, "shape": self.shape,
                       "default_value": self.default_value, "boolean_mask": self.boolean_mask})
        return config
