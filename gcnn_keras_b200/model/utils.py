"""Model-kwargs plumbing of kgcnn/model/utils.py:69-140 and a minimal functional-model container."""
import functools
from copy import deepcopy
from math import inf

import torch

from ..plan import unify_partitions


def update_model_kwargs_logic(default_kwargs=None, user_kwargs=None, update_recursive=inf):
    """kgcnn/model/utils.py:69-117: nested dict update; unknown top-level keys raise ValueError."""
    default_kwargs = default_kwargs or {}
    user_kwargs = user_kwargs or {}
    for key in user_kwargs:
        if key not in default_kwargs:
            raise ValueError("Model kwarg {0} not in default arguments {1}".format(key, default_kwargs.keys()))
    out = deepcopy(default_kwargs)

    def _nested_update(d1, d2, max_depth=inf, depth=0):
        for key, values in d2.items():
            if key not in d1 or not isinstance(d1[key], dict) or not isinstance(values, dict):
                d1[key] = values
                continue
            d1[key] = _nested_update(d1[key], values, max_depth, depth + 1) if depth < max_depth else values
        return d1

    return _nested_update(out, user_kwargs, update_recursive, 0)


def update_model_kwargs(model_default, update_recursive=inf):
    """kgcnn/model/utils.py:120-140."""

    def decorator(func):
        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            updated = update_model_kwargs_logic(model_default, kwargs, update_recursive)
            return func(*args, **updated)

        return wrapper

    return decorator


class Model(torch.nn.Module):
    """Stand-in for ``ks.models.Model(inputs=..., outputs=...)``: holds the layers built by a ``make_model``
    function and the wiring as a python callable.  ``model(inputs)`` takes the list of (ragged) tensors in the
    order of the ``inputs`` config, exactly like the Keras functional model."""

    def __init__(self, layers, wiring, input_config=None, name=None, config=None):
        super().__init__()
        self.name = name
        self.input_config = input_config
        self.model_config = config
        self.layers_by_name = torch.nn.ModuleDict(layers)
        self._wiring = wiring

    def forward(self, inputs, **kwargs):
        inputs = unify_partitions(list(inputs))
        return self._wiring(self.layers_by_name, inputs, **kwargs)

    def get_config(self):
        return deepcopy(self.model_config)

    def get_weights_dict(self):
        """Flat {name: tensor} of all weights with the reference's attribute names (for parity tests)."""
        return {k: v.detach() for k, v in self.state_dict().items()}
