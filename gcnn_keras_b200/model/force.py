"""EnergyForceModel of kgcnn/model/force.py:11-242: wraps an energy model and returns dE/dx beside the energy.

The reference pads the coordinates, runs the energy model inside a GradientTape, takes ``batch_jacobian`` and
transposes (force.py:152-188).  Graphs are independent, so the per-graph Jacobian equals the gradient of the
summed energy with respect to the ragged coordinate values; that is what is computed here (no pad/unpad, no
pfor), with ``create_graph`` on so that the force can be trained on (double backward) exactly like Keras'
outer tape differentiating through the inner one.  Constructor arguments and outputs are unchanged.
"""
import importlib

import torch

from ..layers.casting import ChangeTensorType
from ..ragged import RaggedTensor


def get_model_class(module_name, class_name):
    """kgcnn/model/utils.py:17-39 restricted to this package's literature modules."""
    module_name = module_name.replace("kgcnn.literature", "gcnn_keras_b200.literature")
    if "." not in module_name:
        module_name = "gcnn_keras_b200.literature." + module_name
    return getattr(importlib.import_module(module_name), class_name)


class EnergyForceModel(torch.nn.Module):
    def __init__(self, model_energy=None, coordinate_input=1, esp_input=None, esp_grad_input=None, energy_output=1,
                 output_as_dict=True, ragged_validate=False, output_to_tensor=True, output_squeeze_states=False,
                 nested_model_config=True, is_physical_force=True, **kwargs):
        super().__init__()
        if model_energy is None:
            raise ValueError("Require valid model in `model_energy` for force prediction.")
        self._model_energy = model_energy
        if isinstance(model_energy, torch.nn.Module):
            self.energy_model = model_energy
        elif isinstance(model_energy, dict):
            cls = get_model_class(model_energy["module_name"], model_energy["class_name"])
            self.energy_model = cls(**model_energy["config"])
        else:
            raise TypeError("Input `model_energy` must be dict or `ks.models.Model` .")
        if output_as_dict is True and energy_output != 0:      # force.py:118-120
            print("Kgcnn warning: energy-model returns more than just energy, setting output_as_dict as False")
            output_as_dict = False
        if esp_input is not None or esp_grad_input is not None:
            raise NotImplementedError("esp inputs (QM/MM, kgcnn/model/force.py:153-158) are outside the hot path")
        self.ragged_validate = ragged_validate
        self.coordinate_input = coordinate_input
        self.esp_input, self.esp_grad_input = esp_input, esp_grad_input
        self.energy_output = energy_output
        self.output_as_dict = output_as_dict
        self.output_to_tensor = output_to_tensor
        self.output_squeeze_states = output_squeeze_states
        self.is_physical_force = is_physical_force
        self.nested_model_config = nested_model_config
        self.cast_coordinates = ChangeTensorType(input_tensor_type="ragged", output_tensor_type="mask")

    def forward(self, inputs, training=None, **kwargs):
        x = inputs[self.coordinate_input]
        inputs_energy = list(inputs)
        xv = x.values
        if not xv.requires_grad:
            xv = xv.detach().requires_grad_(True)               # tape.watch(x_pad), force.py:160
        inputs_energy[self.coordinate_input] = x.with_values(xv)
        train = torch.is_grad_enabled() if training is None else bool(training)
        with torch.enable_grad():
            outputs = self.energy_model(inputs_energy, **kwargs)
            eng = outputs[self.energy_output] if isinstance(outputs, (list, tuple)) else outputs
            states = int(eng.shape[-1]) if eng.dim() > 1 else 1
            eng2 = eng.reshape(eng.shape[0], states)
            cols = []
            from ..functional import input_gradients_only
            with input_gradients_only(xv):                       # dE/dx only: no weight gradients in this sweep
                for s in range(states):                          # batch_jacobian, force.py:176
                    (g,) = torch.autograd.grad(eng2[:, s].sum(), xv, create_graph=train, retain_graph=True)
                    cols.append(g)
        de_dr = torch.stack(cols, dim=-1)                        # (N, 3, states) = transpose at force.py:177
        if not train:
            de_dr = de_dr.detach()
            eng = eng.detach()
            if isinstance(outputs, (list, tuple)):
                outputs = [o.detach() if isinstance(o, torch.Tensor) else o for o in outputs]
            else:
                outputs = eng
        if self.is_physical_force:                               # force.py:185-186
            de_dr = -de_dr
        if self.output_squeeze_states:                           # force.py:187-188
            de_dr = de_dr.squeeze(-1)
        force = x.with_values(de_dr)
        if self.output_to_tensor:                                # force.py:189-194
            force = force.to_tensor()
            if isinstance(outputs, (list, tuple)):
                outputs = [o.to_tensor() if isinstance(o, RaggedTensor) else o for o in outputs]
        if self.output_as_dict:
            return {"energy": eng, "force": force}
        if isinstance(outputs, (list, tuple)):
            return list(outputs) + [force]
        return outputs, force

    def get_config(self):
        return {"model_energy": self._model_energy if isinstance(self._model_energy, dict) else
                getattr(self.energy_model, "model_config", None),
                "coordinate_input": self.coordinate_input, "esp_input": self.esp_input,
                "esp_grad_input": self.esp_grad_input, "output_as_dict": self.output_as_dict,
                "ragged_validate": self.ragged_validate, "output_to_tensor": self.output_to_tensor,
                "output_squeeze_states": self.output_squeeze_states,
                "nested_model_config": self.nested_model_config}
