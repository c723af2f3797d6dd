"""Geometry layers of kgcnn/layers/geom.py on the hot path: NodePosition, EuclideanNorm,
NodeDistanceEuclidean, EdgeDirectionNormalized, ScalarProduct, GaussBasisLayer, BesselBasisLayer,
CosCutOffEnvelope, CosCutOff.  Same names, kwargs and get_config keys as the reference.

The coordinate gathers run through the row-gather kernel (closed under differentiation with the CSR
segment-sum); the per-edge scalar math is written with differentiable tensor ops so that forces (first
derivative) and force training (second derivative) come out of the tape exactly as in the reference.
"""
import math

import numpy as np
import torch

from .. import functional as F
from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable
from .gather import GatherNodesSelection

KERAS_EPSILON = 1e-7
FUSED = True     # fused sm_100a edge-geometry kernel when the inputs carry their provenance; False: op-by-op


def _pair_source(a, b):
    """(coordinates ragged, rel_a, rel_b) if a and b are row gathers of the SAME [N,3] fp32 CUDA tensor."""
    pa, pb = getattr(a, "provenance", None), getattr(b, "provenance", None)
    if not FUSED or pa is None or pb is None or pa[0] != "gathered" or pb[0] != "gathered":
        return None
    src = pa[1]
    v = src.values
    if pb[1].values is not v or v.dim() != 2 or v.shape[1] != 3 or v.dtype != torch.float32 or not v.is_cuda:
        return None
    return src, pa[2], pb[2]


def _distance_source(ed):
    """(coordinates ragged, rel_i, rel_j, eps) if `ed` is the untouched output of NodeDistanceEuclidean."""
    p = getattr(ed, "provenance", None)
    if not FUSED or p is None or p[0] != "distance":
        return None
    return p[1], p[2], p[3], p[4]


def _positive_axis(axis, rank):
    """kgcnn/ops/axis.py:2-34."""
    return axis + rank if axis < 0 else axis


@register_keras_serializable(package="kgcnn", name="NodePosition")
class NodePosition(GraphBaseLayer):
    """kgcnn/layers/geom.py:14-73: GatherNodesSelection([0, 1]) on coordinates."""

    def __init__(self, selection_index=None, **kwargs):
        super().__init__(**kwargs)
        self.selection_index = [0, 1] if selection_index is None else selection_index
        self.layer_gather = GatherNodesSelection(self.selection_index)

    def call(self, inputs, **kwargs):
        return self.layer_gather(inputs, **kwargs)

    def get_config(self):
        config = super().get_config()
        config.update({"selection_index": self.selection_index})
        return config


def compute_euclidean_norm(x, axis=-1, keepdims=False, invert_norm=False, add_eps=False, no_nan=True,
                           square_norm=False):
    """EuclideanNorm._compute_euclidean_norm, geom.py:165-193."""
    out = torch.relu(torch.sum(torch.square(x), dim=axis, keepdim=keepdims))
    if add_eps:
        out = out + KERAS_EPSILON
    if not square_norm:
        out = torch.sqrt(out)
    if invert_norm:
        if no_nan:
            safe = torch.where(out == 0, torch.ones_like(out), out)
            out = torch.where(out == 0, torch.zeros_like(out), 1.0 / safe)
        else:
            out = 1.0 / out
    return out


@register_keras_serializable(package="kgcnn", name="EuclideanNorm")
class EuclideanNorm(GraphBaseLayer):
    """kgcnn/layers/geom.py:127-214."""

    def __init__(self, axis=-1, keepdims=False, invert_norm=False, add_eps=False, no_nan=True, square_norm=False,
                 **kwargs):
        super().__init__(**kwargs)
        self.axis = axis
        self.keepdims = keepdims
        self.invert_norm = invert_norm
        self.square_norm = square_norm
        self.add_eps = add_eps
        self.no_nan = no_nan

    def build(self, input_shape):
        super().build(input_shape)
        # input_shape here is the shape of the ragged tensor (batch, None, ...): rank = values rank + 1
        self.axis = _positive_axis(self.axis, len(input_shape))

    def forward(self, inputs, **kwargs):
        if not self.built:
            self.build(inputs.shape if isinstance(inputs, RaggedTensor) else tuple(inputs.shape))
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        return self.map_values(compute_euclidean_norm, inputs, axis=self.axis, keepdims=self.keepdims,
                               invert_norm=self.invert_norm, add_eps=self.add_eps, no_nan=self.no_nan,
                               square_norm=self.square_norm)

    def get_config(self):
        config = super().get_config()
        config.update({"axis": self.axis, "keepdims": self.keepdims, "invert_norm": self.invert_norm,
                       "add_eps": self.add_eps, "no_nan": self.no_nan, "square_norm": self.square_norm})
        return config


@register_keras_serializable(package="kgcnn", name="ScalarProduct")
class ScalarProduct(GraphBaseLayer):
    """kgcnn/layers/geom.py:218-281: sum(a * b) along `axis`."""

    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def forward(self, inputs, **kwargs):
        if not self.built:
            first = inputs[0]
            self.axis = _positive_axis(self.axis, len(first.shape))
            self.built = True
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        return self.map_values(lambda v, axis: torch.sum(v[0] * v[1], dim=axis), list(inputs), axis=self.axis)

    def get_config(self):
        config = super().get_config()
        config.update({"axis": self.axis})
        return config


@register_keras_serializable(package="kgcnn", name="NodeDistanceEuclidean")
class NodeDistanceEuclidean(GraphBaseLayer):
    """kgcnn/layers/geom.py:285-327: ||x1 - x2|| with keepdims -> (batch, [M], 1)."""

    def __init__(self, add_eps=False, no_nan=True, **kwargs):
        super().__init__(**kwargs)
        self.add_eps, self.no_nan = add_eps, no_nan

    def call(self, inputs, **kwargs):
        a, b = inputs
        eps = KERAS_EPSILON if self.add_eps else 0.0
        fused = _pair_source(a, b)
        if fused is not None:                      # gather + subtract + norm in one launch (csrc/geometry.cu)
            src, ri, rj = fused
            (d,) = F.edge_geometry(src.values, F.GeometrySpec(ri, rj, ("d",), eps=eps))
            out = a.with_values(d)
            out.provenance = ("distance", src, ri, rj, eps)
            return out
        diff = a.values - b.values
        return a.with_values(compute_euclidean_norm(diff, axis=1, keepdims=True, add_eps=self.add_eps,
                                                    no_nan=self.no_nan))

    def get_config(self):
        config = super().get_config()
        config.update({"add_eps": self.add_eps, "no_nan": self.no_nan})
        return config


@register_keras_serializable(package="kgcnn", name="EdgeDirectionNormalized")
class EdgeDirectionNormalized(GraphBaseLayer):
    """kgcnn/layers/geom.py:331-378: (r_i - r_j) / ||r_i - r_j|| with divide_no_nan."""

    def __init__(self, add_eps=False, no_nan=True, **kwargs):
        super().__init__(**kwargs)
        self.add_eps, self.no_nan = add_eps, no_nan

    def call(self, inputs, **kwargs):
        a, b = inputs
        fused = _pair_source(a, b)
        if fused is not None:
            src, ri, rj = fused
            (u,) = F.edge_geometry(src.values, F.GeometrySpec(ri, rj, ("u",), eps=KERAS_EPSILON if self.add_eps else 0.0,
                                                              no_nan=self.no_nan))
            return a.with_values(u)
        diff = a.values - b.values
        inv = compute_euclidean_norm(diff, axis=1, keepdims=True, invert_norm=True, add_eps=self.add_eps,
                                     no_nan=self.no_nan)
        return a.with_values(diff * inv)

    def get_config(self):
        config = super().get_config()
        config.update({"add_eps": self.add_eps, "no_nan": self.no_nan})
        return config


@register_keras_serializable(package="kgcnn", name="GaussBasisLayer")
class GaussBasisLayer(GraphBaseLayer):
    """kgcnn/layers/geom.py:514-592: exp(-gamma (d - offset - mu_k)^2), mu_k = k/bins*distance."""

    def __init__(self, bins=20, distance=4.0, sigma=0.4, offset=0.0, **kwargs):
        super().__init__(**kwargs)
        self.bins = int(bins)
        self.distance = float(distance)
        self.offset = float(offset)
        self.sigma = float(sigma)
        self.gamma = 1 / sigma / sigma / 2

    @staticmethod
    def _compute_gauss_basis(inputs, offset, gamma, bins, distance):
        gbs = torch.arange(0, bins, 1, dtype=inputs.dtype, device=inputs.device) / float(bins) * distance
        out = inputs - offset
        out = torch.square(out - gbs) * (gamma * (-1.0))
        return torch.exp(out)

    def call(self, inputs, **kwargs):
        fused = _distance_source(inputs) if isinstance(inputs, RaggedTensor) else None
        if fused is not None and self.bins <= 64:  # straight from the coordinates: d never becomes a graph node
            src, ri, rj, eps = fused
            spec = F.GeometrySpec(ri, rj, ("rbf",), basis=1, R=self.bins, eps=eps,
                                  basis_params=(self.gamma, self.offset, self.distance / float(self.bins)))
            out = inputs.with_values(F.edge_geometry(src.values, spec)[0])
        elif isinstance(inputs, RaggedTensor) and FUSED and inputs.values.is_cuda \
                and inputs.values.dtype == torch.float32 and inputs.values.numel() == inputs.values.shape[0]:
            # distances that are a model input (make_distance=False): one launch of kgcnn_gauss_basis
            out = inputs.with_values(F.gauss_basis(inputs.values, self.bins, self.gamma, self.offset,
                                                   self.distance / float(self.bins)))
        else:
            out = self.map_values(self._compute_gauss_basis, inputs, offset=self.offset, gamma=self.gamma,
                                  bins=self.bins, distance=self.distance)
        if isinstance(inputs, RaggedTensor) and isinstance(out, RaggedTensor):
            # lets SchNetCFconv evaluate its filter network from the distances inside the fused kernel (schnet_conv.py)
            out.provenance = ("gauss", inputs, self.gamma, self.offset, self.distance / float(self.bins), self.bins)
        return out

    def get_config(self):
        config = super().get_config()
        config.update({"bins": self.bins, "distance": self.distance, "offset": self.offset, "sigma": self.sigma})
        return config


@register_keras_serializable(package="kgcnn", name="BesselBasisLayer")
class BesselBasisLayer(GraphBaseLayer):
    """kgcnn/layers/geom.py:717-805: env(d/c) * sin(f_n d/c), trainable frequencies f_n = n pi."""

    def __init__(self, num_radial, cutoff, envelope_exponent=5, envelope_type="poly", **kwargs):
        super().__init__(**kwargs)
        self.num_radial = num_radial
        self.cutoff = cutoff
        self.inv_cutoff = float(np.float32(1 / cutoff))                # tf.constant(..., float32), geom.py:758
        self.envelope_exponent = envelope_exponent
        self.envelope_type = str(envelope_type)
        if self.envelope_type not in ["poly"]:
            raise ValueError("Unknown envelope type '%s' in `BesselBasisLayer`." % self.envelope_type)
        self.frequencies = torch.nn.Parameter(
            torch.tensor(np.pi * np.arange(1, num_radial + 1, dtype=np.float32)))   # geom.py:766-770

    def envelope(self, inputs):
        p = self.envelope_exponent + 1
        a = -(p + 1) * (p + 2) / 2
        b = p * (p + 2)
        c = -p * (p + 1) / 2
        env_val = 1.0 / inputs + a * inputs ** (p - 1) + b * inputs ** p + c * inputs ** (p + 1)
        return torch.where(inputs < 1, env_val, torch.zeros_like(inputs))

    def expand_bessel_basis(self, inputs):
        if self.frequencies.device != inputs.device:
            self.to(inputs.device)
        d_scaled = inputs * self.inv_cutoff
        d_cutoff = self.envelope(d_scaled)
        return d_cutoff * torch.sin(self.frequencies * d_scaled)

    def call(self, inputs, **kwargs):
        fused = _distance_source(inputs) if isinstance(inputs, RaggedTensor) else None
        trainable = self.frequencies.requires_grad and torch.is_grad_enabled()
        if fused is not None and self.num_radial <= 64:
            src, ri, rj, eps = fused
            if self.frequencies.device != src.values.device:
                self.to(src.values.device)
            spec = F.GeometrySpec(ri, rj, ("rbf",), basis=2, R=self.num_radial, eps=eps,
                                  basis_params=(self.inv_cutoff, self.envelope_exponent + 1, 0.0),
                                  frequencies=self.frequencies.detach())
            # trainable frequencies (geom.py:766-770): d/d f_n through the energy and through the forces comes from
            # kgcnn_bessel_frequency_grad (edge reduction inside the kernel), not from an eager sin / cos graph
            return inputs.with_values(F.edge_geometry(src.values, spec, self.frequencies if trainable else None)[0])
        return self.map_values(self.expand_bessel_basis, inputs)   # distances not produced by the fused geometry

    def get_config(self):
        config = super().get_config()
        config.update({"num_radial": self.num_radial, "cutoff": self.cutoff,
                       "envelope_exponent": self.envelope_exponent, "envelope_type": self.envelope_type})
        return config


@register_keras_serializable(package="kgcnn", name="CosCutOffEnvelope")
class CosCutOffEnvelope(GraphBaseLayer):
    """kgcnn/layers/geom.py:809-856: 0.5 (cos(pi clip(d, -c, c) / c) + 1); cutoff None -> 1e8."""

    def __init__(self, cutoff, **kwargs):
        super().__init__(**kwargs)
        self.cutoff = float(np.abs(cutoff)) if cutoff is not None else 1e8

    @staticmethod
    def _compute_cutoff_envelope(inputs, cutoff):
        fc = torch.clamp(inputs, -cutoff, cutoff)
        return (torch.cos(fc * np.pi / cutoff) + 1) * 0.5

    def call(self, inputs, **kwargs):
        fused = _distance_source(inputs) if isinstance(inputs, RaggedTensor) else None
        if fused is not None:
            src, ri, rj, eps = fused
            spec = F.GeometrySpec(ri, rj, ("fc",), eps=eps, cos_cutoff=self.cutoff)
            return inputs.with_values(F.edge_geometry(src.values, spec)[0])
        return self.map_values(self._compute_cutoff_envelope, inputs, cutoff=self.cutoff)

    def get_config(self):
        config = super().get_config()
        config.update({"cutoff": self.cutoff})
        return config


@register_keras_serializable(package="kgcnn", name="CosCutOff")
class CosCutOff(GraphBaseLayer):
    """kgcnn/layers/geom.py:860-907: envelope(d) * d."""

    def __init__(self, cutoff, **kwargs):
        super().__init__(**kwargs)
        self.cutoff = float(np.abs(cutoff)) if cutoff is not None else 1e8

    def call(self, inputs, **kwargs):
        return self.map_values(
            lambda v, cutoff: CosCutOffEnvelope._compute_cutoff_envelope(v, cutoff) * v, inputs, cutoff=self.cutoff)

    def get_config(self):
        config = super().get_config()
        config.update({"cutoff": self.cutoff})
        return config
