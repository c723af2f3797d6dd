"""wACSFRad / wACSFAng of kgcnn/layers/conv/wacsf_conv.py (weighted atom-centred symmetry functions).

Same constructor arguments, ``call([z, xyz, ij|ijk(, w)])`` and output layout (batch, [N], m) as the reference.
Differences to ACSFG2/G4 (acsf_conv.py): the parameter set is looked up by the CENTRE atom Z_i in a
(118, m, .) table, neighbours are not split by element but weighted by g = Z_j (radial) / h = Z_j Z_k (angular)
or an external weight, the angular Gaussians are shifted by mu, and pooling is a plain receiver sum.  The fused
path runs the ACSF kernels with one relation, ``tset_atom = Z_i`` and a per-pair / per-triple weight, so no
[M, m] / [A, m] tensor reaches HBM; ``acsf_conv.FUSED = False`` selects the op-by-op chain.
"""
import math
import os

import numpy as np
import torch

from ... import functional as F
from ...plan import get_plan
from ..base import GraphBaseLayer, register_keras_serializable
from . import acsf_conv

_DEFAULTS = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "wacsf_defaults.npz"))
radial_eta_mu_defaults = _DEFAULTS["radial_eta_mu_rc"]                                 # wacsf_conv.py:12-19 (+ per-element rows)
angular_eta_mu_lambda_zeta_defaults = _DEFAULTS["angular_eta_mu_lambda_zeta_rc"]       # wacsf_conv.py:20-25


def _fc(r, cutoff):
    """wacsf_conv.py:146-151."""
    return (torch.cos(torch.clamp(r, -cutoff, cutoff) * math.pi / cutoff) + 1.0) * 0.5


def _distance(a, b, eps):
    d = a - b
    return torch.sqrt(torch.relu(torch.sum(d * d, dim=1, keepdim=True)) + eps)


def _centre_set(zv, n_sets):
    """Row of the parameter table for every atom as a centre; atoms outside the table contribute nothing."""
    bad = (zv < 0) | (zv >= n_sets)
    return torch.where(bad, torch.full_like(zv, -1), zv).to(torch.int32), bad


@register_keras_serializable(package="kgcnn", name="wACSFRad")
class wACSFRad(GraphBaseLayer):
    """kgcnn/layers/conv/wacsf_conv.py:97-213: W_i[k] = sum_j g(Z_j) exp(-eta_k(Z_i) (r_ij - mu_k(Z_i))^2) f_c(r_ij)."""

    def __init__(self, eta_mu=None, cutoff=8.0, add_eps=False, use_external_weights=False, param_constraint=None,
                 param_regularizer=None, param_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.cutoff = cutoff
        self.add_eps = add_eps
        if eta_mu is None:
            eta_mu = radial_eta_mu_defaults[:, :, :2]
        self.eta_mu = np.array(eta_mu, dtype="float").tolist()
        self.use_external_weights = use_external_weights
        self._cfg = {"param_constraint": param_constraint, "param_regularizer": param_regularizer,
                     "param_initializer": param_initializer}
        table = torch.tensor(np.array(eta_mu), dtype=torch.float32)
        self._weight_eta_mu = torch.nn.Parameter(table, requires_grad=False)          # trainable=False, :140-148
        fused = torch.cat([table, torch.full_like(table[..., :1], float(cutoff))], dim=-1)
        self.register_buffer("_table_eta_mu_rc", fused.contiguous(), persistent=False)

    def call(self, inputs, mask=None, **kwargs):
        if self.use_external_weights:
            z, xyz, eij, w = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
        else:
            z, xyz, eij = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
            w = None
        if self._weight_eta_mu.device != xyz.values.device:
            self.to(xyz.values.device)
        plan = get_plan(eij, xyz)
        ri, rj = plan.relation(0), plan.relation(1)
        zv = z.values.to(torch.int64)
        eps = 1e-7 if self.add_eps else 0.0
        if w is None:                                                               # :186-189  g(Z_j) = Z_j
            wv = F.gather(zv.to(torch.float32).reshape(-1, 1), rj)
        else:
            wv = w.values.to(torch.float32).reshape(-1, 1)
        n_sets, m = int(self._weight_eta_mu.shape[0]), int(self._weight_eta_mu.shape[1])
        if acsf_conv.FUSED and not wv.requires_grad:
            tset, _ = _centre_set(zv, n_sets)
            spec = F.ACSFSpec(2, ri, self._table_eta_mu_rc, 1, m, eps, rel_atom=torch.zeros_like(tset),
                              tset_atom=tset, rel_send=rj, weight=wv.reshape(-1).contiguous())
            return xyz.with_values(F.acsf(xyz.values, spec))
        xi, xj = F.gather(xyz.values, ri), F.gather(xyz.values, rj)                 # :191
        rij = _distance(xi, xj, eps)                                                # :192
        fc = _fc(rij, self.cutoff)                                                  # :193
        zi = F.gather(zv, ri)                                                       # :194
        ok = (zi >= 0) & (zi < n_sets)
        params = self._weight_eta_mu[torch.where(ok, zi, torch.zeros_like(zi))]     # :154  [M,m,2]
        gij = torch.exp(-(torch.square(rij - params[..., 1]) * params[..., 0]))     # :157-158
        rep = gij * fc * wv                                                         # :196
        rep = torch.where(ok.unsqueeze(-1), rep, torch.zeros_like(rep))
        return xyz.with_values(F.segment_reduce(rep, ri, F.OP_SUM))                 # :197

    def get_config(self):
        config = super().get_config()
        config.update({"eta_mu": self.eta_mu, "cutoff": self.cutoff,
                       "use_external_weights": self.use_external_weights, "add_eps": self.add_eps})
        config.update(self._cfg)
        return config


@register_keras_serializable(package="kgcnn", name="wACSFAng")
class wACSFAng(GraphBaseLayer):
    """kgcnn/layers/conv/wacsf_conv.py:217-367: angular wACSF on triples (i, j, k), weight h = Z_j Z_k."""

    def __init__(self, eta_mu_lambda_zeta=None, cutoff=8.0, add_eps=False, use_external_weights=False,
                 param_initializer="zeros", param_regularizer=None, param_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self.add_eps = add_eps
        self.cutoff = cutoff
        if eta_mu_lambda_zeta is None:
            eta_mu_lambda_zeta = angular_eta_mu_lambda_zeta_defaults[:, :, :4]
        self.eta_mu_lambda_zeta = np.array(eta_mu_lambda_zeta, dtype="float").tolist()
        self.use_external_weights = use_external_weights
        self._cfg = {"param_constraint": param_constraint, "param_regularizer": param_regularizer,
                     "param_initializer": param_initializer}
        table = torch.tensor(np.array(eta_mu_lambda_zeta), dtype=torch.float32)      # (118, m, 4): eta, mu, lambda, zeta
        self._weight_eta_mu_lambda_zeta = torch.nn.Parameter(table, requires_grad=False)
        # kernel layout (eta, zeta, lambda, Rc) + separate mu table
        fused = torch.stack([table[..., 0], table[..., 3], table[..., 2],
                             torch.full_like(table[..., 0], float(cutoff))], dim=-1)
        self.register_buffer("_table_eta_zeta_lambda_rc", fused.contiguous(), persistent=False)
        self.register_buffer("_table_mu", table[..., 1].contiguous(), persistent=False)
        fac = F.g4_factor_tables(fused.numpy(), table[..., 1].numpy())
        self.register_buffer("_factor_table", None if fac is None else torch.from_numpy(fac[0]), persistent=False)
        self.register_buffer("_factor_index", None if fac is None else torch.from_numpy(fac[1]), persistent=False)

    def call(self, inputs, mask=None, **kwargs):
        if self.use_external_weights:
            z, xyz, ijk, w = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
        else:
            z, xyz, ijk = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
            w = None
        if self._weight_eta_mu_lambda_zeta.device != xyz.values.device:
            self.to(xyz.values.device)
        plan = get_plan(ijk, xyz)
        ri, rj, rk = plan.relation(0), plan.relation(1), plan.relation(2)
        zv = z.values.to(torch.int64)
        eps = 1e-7 if self.add_eps else 0.0
        if w is None:                                                               # :332-338  h = Z_j Z_k
            zf = zv.to(torch.float32).reshape(-1, 1)
            wv = F.gather(zf, rj) * F.gather(zf, rk)
        else:
            wv = w.values.to(torch.float32).reshape(-1, 1)
        n_sets, m = int(self._table_mu.shape[0]), int(self._table_mu.shape[1])
        if acsf_conv.FUSED and not wv.requires_grad:
            tset, _ = _centre_set(zv, n_sets)
            spec = F.ACSFSpec(4, ri, self._table_eta_zeta_lambda_rc, 1, m, eps, tset_atom=tset,
                              z_atom=torch.zeros_like(tset), zmax=1,
                              rev_pair=torch.zeros(1, dtype=torch.int32, device=tset.device), rel_j=rj, rel_k=rk,
                              multiplicity=0.0, uniform_rc=True, weight=wv.reshape(-1).contiguous(),
                              mu_table=self._table_mu,
                              factor=None if self._factor_table is None else (self._factor_table, self._factor_index))
            return xyz.with_values(F.acsf(xyz.values, spec))
        xi, xj, xk = (F.gather(xyz.values, r) for r in (ri, rj, rk))               # :339
        rij, rik, rjk = _distance(xi, xj, eps), _distance(xi, xk, eps), _distance(xj, xk, eps)   # :340-342
        fij, fik, fjk = _fc(rij, self.cutoff), _fc(rik, self.cutoff), _fc(rjk, self.cutoff)      # :343-345
        zi = F.gather(zv, ri)                                                       # :346
        ok = (zi >= 0) & (zi < n_sets)
        params = self._weight_eta_mu_lambda_zeta[torch.where(ok, zi, torch.zeros_like(zi))]     # [A,m,4]
        eta, mu, lamda, zeta = params[..., 0], params[..., 1], params[..., 2], params[..., 3]
        gij = torch.exp(-(torch.square(rij - mu) * eta))                            # :347-349
        gik = torch.exp(-(torch.square(rik - mu) * eta))
        gjk = torch.exp(-(torch.square(rjk - mu) * eta))
        vij, vik = xi - xj, xi - xk                                                 # :350-351
        cos_theta = torch.sum(vij * vik, dim=-1, keepdim=True) / rij / rik          # :288
        pow_cos = torch.pow(cos_theta * lamda + 1.0, zeta)                          # :289-290
        rep = pow_cos * gij * gik * gjk * fij * fik * fjk * wv                      # :353
        rep = torch.where(ok.unsqueeze(-1), rep, torch.zeros_like(rep))
        pooled = F.segment_reduce(rep, ri, F.OP_SUM)                                # :354
        okn = (zv >= 0) & (zv < n_sets)
        zeta_n = self._weight_eta_mu_lambda_zeta[torch.where(okn, zv, torch.zeros_like(zv))][..., 3]
        return xyz.with_values(torch.pow(torch.full_like(pooled, 2.0), 1.0 - zeta_n) * pooled)   # :297-302, :356

    def get_config(self):
        config = super().get_config()
        config.update({"eta_mu_lambda_zeta": self.eta_mu_lambda_zeta, "cutoff": self.cutoff,
                       "use_external_weights": self.use_external_weights, "add_eps": self.add_eps})
        config.update(self._cfg)
        return config
