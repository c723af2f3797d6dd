"""ACSFG2 / ACSFG4 / ACSFConstNormalization of kgcnn/layers/conv/acsf_conv.py.

Same constructor arguments, ``make_param_table`` helpers, call signatures ``call([z, xyz, ij|ijk])`` and output
layout (batch, [N], R*m) as the reference.  The reference pools with ``tensor_scatter_nd_add`` on (receiver,
relation) (atomic scatter on GPU TF, pooling.py:658-663); here the composite key receiver*R + relation gets a
CSR and the deterministic segment kernel runs.  Neighbours whose element is not in ``element_mapping`` contribute
nothing (what TF-GPU's out-of-range gather/scatter does; TF-CPU raises) -- see oracle/conv.py:_drop_unmapped.
"""
import math

import numpy as np
import torch

from ... import functional as F
from ...plan import Relation, get_plan
from ..base import GraphBaseLayer, register_keras_serializable

_MAX_ATOMIC_NUMBER = 31
FUSED = True    # fused sm_100a kernels (G, J^T v, J u); False -> op-by-op chain (trainable parameters, Hessians)


def _reverse_mapping(element_mapping):
    rev = np.full(_MAX_ATOMIC_NUMBER, np.iinfo(np.int64).max, dtype=np.int64)      # acsf_conv.py:88-91
    for i, pos in enumerate(element_mapping):
        rev[pos] = i
    return rev


def _cos_cutoff(r, cutoff):
    fc = torch.minimum(torch.maximum(r.expand_as(cutoff), -cutoff), cutoff)         # clip_by_value
    return (torch.cos(fc * math.pi / cutoff) + 1.0) * 0.5


def _relational_sum(rep, recv, rel_map, valid, num_relations, n_nodes):
    """RelationalPoolingLocalEdges(sum) with a composite-key CSR (deterministic)."""
    rep = torch.where(valid.unsqueeze(-1), rep, torch.zeros_like(rep))
    key = (recv.ids.to(torch.int64) * num_relations + torch.where(valid, rel_map, torch.zeros_like(rel_map)))
    rel = Relation(key.to(torch.int32), n_nodes * num_relations)
    out = F.segment_reduce(rep, rel, F.OP_SUM)
    return out.reshape(n_nodes, num_relations * rep.shape[1])                       # _flatten_relations


@register_keras_serializable(package="kgcnn", name="ACSFG2")
class ACSFG2(GraphBaseLayer):
    """kgcnn/layers/conv/acsf_conv.py:17-224: G2_i[Z_j, k] = sum_j exp(-eta_k (r_ij - mu_k)^2) f_c(r_ij)."""

    _max_atomic_number = _MAX_ATOMIC_NUMBER

    def __init__(self, eta_rs_rc, element_mapping, add_eps=False, param_constraint=None, param_regularizer=None,
                 param_initializer="zeros", param_trainable=False, **kwargs):
        super().__init__(**kwargs)
        self.eta_rs_rc = np.array(eta_rs_rc, dtype="float")
        assert len(self.eta_rs_rc.shape) in [3, 4], "Require `eta_rs_rc` of shape `(N, N, m, 3)` or `(N, m, 3)`"
        self.use_target_set = (len(self.eta_rs_rc.shape) == 4)
        self.num_relations = self.eta_rs_rc.shape[1] if self.use_target_set else self.eta_rs_rc.shape[0]
        self.element_mapping = np.array(element_mapping, dtype="int")
        self.reverse_mapping = _reverse_mapping(self.element_mapping)
        self.add_eps = add_eps
        self.param_trainable = param_trainable
        self._cfg = {"param_constraint": param_constraint, "param_regularizer": param_regularizer,
                     "param_initializer": param_initializer}
        self.weight_eta_rs_rc = torch.nn.Parameter(torch.tensor(self.eta_rs_rc, dtype=torch.float32),
                                                   requires_grad=bool(param_trainable))
        self.register_buffer("weight_reverse_mapping", torch.tensor(self.reverse_mapping))

    @staticmethod
    def make_param_table(eta, rs, rc, elements, **kwargs):
        """acsf_conv.py:127-144."""
        eta_rs_rc = [(et, Rs, rc) for Rs in rs for et in eta]
        elements = np.sort(elements)
        params = np.broadcast_to(eta_rs_rc, (len(elements), len(eta_rs_rc), 3))
        return {"eta_rs_rc": params, "element_mapping": elements, **kwargs}

    def call(self, inputs, mask=None, **kwargs):
        z, xyz, eij = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
        if self.weight_eta_rs_rc.device != xyz.values.device:
            self.to(xyz.values.device)
        plan = get_plan(eij, xyz)
        ri, rj = plan.relation(0), plan.relation(1)
        zv = z.values.to(torch.int64)                                               # acsf_conv.py:200
        if FUSED and not self.param_trainable:
            zc = torch.clamp(zv, 0, _MAX_ATOMIC_NUMBER - 1)
            rel_atom = self.weight_reverse_mapping[zc]
            bad = (zv < 0) | (zv >= _MAX_ATOMIC_NUMBER) | (rel_atom < 0) | (rel_atom >= _MAX_ATOMIC_NUMBER)
            rel_atom = torch.where(bad, torch.full_like(rel_atom, -1), rel_atom).to(torch.int32)
            spec = F.ACSFSpec(2, ri, self.weight_eta_rs_rc.detach().contiguous(), self.num_relations,
                              int(self.eta_rs_rc.shape[-2]), 1e-7 if self.add_eps else 0.0, rel_atom=rel_atom,
                              tset_atom=rel_atom if self.use_target_set else None, rel_send=rj)
            return xyz.with_values(F.acsf(xyz.values, spec))
        xi, xj = F.gather(xyz.values, ri), F.gather(xyz.values, rj)                 # :201
        diff = xi - xj
        rij = torch.sqrt(torch.relu(torch.sum(diff * diff, dim=1, keepdim=True))
                         + (1e-7 if self.add_eps else 0.0))                         # :202
        zi, zj = F.gather(zv, ri), F.gather(zv, rj)                                 # :203
        zi_map, zj_map = self.weight_reverse_mapping[zi], self.weight_reverse_mapping[zj]   # :204-205
        valid = (zj_map >= 0) & (zj_map < self.num_relations)
        zj_safe = torch.where(valid, zj_map, torch.zeros_like(zj_map))
        if self.use_target_set:                                                     # :151-156
            vi = (zi_map >= 0) & (zi_map < self.weight_eta_rs_rc.shape[0])
            valid = valid & vi
            params = self.weight_eta_rs_rc[torch.where(vi, zi_map, torch.zeros_like(zi_map)), zj_safe]
        else:
            params = self.weight_eta_rs_rc[zj_safe]
        fc = _cos_cutoff(rij, params[..., 2])                                       # :207
        gij = torch.exp(-(torch.square(rij - params[..., 1]) * params[..., 0]))     # :208
        rep = gij * fc                                                              # :209
        out = _relational_sum(rep, ri, zj_map, valid, self.num_relations, plan.N)   # :210-211
        return xyz.with_values(out)

    def get_config(self):
        config = super().get_config()
        config.update({"eta_rs_rc": self.eta_rs_rc.tolist(), "element_mapping": self.element_mapping.tolist(),
                       "add_eps": self.add_eps, "param_trainable": self.param_trainable})
        config.update(self._cfg)
        return config


@register_keras_serializable(package="kgcnn", name="ACSFG4")
class ACSFG4(GraphBaseLayer):
    """kgcnn/layers/conv/acsf_conv.py:228-510: angular symmetry functions on triples (i, j, k)."""

    _max_atomic_number = _MAX_ATOMIC_NUMBER

    def __init__(self, eta_zeta_lambda_rc, element_mapping, element_pair_mapping=None, add_eps=False,
                 keep_pair_order=False, multiplicity=None, param_initializer="zeros", param_regularizer=None,
                 param_constraint=None, param_trainable=False, **kwargs):
        super().__init__(**kwargs)
        self.add_eps = add_eps
        self.multiplicity = multiplicity
        self.keep_pair_order = keep_pair_order
        self.eta_zeta_lambda_rc = np.array(eta_zeta_lambda_rc, dtype="float")
        assert len(self.eta_zeta_lambda_rc.shape) in [3, 4], "Require `eta_zeta_lambda_rc` rank 3 or 4."
        self.use_target_set = (len(self.eta_zeta_lambda_rc.shape) == 4)
        self.num_relations = self.eta_zeta_lambda_rc.shape[1 if self.use_target_set else 0]
        self.element_mapping = np.array(element_mapping, dtype="int")
        if element_pair_mapping is None:                                            # acsf_conv.py:314-325
            idx = np.expand_dims(self.element_mapping, axis=-1)
            pairs = np.concatenate([np.repeat(np.expand_dims(idx, axis=0), len(self.element_mapping), axis=0),
                                    np.repeat(np.expand_dims(idx, axis=1), len(self.element_mapping), axis=1)],
                                   axis=-1).reshape((-1, 2))
            if not self.keep_pair_order:
                pairs = np.sort(pairs, axis=-1)
                pairs = pairs[np.sort(np.unique(pairs, axis=0, return_index=True)[1])]
            self.element_pair_mapping = pairs
        else:
            self.element_pair_mapping = np.array(element_pair_mapping, dtype="int")
        assert len(self.element_pair_mapping.shape) == 2 and self.element_pair_mapping.shape[1] == 2
        assert self.element_pair_mapping.shape[0] == self.num_relations
        self.reverse_mapping = _reverse_mapping(self.element_mapping)
        rev_pair = np.full((_MAX_ATOMIC_NUMBER, _MAX_ATOMIC_NUMBER), np.iinfo(np.int64).max, dtype=np.int64)
        for i, pos in enumerate(self.element_pair_mapping):                         # :336-341
            rev_pair[pos[0], pos[1]] = i
            if not self.keep_pair_order:
                rev_pair[pos[1], pos[0]] = i
        self.reverse_pair_mapping = rev_pair
        self.param_trainable = param_trainable
        self._cfg = {"param_constraint": param_constraint, "param_regularizer": param_regularizer,
                     "param_initializer": param_initializer}
        self.weight_eta_zeta_lambda_rc = torch.nn.Parameter(
            torch.tensor(self.eta_zeta_lambda_rc, dtype=torch.float32), requires_grad=bool(param_trainable))
        self.register_buffer("weight_reverse_mapping", torch.tensor(self.reverse_mapping))
        self.register_buffer("weight_reverse_pair_mapping", torch.tensor(self.reverse_pair_mapping))
        fac = None if param_trainable else F.g4_factor_tables(self.eta_zeta_lambda_rc)
        self.register_buffer("_factor_table", None if fac is None else torch.from_numpy(fac[0]), persistent=False)
        self.register_buffer("_factor_index", None if fac is None else torch.from_numpy(fac[1]), persistent=False)

    @staticmethod
    def make_param_table(eta, zeta, lamda, rc, elements, **kwargs):
        """acsf_conv.py:382-401."""
        ezl = [[e, z, la, rc] for e in eta for z in zeta for la in lamda]
        elements = np.sort(elements)
        params = np.broadcast_to(ezl, (int(len(elements) * (len(elements) + 1) / 2), len(ezl), 4))
        return {"eta_zeta_lambda_rc": params, "element_mapping": elements, "element_pair_mapping": None, **kwargs}

    def call(self, inputs, mask=None, **kwargs):
        z, xyz, ijk = self.assert_ragged_input_rank(inputs, mask=mask, ragged_rank=1)
        if self.weight_eta_zeta_lambda_rc.device != xyz.values.device:
            self.to(xyz.values.device)
        plan = get_plan(ijk, xyz)
        ri, rj, rk = plan.relation(0), plan.relation(1), plan.relation(2)
        zv = z.values.to(torch.int64)                                               # acsf_conv.py:474
        if FUSED and not self.param_trainable:
            big = torch.iinfo(torch.int32).max
            rev_pair = torch.where(self.weight_reverse_pair_mapping > big, torch.full_like(
                self.weight_reverse_pair_mapping, -1), self.weight_reverse_pair_mapping).to(torch.int32).contiguous()
            tset = None
            if self.use_target_set:
                zc = torch.clamp(zv, 0, _MAX_ATOMIC_NUMBER - 1)
                t = self.weight_reverse_mapping[zc]
                bad = (zv < 0) | (zv >= _MAX_ATOMIC_NUMBER) | (t < 0) | (t >= _MAX_ATOMIC_NUMBER)
                tset = torch.where(bad, torch.full_like(t, -1), t).to(torch.int32)
            spec = F.ACSFSpec(4, ri, self.weight_eta_zeta_lambda_rc.detach().contiguous(), self.num_relations,
                              int(self.eta_zeta_lambda_rc.shape[-2]), 1e-7 if self.add_eps else 0.0,
                              tset_atom=tset, z_atom=zv.to(torch.int32), rev_pair=rev_pair, zmax=_MAX_ATOMIC_NUMBER,
                              rel_j=rj, rel_k=rk, multiplicity=self.multiplicity or 0.0,
                              uniform_rc=bool(np.all(self.eta_zeta_lambda_rc[..., 3] == self.eta_zeta_lambda_rc.reshape(-1, 4)[0, 3])),
                              factor=None if self._factor_table is None else (self._factor_table, self._factor_index))
            return xyz.with_values(F.acsf(xyz.values, spec))
        zi, zj, zk = F.gather(zv, ri), F.gather(zv, rj), F.gather(zv, rk)           # :475
        xi, xj, xk = (F.gather(xyz.values, r) for r in (ri, rj, rk))               # :476
        valid_z = (zj >= 0) & (zj < _MAX_ATOMIC_NUMBER) & (zk >= 0) & (zk < _MAX_ATOMIC_NUMBER)
        zjk_map = self.weight_reverse_pair_mapping[torch.where(valid_z, zj, torch.zeros_like(zj)),
                                                   torch.where(valid_z, zk, torch.zeros_like(zk))]   # :478
        valid = valid_z & (zjk_map >= 0) & (zjk_map < self.num_relations)
        zjk_safe = torch.where(valid, zjk_map, torch.zeros_like(zjk_map))
        if self.use_target_set:                                                     # :410-416
            zi_map = self.weight_reverse_mapping[zi]
            vi = (zi_map >= 0) & (zi_map < self.weight_eta_zeta_lambda_rc.shape[0])
            valid = valid & vi
            params = self.weight_eta_zeta_lambda_rc[torch.where(vi, zi_map, torch.zeros_like(zi_map)), zjk_safe]
        else:
            params = self.weight_eta_zeta_lambda_rc[zjk_safe]
        eps = 1e-7 if self.add_eps else 0.0

        def dist(a, b):                                                             # :480-482
            d = a - b
            return torch.sqrt(torch.relu(torch.sum(d * d, dim=1, keepdim=True)) + eps)

        rij, rik, rjk = dist(xi, xj), dist(xi, xk), dist(xj, xk)
        eta, zeta, lamda, rc = params[..., 0], params[..., 1], params[..., 2], params[..., 3]
        fij, fik, fjk = _cos_cutoff(rij, rc), _cos_cutoff(rik, rc), _cos_cutoff(rjk, rc)   # :483-485
        gij = torch.exp(-(torch.square(rij) * eta))                                 # :486-488
        gik = torch.exp(-(torch.square(rik) * eta))
        gjk = torch.exp(-(torch.square(rjk) * eta))
        vij, vik = xi - xj, xi - xk                                                 # :489-490
        cos_theta = torch.sum(vij * vik, dim=-1, keepdim=True) / rij / rik          # :439
        cos_term = torch.pow(cos_theta * lamda + 1.0, zeta)                         # :440-441
        scaled = torch.pow(torch.full_like(cos_term, 2.0), 1.0 - zeta) * cos_term   # :442-443
        if self.multiplicity is not None:
            scaled = scaled / self.multiplicity
        rep = scaled * gij * gik * gjk * fij * fik * fjk                            # :492
        out = _relational_sum(rep, ri, zjk_map, valid, self.num_relations, plan.N)  # :493-494
        return xyz.with_values(out)

    def get_config(self):
        config = super().get_config()
        config.update({"eta_zeta_lambda_rc": self.eta_zeta_lambda_rc, "add_eps": self.add_eps,
                       "element_mapping": self.element_mapping, "keep_pair_order": self.keep_pair_order,
                       "multiplicity": self.multiplicity, "element_pair_mapping": self.element_pair_mapping,
                       "param_trainable": self.param_trainable})
        config.update(self._cfg)
        return config


@register_keras_serializable(package="kgcnn", name="ACSFConstNormalization")
class ACSFConstNormalization(GraphBaseLayer):
    """kgcnn/layers/conv/acsf_conv.py:514-546: (x - mean) / std with constants."""

    def __init__(self, std=1.0, mean=0.0, **kwargs):
        super().__init__(**kwargs)
        self._np_std = np.array(std)
        self._np_mean = np.array(mean)

    def call(self, inputs, mask=None, **kwargs):
        def scale(v):
            mean = torch.as_tensor(self._np_mean, dtype=v.dtype, device=v.device)
            std = torch.as_tensor(self._np_std, dtype=v.dtype, device=v.device)
            return (v - mean) / std
        return self.map_values(scale, inputs)

    def get_config(self):
        config = super().get_config()
        config.update({"mean": self._np_mean.tolist(), "std": self._np_std.tolist()})
        return config
