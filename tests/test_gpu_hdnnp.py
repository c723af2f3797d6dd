"""ACSF G2/G4 through the CUDA path against the reference's golden vectors, and the HDNNP2nd energy+force model
against the fp64 oracle (BASELINE configs[3])."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from helpers import RTOL, device_ragged, oracle_ragged, rel_err
from oracle import models as OM

pytestmark = pytest.mark.gpu


def _golden_inputs(angle_key="angle_index"):
    d = np.load(os.path.join(GOLDEN, "acsf.npz"))
    z = device_ragged(d["atomic_number"], d["atomic_number_lengths"])
    xyz = device_ragged(d["positions"].astype(np.float32), d["position_lengths"])
    eij = device_ragged(d["edge_index"], d["edge_index_lengths"])
    ijk = device_ragged(d[angle_key], d[angle_key + "_lengths"])
    return d, z, xyz, eij, ijk


def test_acsf_g2_reference_golden_vector():
    # test/test_hdnnp.py:103-111
    from gcnn_keras_b200.layers.conv.acsf_conv import ACSFG2
    d, z, xyz, eij, _ = _golden_inputs()
    layer = ACSFG2(**ACSFG2.make_param_table(eta=d["g2_eta"].tolist(), rs=d["g2_rs"].tolist(), rc=float(d["g2_rc"]),
                                             elements=d["g2_elements"].tolist()))
    out = layer([z, xyz, eij])
    assert np.all(np.abs(out[2][0].cpu().numpy() - d["expected_g2"]) < 1e-4)


def test_acsf_g4_reference_golden_vector():
    # test/test_hdnnp.py:113-134
    from gcnn_keras_b200.layers.conv.acsf_conv import ACSFG4
    d, z, xyz, _, ijk = _golden_inputs()
    layer = ACSFG4(**ACSFG4.make_param_table(eta=d["g4_eta"].tolist(), zeta=d["g4_zeta"].tolist(),
                                             lamda=d["g4_lamda"].tolist(), rc=float(d["g4_rc"]),
                                             elements=d["g4_elements"].tolist(),
                                             multiplicity=float(d["g4_multiplicity"])))
    out = layer([z, xyz, ijk])
    assert np.all(np.abs(out[2][0].cpu().numpy() - d["expected_g4"]) < 1e-4)


def _setup(batch=3, n_atoms=12, seed=0):
    from gcnn_keras_b200.literature.HDNNP2nd import make_model_behler
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import HDNNP_C4_CONFIG, hdnnp_c4_batch
    torch.manual_seed(seed)
    torch.backends.cuda.matmul.allow_tf32 = False
    b = hdnnp_c4_batch(batch, seed, n_atoms=n_atoms, max_neighbours=8)
    model = EnergyForceModel(model_energy=make_model_behler(**HDNNP_C4_CONFIG), energy_output=0,
                             output_to_tensor=False, output_as_dict=False, output_squeeze_states=True,
                             is_physical_force=False)
    inputs = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
              device_ragged(b["index"], b["edge_len"]), device_ragged(b["angle_index"], b["angle_len"])]
    return HDNNP_C4_CONFIG, b, model, inputs


def test_hdnnp2nd_energy_forces_and_training_gradients():
    cfg, b, model, inputs = _setup()
    eng, force = model(inputs, training=True)
    params = {k: v.detach().cpu().double().requires_grad_(True)
              for k, v in model.energy_model.state_dict().items() if v.is_floating_point()}
    z = oracle_ragged(b["z"], b["node_len"])
    eij = oracle_ragged(b["index"], b["edge_len"])
    ijk = oracle_ragged(b["angle_index"], b["angle_len"])
    xyz = oracle_ragged(b["xyz"], b["node_len"], torch.float64)

    def energy_fn(x):
        return OM.hdnnp2nd_energy(params, z, x, eij, ijk, cfg["g2_kwargs"], cfg["g4_kwargs"],
                                  cfg["mlp_kwargs"]["activation"])

    o_eng, o_force = OM.energy_force(energy_fn, xyz)
    o_force = o_force.squeeze(-1)
    assert rel_err(eng.detach().cpu().numpy(), o_eng.detach().numpy()) < RTOL
    assert rel_err(force.values.detach().cpu().numpy(), o_force.detach().numpy()) < RTOL
    e_t, f_t = torch.tensor(b["energy"]).cuda(), torch.tensor(b["force"]).cuda()
    loss = (1 / 200) * torch.mean((eng - e_t) ** 2) + (199 / 200) * torch.mean((force.values - f_t) ** 2)
    named = [(n, p) for n, p in model.energy_model.named_parameters() if p.requires_grad]
    from gcnn_keras_b200.functional import no_coordinate_hessian
    with no_coordinate_hessian():
        grads = torch.autograd.grad(loss, [p for _, p in named], allow_unused=True)
    o_loss = OM.force_training_loss(o_eng, o_force, torch.tensor(b["energy"], dtype=torch.float64),
                                    torch.tensor(b["force"], dtype=torch.float64))
    o_grads = torch.autograd.grad(o_loss, [params[n] for n, _ in named], allow_unused=True)
    # the same op chain in fp32 (the "TF-CPU-like" comparand): where fp32 itself cannot reach 1e-5 of the truth
    # (sums over 36 atoms with cancellation), we must be no worse than twice its error
    p32 = {k: v.detach().float().requires_grad_(True) for k, v in params.items()}
    xyz32 = oracle_ragged(b["xyz"], b["node_len"], torch.float32)
    e32, f32 = OM.energy_force(lambda x: OM.hdnnp2nd_energy(p32, z, x, eij, ijk, cfg["g2_kwargs"], cfg["g4_kwargs"],
                                                             cfg["mlp_kwargs"]["activation"]), xyz32)
    l32 = OM.force_training_loss(e32, f32.squeeze(-1), torch.tensor(b["energy"]), torch.tensor(b["force"]))
    g32 = torch.autograd.grad(l32, [p32[n] for n, _ in named], allow_unused=True)
    for (n, _), g, og, o32 in zip(named, grads, o_grads, g32):
        err, err32 = rel_err(g.cpu().numpy(), og.numpy()), rel_err(o32.numpy(), og.numpy())
        assert err < max(RTOL, 2 * err32), (n, err, err32)

def test_fused_acsf_matches_unfused_chain_and_guards_hessian():
    """Fused G / J^T v / J u kernels vs the op-by-op chain (same GPU), incl. unmapped elements and unsorted triples."""
    from gcnn_keras_b200.layers.conv import acsf_conv
    from gcnn_keras_b200.layers.conv.acsf_conv import ACSFG2, ACSFG4
    d, z, xyz, eij, ijk = _golden_inputs()
    rng = np.random.default_rng(3)
    g2 = ACSFG2(**ACSFG2.make_param_table(eta=[0.1, 0.3, 0.9], rs=[0.0, 1.5, 3.0], rc=6.0, elements=[1, 6, 16]))
    g4 = ACSFG4(**ACSFG4.make_param_table(eta=[0.05, 0.3], zeta=[1.0, 2.0, 8.0], lamda=[-1.0, 1.0], rc=6.0,
                                          elements=[1, 6, 16], multiplicity=2.0))
    perm = torch.tensor(np.concatenate([o + rng.permutation(n) for o, n in zip(
        np.concatenate([[0], np.cumsum(d["angle_index_lengths"])[:-1]]), d["angle_index_lengths"])])).cuda()
    ijk_shuffled = ijk.with_values(ijk.values[perm])
    for layer, idx in ((g2, eij), (g4, ijk), (g4, ijk_shuffled)):
        res = {}
        for fused in (True, False):
            acsf_conv.FUSED = fused
            x = xyz.values.clone().requires_grad_(True)
            out = layer([z, xyz.with_values(x), idx]).values
            cot = torch.tensor(np.random.default_rng(4).standard_normal(tuple(out.shape)), dtype=torch.float32).cuda()
            wgt = cot.clone().requires_grad_(True)
            (fx,) = torch.autograd.grad(out, x, wgt, create_graph=True)          # J^T w
            h = torch.tensor(np.random.default_rng(5).standard_normal(tuple(fx.shape)), dtype=torch.float32).cuda()
            if fused:
                with pytest.raises(NotImplementedError):                         # d2/dx2 is guarded, not silently 0
                    torch.autograd.grad(fx, [wgt], h, retain_graph=True)
                from gcnn_keras_b200.functional import no_coordinate_hessian
                with no_coordinate_hessian():
                    (jh,) = torch.autograd.grad(fx, [wgt], h)                    # J h
            else:
                (jh,) = torch.autograd.grad(fx, [wgt], h)
            res[fused] = (out.detach(), fx.detach(), jh)
        acsf_conv.FUSED = True
        for a, b in zip(res[True], res[False]):
            assert rel_err(a.cpu().numpy(), b.cpu().numpy()) < 2e-5


def test_g4_factorised_table_matches_general_loop():
    """The factorised parameter loop (distinct Gaussian / angular terms evaluated once per triple) against the
    general per-parameter loop of the same kernel: G, J^T v and J u, ACSF (R = 10, m = 50) configuration."""
    from gcnn_keras_b200 import functional as F
    cfg, b, model, inputs = _setup(batch=4, n_atoms=20, seed=3)
    layer = [m for m in model.energy_model.modules() if type(m).__name__ == "ACSFG4"][0]
    z, xyz, _, ijk = inputs

    def three_maps():
        x = xyz.values.clone().requires_grad_(True)
        out = layer([z, xyz.with_values(x), ijk]).values
        g = torch.Generator(device="cuda").manual_seed(0)
        cot = torch.randn(out.shape, device="cuda", generator=g)
        vjp, = torch.autograd.grad(out, x, cot)
        return out.detach(), vjp.detach()

    F.G4_FACTORISE = True
    fac = three_maps()
    F.G4_FACTORISE = False
    try:
        gen = three_maps()
    finally:
        F.G4_FACTORISE = True
    for a, c in zip(fac, gen):
        assert rel_err(a.cpu().numpy(), c.cpu().numpy()) < 2e-6
    # J u through the training path: force-loss weight gradients agree between the two loops
    def weight_grads():
        eng, force = model(inputs, training=True)
        loss = (force.values ** 2).mean() + (eng ** 2).mean()
        params = [p for p in model.energy_model.parameters() if p.requires_grad]
        with F.no_coordinate_hessian():
            return [g for g in torch.autograd.grad(loss, params, allow_unused=True) if g is not None]
    g_fac = weight_grads()
    F.G4_FACTORISE = False
    try:
        g_gen = weight_grads()
    finally:
        F.G4_FACTORISE = True
    assert len(g_fac) == len(g_gen) and len(g_fac) > 0
    for a, c in zip(g_fac, g_gen):
        assert rel_err(a.cpu().numpy(), c.cpu().numpy()) < 1e-5
