"""wACSFRad / wACSFAng (kgcnn/layers/conv/wacsf_conv.py) through the fused CUDA kernels against the fp64 oracle:
values, forces (VJP) and the force-training direction (VJP of the VJP with respect to the cotangent)."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, rel_err
from oracle import conv as OC

pytestmark = pytest.mark.gpu


def _batch(seed=0, batch=3, n_atoms=10):
    from gcnn_keras_b200.synthetic import hdnnp_c4_batch
    b = hdnnp_c4_batch(batch, seed, n_atoms=n_atoms, max_neighbours=6)
    b["xyz"] = (b["xyz"] * 0.6).astype(np.float32)       # distances around the default mu grid (0.5 .. 7.5)
    return b


def _three_maps(layer_out_fn, xyz_leaf, cot, h):
    """G, J^T cot, and d/dcot <J^T cot, h> = J h."""
    G = layer_out_fn()
    (vjp,) = torch.autograd.grad(G, xyz_leaf, cot, create_graph=True)
    (jvp,) = torch.autograd.grad(vjp, cot, h)
    return G.detach(), vjp.detach(), jvp.detach()


@pytest.mark.parametrize("kind", ["rad", "ang"])
@pytest.mark.parametrize("external", [False, True])
def test_wacsf_matches_oracle(kind, external):
    from gcnn_keras_b200.functional import no_coordinate_hessian
    from gcnn_keras_b200.layers.conv.wacsf_conv import wACSFAng, wACSFRad
    b = _batch()
    rng = np.random.default_rng(1)
    idx_key, len_key = ("index", "edge_len") if kind == "rad" else ("angle_index", "angle_len")
    n_el = len(b[idx_key])
    w_np = rng.uniform(0.5, 2.0, size=(n_el, 1)).astype(np.float32) if external else None
    if kind == "rad":
        layer = wACSFRad(use_external_weights=external)
    else:   # the published angular defaults all have mu = 0: shift them so the mu path of the kernel is exercised
        from gcnn_keras_b200.layers.conv.wacsf_conv import angular_eta_mu_lambda_zeta_defaults
        tab = angular_eta_mu_lambda_zeta_defaults[:, :, :4].copy()
        tab[..., 1] = rng.uniform(0.0, 1.5, size=tab[..., 1].shape)
        layer = wACSFAng(eta_mu_lambda_zeta=tab, use_external_weights=external)
    m = 22 if kind == "rad" else 10
    N = len(b["z"])
    cot_np, h_np = rng.standard_normal((N, m)), rng.standard_normal((N, 3))

    # ---- CUDA path
    x_d = torch.tensor(b["xyz"], device="cuda", requires_grad=True)
    from gcnn_keras_b200 import RaggedTensor
    lens = torch.as_tensor(b["node_len"]).cuda()
    inputs = [device_ragged(b["z"], b["node_len"]), RaggedTensor.from_row_lengths(x_d, lens),
              device_ragged(b[idx_key], b[len_key])]
    if external:
        inputs.append(device_ragged(w_np, b[len_key]))
    cot = torch.tensor(cot_np, dtype=torch.float32, device="cuda", requires_grad=True)
    h = torch.tensor(h_np, dtype=torch.float32, device="cuda")
    with no_coordinate_hessian():
        G, vjp, jvp = _three_maps(lambda: layer(inputs).values, x_d, cot, h)

    # ---- fp64 oracle
    x_o = torch.tensor(b["xyz"], dtype=torch.float64, requires_grad=True)
    z_o = oracle_ragged(b["z"], b["node_len"])
    i_o = oracle_ragged(b[idx_key], b[len_key])
    w_o = oracle_ragged(w_np, b[len_key], torch.float64) if external else None
    xr = oracle_ragged(b["xyz"], b["node_len"], torch.float64).with_values(x_o)
    fn = OC.wacsf_rad if kind == "rad" else OC.wacsf_ang
    table = np.array(layer.eta_mu if kind == "rad" else layer.eta_mu_lambda_zeta)
    cot_o = torch.tensor(cot_np, dtype=torch.float64, requires_grad=True)
    oG, ovjp, ojvp = _three_maps(lambda: fn(z_o, xr, i_o, table, cutoff=8.0, w=w_o).values, x_o, cot_o,
                                 torch.tensor(h_np, dtype=torch.float64))
    assert float(oG.abs().max()) > 1e-3            # the comparison is not vacuous
    assert rel_err(G.cpu().numpy(), oG.numpy()) < RTOL
    assert rel_err(vjp.cpu().numpy(), ovjp.numpy()) < RTOL
    assert rel_err(jvp.cpu().numpy(), ojvp.numpy()) < RTOL


@pytest.mark.parametrize("kind", ["rad", "ang"])
def test_wacsf_fused_matches_unfused_chain(kind):
    from gcnn_keras_b200.layers.conv import acsf_conv
    from gcnn_keras_b200.layers.conv.wacsf_conv import wACSFAng, wACSFRad
    from gcnn_keras_b200 import RaggedTensor
    b = _batch(seed=3)
    idx_key, len_key = ("index", "edge_len") if kind == "rad" else ("angle_index", "angle_len")
    layer = (wACSFRad if kind == "rad" else wACSFAng)(add_eps=True)
    lens = torch.as_tensor(b["node_len"]).cuda()
    outs = []
    try:
        for fused in (True, False):
            acsf_conv.FUSED = fused
            x = torch.tensor(b["xyz"], device="cuda", requires_grad=True)
            inputs = [device_ragged(b["z"], b["node_len"]), RaggedTensor.from_row_lengths(x, lens),
                      device_ragged(b[idx_key], b[len_key])]
            G = layer(inputs).values
            (g,) = torch.autograd.grad((G ** 2).sum(), x)
            outs.append((G.detach().cpu().numpy(), g.cpu().numpy()))
    finally:
        acsf_conv.FUSED = True
    assert rel_err(outs[0][0], outs[1][0]) < RTOL
    assert rel_err(outs[0][1], outs[1][1]) < 5 * RTOL     # both sides are fp32 here


def test_wacsf_atoms_outside_the_table_contribute_nothing():
    from gcnn_keras_b200.layers.conv.wacsf_conv import wACSFRad
    b = _batch(seed=5)
    z = b["z"].copy()
    z[0] = 200                                            # outside the (118, m, 2) table
    layer = wACSFRad()
    out = layer([device_ragged(z, b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
                 device_ragged(b["index"], b["edge_len"])]).values
    assert float(out[0].abs().max()) == 0.0
    assert torch.isfinite(out).all()


def test_hdnnp2nd_weighted_model_energy_and_force_run_and_are_deterministic():
    # kgcnn/literature/HDNNP2nd.py:46-129 (make_model_weighted) under EnergyForceModel
    from gcnn_keras_b200.literature.HDNNP2nd import make_model_weighted
    from gcnn_keras_b200.model.force import EnergyForceModel
    b = _batch(seed=7)
    torch.manual_seed(0)
    energy = make_model_weighted(mlp_kwargs={"units": [16, 16, 1], "num_relations": 10,
                                             "activation": ["swish", "swish", "linear"]})
    model = EnergyForceModel(model_energy=energy, energy_output=0, output_to_tensor=False, output_as_dict=False,
                             output_squeeze_states=True, is_physical_force=False)
    inputs = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
              device_ragged(b["index"], b["edge_len"]), device_ragged(b["angle_index"], b["angle_len"])]
    e1, f1 = model(inputs, training=True)
    e2, f2 = model(inputs, training=True)
    assert e1.shape == (3, 1) and f1.values.shape == (30, 3)
    assert torch.isfinite(e1).all() and torch.isfinite(f1.values).all() and float(f1.values.abs().max()) > 0
    assert torch.equal(e1, e2) and torch.equal(f1.values, f2.values)          # no atomics: bitwise repeatable
    # central finite difference of the energy along one coordinate agrees with the returned dE/dx
    x = b["xyz"].copy()
    vals = []
    for s in (+1e-2, -1e-2):
        x2 = x.copy()
        x2[4, 1] += s
        inp = [inputs[0], device_ragged(x2, b["node_len"]), inputs[2], inputs[3]]
        vals.append(float(model(inp, training=False)[0].sum()))
    fd = (vals[0] - vals[1]) / 2e-2
    assert abs(fd - float(f1.values[4, 1])) < 2e-2 * max(1.0, abs(fd))
