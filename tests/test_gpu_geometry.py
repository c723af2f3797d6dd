"""Fused edge-geometry kernel (kgcnn_edge_geometry: distance, direction, Gauss / Bessel basis, cosine cutoff) through
the reference layer API against the fp64 oracle: values, forces (VJP) and the force-training direction (JVP)."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, rel_err
from oracle import layers as OL

pytestmark = pytest.mark.gpu


def _mol_batch(seed=0, n_graphs=4, n_atoms=7):
    rng = np.random.default_rng(seed)
    xyz, idx, nl, el = [], [], [], []
    for _ in range(n_graphs):
        x = rng.uniform(0, 3.0, size=(n_atoms, 3))
        pairs = np.array([[i, j] for i in range(n_atoms) for j in range(n_atoms) if i != j])
        pairs = pairs[rng.permutation(len(pairs))[: len(pairs) * 2 // 3]]          # unsorted, ragged
        xyz.append(x), idx.append(pairs), nl.append(n_atoms), el.append(len(pairs))
    return (np.concatenate(xyz).astype(np.float32), np.concatenate(idx).astype(np.int64),
            np.array(nl, dtype=np.int64), np.array(el, dtype=np.int64))


def _layers(kind):
    from gcnn_keras_b200.layers import geom as G
    if kind == "gauss":
        return G.GaussBasisLayer(bins=25, distance=5.0, sigma=0.4, offset=0.0)
    if kind == "bessel":
        lay = G.BesselBasisLayer(num_radial=20, cutoff=5.0, envelope_exponent=5)
        lay.frequencies.requires_grad_(False)
        return lay
    if kind == "cos":
        return G.CosCutOffEnvelope(cutoff=4.0)
    return None


def _device_outputs(kind, xyz_leaf, nl, idx, el, add_eps):
    from gcnn_keras_b200 import RaggedTensor
    from gcnn_keras_b200.layers import geom as G
    xyz = RaggedTensor.from_row_lengths(xyz_leaf, torch.as_tensor(nl).cuda())
    edi = device_ragged(idx, el)
    pos1, pos2 = G.NodePosition()([xyz, edi])
    if kind == "direction":
        return G.EdgeDirectionNormalized(add_eps=add_eps)([pos1, pos2]).values
    d = G.NodeDistanceEuclidean(add_eps=add_eps)([pos1, pos2])
    if kind == "distance":
        return d.values
    return _layers(kind).cuda()(d).values if kind == "bessel" else _layers(kind)(d).values


def _oracle_outputs(kind, xyz_leaf, nl, idx, el, add_eps):
    xyz = oracle_ragged(np.zeros((len(xyz_leaf), 3)), nl, torch.float64).with_values(xyz_leaf)
    edi = oracle_ragged(idx, el)
    p1, p2 = OL.node_position(xyz, edi)
    if kind == "direction":
        return OL.edge_direction_normalized(p1, p2, add_eps=add_eps).values
    d = OL.node_distance_euclidean(p1, p2, add_eps=add_eps)
    if kind == "distance":
        return d.values
    if kind == "gauss":
        return OL.gauss_basis(d, bins=25, dist=5.0, sigma=0.4, offset=0.0).values
    if kind == "bessel":
        return OL.bessel_basis(d, 20, 5.0, 5).values
    return OL.cos_cutoff_envelope(d, 4.0).values


def _three(fn, leaf, cot, h):
    out = fn()
    (vjp,) = torch.autograd.grad(out, leaf, cot, create_graph=True)
    (jvp,) = torch.autograd.grad(vjp, cot, h)
    return out.detach(), vjp.detach(), jvp.detach()


@pytest.mark.parametrize("kind", ["distance", "direction", "gauss", "bessel", "cos"])
@pytest.mark.parametrize("add_eps", [False, True])
def test_fused_geometry_layers_match_oracle(kind, add_eps):
    from gcnn_keras_b200 import _lib
    from gcnn_keras_b200.functional import no_coordinate_hessian
    x_np, idx, nl, el = _mol_batch()
    rng = np.random.default_rng(3)
    x_d = torch.tensor(x_np, device="cuda", requires_grad=True)
    _lib.launch_count_reset()
    probe = _device_outputs(kind, x_d, nl, idx, el, add_eps)
    cot_np, h_np = rng.standard_normal(tuple(probe.shape)), rng.standard_normal(x_np.shape)
    cot = torch.tensor(cot_np, dtype=torch.float32, device="cuda", requires_grad=True)
    with no_coordinate_hessian():
        out, vjp, jvp = _three(lambda: _device_outputs(kind, x_d, nl, idx, el, add_eps), x_d, cot,
                               torch.tensor(h_np, dtype=torch.float32, device="cuda"))
    x_o = torch.tensor(x_np, dtype=torch.float64, requires_grad=True)
    cot_o = torch.tensor(cot_np, dtype=torch.float64, requires_grad=True)
    o_out, o_vjp, o_jvp = _three(lambda: _oracle_outputs(kind, x_o, nl, idx, el, add_eps), x_o, cot_o,
                                 torch.tensor(h_np, dtype=torch.float64))
    assert rel_err(out.cpu().numpy(), o_out.numpy()) < RTOL
    assert rel_err(vjp.cpu().numpy(), o_vjp.numpy()) < RTOL
    assert rel_err(jvp.cpu().numpy(), o_jvp.numpy()) < RTOL


def test_fused_geometry_runs_the_kernel_and_falls_back_without_provenance():
    from gcnn_keras_b200 import RaggedTensor, _lib
    from gcnn_keras_b200.layers import geom as G
    x_np, idx, nl, el = _mol_batch(seed=1)
    xyz, edi = device_ragged(x_np, nl), device_ragged(idx, el)
    pos1, pos2 = G.NodePosition()([xyz, edi])
    d = G.NodeDistanceEuclidean()([pos1, pos2])
    assert d.provenance is not None and d.provenance[0] == "distance"
    fused = G.GaussBasisLayer(bins=25, distance=5.0)(d).values
    plain = G.GaussBasisLayer(bins=25, distance=5.0)(d.with_values(d.values)).values      # provenance dropped
    assert rel_err(fused.cpu().numpy(), plain.cpu().numpy()) < RTOL
    # same answer with the switch off (op-by-op chain)
    G.FUSED = False
    try:
        d2 = G.NodeDistanceEuclidean()([pos1, pos2])
        assert d2.provenance is None
    finally:
        G.FUSED = True
    assert rel_err(d.values.cpu().numpy(), d2.values.cpu().numpy()) < RTOL


def test_coordinate_hessian_term_through_the_fused_geometry_matches_op_chain():
    # d/dx of a function of the force: needs d2/dx2 through the geometry (not used by training, but exact)
    from gcnn_keras_b200 import RaggedTensor
    from gcnn_keras_b200.layers import geom as G
    x_np, idx, nl, el = _mol_batch(seed=2)
    w = torch.randn(25, device="cuda")
    res = []
    try:
        for fused in (True, False):
            G.FUSED = fused
            x = torch.tensor(x_np, device="cuda", requires_grad=True)
            xyz = RaggedTensor.from_row_lengths(x, torch.as_tensor(nl).cuda())
            pos1, pos2 = G.NodePosition()([xyz, device_ragged(idx, el)])
            rbf = G.GaussBasisLayer(bins=25, distance=5.0)(G.NodeDistanceEuclidean()([pos1, pos2])).values
            energy = (torch.tanh(rbf) * w).sum()
            (force,) = torch.autograd.grad(energy, x, create_graph=True)
            (hess,) = torch.autograd.grad((force ** 2).sum(), x)
            res.append(hess.cpu().numpy())
    finally:
        G.FUSED = True
    assert rel_err(res[0], res[1]) < 1e-4          # both fp32


def test_trainable_bessel_frequencies_get_their_gradient_from_the_kernel():
    """kgcnn/layers/geom.py:766-770: the frequencies are a trainable variable.  A force-training loss reaches them
    through the energy (d rbf / d f) and through the forces (d2 rbf / d d d f); both come from
    kgcnn_bessel_frequency_grad while the basis itself stays on the fused geometry kernel."""
    from gcnn_keras_b200 import RaggedTensor, _lib
    from gcnn_keras_b200.functional import no_coordinate_hessian
    from gcnn_keras_b200.layers import geom as G
    x_np, idx, nl, el = _mol_batch(seed=4, n_graphs=6, n_atoms=9)
    rng = np.random.default_rng(9)
    w_e = rng.normal(size=(len(idx), 20))                    # a fixed "network": energy = sum(rbf * w_e)
    w_f = rng.normal(size=x_np.shape)                        # force loss direction

    def loss_and_grad(rbf_of, x, freq, dt):
        rbf = rbf_of(x)
        energy = (rbf * torch.as_tensor(w_e, dtype=dt, device=x.device)).sum()
        (force,) = torch.autograd.grad(energy, x, create_graph=True)
        loss = 0.3 * energy + (force * torch.as_tensor(w_f, dtype=dt, device=x.device)).sum()
        with no_coordinate_hessian():
            (g,) = torch.autograd.grad(loss, freq)
        return float(loss), g

    lay = G.BesselBasisLayer(num_radial=20, cutoff=5.0, envelope_exponent=5).cuda()
    x_dev = torch.tensor(x_np, device="cuda", requires_grad=True)

    def device_rbf(x):
        xyz = RaggedTensor.from_row_lengths(x, torch.as_tensor(nl).cuda())
        p1, p2 = G.NodePosition()([xyz, device_ragged(idx, el)])
        return lay(G.NodeDistanceEuclidean()([p1, p2])).values

    _lib.launch_count_reset()
    loss_d, g_d = loss_and_grad(device_rbf, x_dev, lay.frequencies, torch.float32)
    assert _lib.launch_count() > 0

    f64 = lay.frequencies.detach().double().cpu().requires_grad_(True)
    x64 = torch.tensor(x_np, dtype=torch.float64, requires_grad=True)

    def oracle_rbf(x):
        xyz = oracle_ragged(np.zeros((len(x_np), 3)), nl, torch.float64).with_values(x)
        p1, p2 = OL.node_position(xyz, oracle_ragged(idx, el))
        return OL.bessel_basis(OL.node_distance_euclidean(p1, p2), 20, 5.0, 5, frequencies=f64).values

    loss_o, g_o = loss_and_grad(oracle_rbf, x64, f64, torch.float64)
    assert abs(loss_d - loss_o) <= RTOL * abs(loss_o)
    assert rel_err(g_d.cpu().numpy(), g_o.numpy()) < RTOL


@pytest.mark.parametrize("kind", ["distance", "direction", "gauss", "bessel", "cos"])
@pytest.mark.parametrize("add_eps", [False, True])
def test_coordinate_hessian_kernel_matches_fp64_oracle(kind, add_eps):
    """kgcnn_edge_geometry mode 3 (the VJP of the coordinate VJP, kgcnn/layers/geom.py:181-190 under two tapes):
    d/dx <J(x)^T c, h> against the fp64 restatement, for every output the fused geometry produces."""
    from gcnn_keras_b200 import functional as F
    assert F.GEOMETRY_HESSIAN_KERNEL
    x_np, idx, nl, el = _mol_batch(seed=4)
    x = torch.tensor(x_np, device="cuda", requires_grad=True)
    ox = torch.tensor(x_np, dtype=torch.float64, requires_grad=True)
    out = _device_outputs(kind, x, nl, idx, el, add_eps)
    o_out = _oracle_outputs(kind, ox, nl, idx, el, add_eps)
    rng = np.random.default_rng(7)
    cot = rng.standard_normal(tuple(o_out.shape))
    h = rng.standard_normal(x_np.shape)
    launches = F._lib.launch_count()
    (vjp,) = torch.autograd.grad(out, x, torch.tensor(cot, dtype=torch.float32, device="cuda"), create_graph=True)
    (hess,) = torch.autograd.grad(vjp, x, torch.tensor(h, dtype=torch.float32, device="cuda"))
    assert F._lib.launch_count() > launches
    (o_vjp,) = torch.autograd.grad(o_out, ox, torch.tensor(cot), create_graph=True)
    (o_hess,) = torch.autograd.grad(o_vjp, ox, torch.tensor(h))
    assert rel_err(hess.cpu().numpy(), o_hess.numpy()) < RTOL


def test_gauss_basis_on_input_distances_matches_fp64_through_two_derivatives():
    """GaussBasisLayer on distances that are a model INPUT (kgcnn/literature/Schnet.py make_distance=False): one launch
    of kgcnn_gauss_basis, its VJP and the VJP of that, against the fp64 expansion."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.layers import geom as G
    rng = np.random.default_rng(3)
    el = np.array([7, 0, 12, 30], dtype=np.int64)
    d_np = rng.uniform(0.5, 4.9, size=(int(el.sum()), 1))
    d = torch.tensor(d_np, dtype=torch.float32, device="cuda", requires_grad=True)
    od = torch.tensor(d_np, dtype=torch.float64, requires_grad=True)
    launches = F._lib.launch_count()
    from gcnn_keras_b200 import RaggedTensor
    out = G.GaussBasisLayer(bins=25, distance=5.0, sigma=0.4, offset=0.1)(
        RaggedTensor.from_row_lengths(d, torch.as_tensor(el).cuda())).values
    assert F._lib.launch_count() == launches + 1 and tuple(out.shape) == (int(el.sum()), 25)
    want = OL.gauss_basis(oracle_ragged(np.zeros_like(d_np), el, torch.float64).with_values(od), bins=25, dist=5.0,
                          sigma=0.4, offset=0.1).values
    cot, h = rng.standard_normal(tuple(want.shape)), rng.standard_normal(d_np.shape)
    (g,) = torch.autograd.grad(out, d, torch.tensor(cot, dtype=torch.float32, device="cuda"), create_graph=True)
    (gg,) = torch.autograd.grad(g, d, torch.tensor(h, dtype=torch.float32, device="cuda"))
    (og,) = torch.autograd.grad(want, od, torch.tensor(cot), create_graph=True)
    (ogg,) = torch.autograd.grad(og, od, torch.tensor(h))
    assert rel_err(out.detach().cpu().numpy(), want.detach().numpy()) < RTOL
    assert rel_err(g.detach().cpu().numpy(), og.detach().numpy()) < RTOL
    assert rel_err(gg.cpu().numpy(), ogg.numpy()) < RTOL
