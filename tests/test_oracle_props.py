"""Property checks that pin the UNPINNED parts of the oracle (no reference golden vectors exist for them):
fp64 gradcheck / gradgradcheck of the op chains, finite-difference forces, invariances."""
import numpy as np
import torch

from helpers import c1_batch, oracle_ragged, random_graph_batch, schnet_params
from oracle import R
from oracle import conv as OC
from oracle import layers as OL
from oracle import models as OM
from oracle import ops as OO


def test_partition_row_indexing_docstring_example():
    # kgcnn/ops/partition.py:110-121
    out = OO.partition_row_indexing(torch.tensor([0, 0, 1, 1]), torch.tensor([2, 2]), torch.tensor([3, 1]),
                                    "row_lengths", "row_lengths")
    assert out.tolist() == [0, 0, 1, 3]


def test_pooling_empty_segments_are_zero_for_every_method():
    # F4 in SURVEY.md: sorted segment op + scatter_nd padding -> 0, also for max/min
    nodes = oracle_ragged(np.zeros((5, 1)), [3, 2])
    edges = oracle_ragged(np.array([[1.0], [-2.0], [3.0]]), [2, 1])
    index = oracle_ragged(np.array([[0, 1], [0, 2], [0, 1]]), [2, 1])
    for method, want in [("sum", [-1, 0, 0, 3, 0]), ("mean", [-0.5, 0, 0, 3, 0]), ("max", [1, 0, 0, 3, 0]),
                         ("min", [-2, 0, 0, 3, 0])]:
        out = OL.pooling_local_edges(nodes, edges, index, method)
        assert out.values.flatten().tolist() == [float(w) for w in want]


def test_pooling_is_invariant_under_edge_permutation():
    rng = np.random.default_rng(0)
    node_len, edge_len, index = c1_batch(rng, n_graphs=3, n_nodes=7, deg=4)
    e = rng.standard_normal((len(index), 5))
    nodes = oracle_ragged(np.zeros((21, 1)), node_len)
    a = OL.pooling_local_edges(nodes, oracle_ragged(e, edge_len), oracle_ragged(index, edge_len), "sum").values
    perm = np.concatenate([g * 28 + rng.permutation(28) for g in range(3)])
    b = OL.pooling_local_edges(nodes, oracle_ragged(e[perm], edge_len), oracle_ragged(index[perm], edge_len),
                               "sum").values
    assert torch.allclose(a, b, atol=1e-12)


def test_cfconv_gradcheck_and_gradgradcheck():
    rng = np.random.default_rng(1)
    node_len, edge_len, index = c1_batch(rng, n_graphs=2, n_nodes=4, deg=2)
    o_idx = oracle_ragged(index, edge_len)
    x = torch.tensor(rng.standard_normal((8, 3)), requires_grad=True)
    w = torch.tensor(rng.standard_normal((len(index), 3)), requires_grad=True)

    def fn(x, w):
        nodes = R.from_row_lengths(x, node_len)
        msg = OL.gather_nodes_outgoing(nodes, o_idx).values * w
        return OL.pooling_local_edges(nodes, o_idx.with_values(msg), o_idx, "sum").values

    assert torch.autograd.gradcheck(fn, (x, w))
    assert torch.autograd.gradgradcheck(fn, (x, w))


def _schnet_setup(batch=2, depth=1):
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    cfg = dict(SCHNET_C2_CONFIG, depth=depth)
    cfg["interaction_args"] = dict(cfg["interaction_args"], units=16)
    cfg["input_embedding"] = {"node": {"input_dim": 95, "output_dim": 16}}
    cfg["last_mlp"] = {"use_bias": [True, True], "units": [8, 1], "activation": ["shifted_softplus", "linear"]}
    b = schnet_c2_batch(batch, 0)
    params = schnet_params(cfg, np.random.default_rng(2))
    z = oracle_ragged(b["z"], b["node_len"])
    edi = oracle_ragged(b["index"], b["edge_len"])

    def energy_fn(x):
        return OM.schnet_energy(params, z, x, edi, cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                cfg["last_mlp"]["activation"], "sum", None)

    return b, energy_fn


def test_schnet_oracle_forces_match_finite_differences():
    b, energy_fn = _schnet_setup()
    xyz = oracle_ragged(b["xyz"], b["node_len"], torch.float64)
    eng, force = OM.energy_force(energy_fn, xyz, create_graph=False)
    assert eng.shape == (2, 1) and force.shape == (18, 3, 1)
    h = 1e-6
    x0 = xyz.values.clone()
    for a, c in [(0, 0), (5, 2), (13, 1)]:
        xp, xm = x0.clone(), x0.clone()
        xp[a, c] += h
        xm[a, c] -= h
        fd = (energy_fn(xyz.with_values(xp)).sum() - energy_fn(xyz.with_values(xm)).sum()) / (2 * h)
        assert abs(force[a, c, 0].item() - fd.item()) < 1e-6 * max(1.0, abs(fd.item()))


def test_schnet_oracle_energy_is_rotation_and_translation_invariant():
    b, energy_fn = _schnet_setup()
    xyz = oracle_ragged(b["xyz"], b["node_len"], torch.float64)
    q, _ = np.linalg.qr(np.random.default_rng(3).standard_normal((3, 3)))
    rot = xyz.with_values(xyz.values @ torch.tensor(q) + torch.tensor([0.3, -1.0, 2.0]))
    with torch.no_grad():
        assert torch.allclose(energy_fn(xyz), energy_fn(rot), atol=1e-10)


def _wacsf_case(seed=0):
    rng = np.random.default_rng(seed)
    n = 6
    xyz = rng.uniform(0.0, 3.0, size=(n, 3))
    z = np.array([1, 6, 8, 1, 7, 6])
    eij = np.array([[i, j] for i in range(n) for j in range(n) if i != j])
    ijk = np.array([[i, j, k] for i in range(n) for j in range(n) for k in range(n) if len({i, j, k}) == 3][::3])
    return z, xyz, eij, ijk


def test_wacsf_oracle_brute_force_formula_and_gradcheck():
    # wacsf_conv.py docstrings: W_rad_i = sum_j Z_j exp(-eta (r_ij-mu)^2) fc ;  W_ang_i = 2^(1-zeta) sum_jk Z_j Z_k
    # (1 + lambda cos)^zeta exp(-eta sum (r-mu)^2) fc fc fc  -- restated here with Python loops (small case)
    z, xyz, eij, ijk = _wacsf_case()
    rng = np.random.default_rng(1)
    rad = np.zeros((118, 3, 2))
    rad[..., 0], rad[..., 1] = rng.uniform(0.5, 2, (118, 3)), rng.uniform(0.5, 3, (118, 3))
    ang = np.zeros((118, 2, 4))
    ang[..., 0], ang[..., 1] = rng.uniform(0.05, 0.5, (118, 2)), rng.uniform(0, 1, (118, 2))
    ang[..., 2], ang[..., 3] = rng.choice([-1.0, 1.0], (118, 2)), rng.choice([1.0, 2.0, 3.0], (118, 2))
    zr = oracle_ragged(z, [len(z)])
    xr = oracle_ragged(xyz, [len(z)], torch.float64)
    fc = lambda r: 0.5 * (np.cos(np.pi * min(r, 8.0) / 8.0) + 1)
    want_r = np.zeros((len(z), 3))
    for i, j in eij:
        r = np.linalg.norm(xyz[i] - xyz[j])
        want_r[i] += z[j] * np.exp(-rad[z[i], :, 0] * (r - rad[z[i], :, 1]) ** 2) * fc(r)
    got_r = OC.wacsf_rad(zr, xr, oracle_ragged(eij, [len(eij)]), rad).values.numpy()
    assert np.allclose(got_r, want_r, rtol=1e-12, atol=1e-14)
    want_a = np.zeros((len(z), 2))
    for i, j, k in ijk:
        vij, vik = xyz[i] - xyz[j], xyz[i] - xyz[k]
        rij, rik, rjk = np.linalg.norm(vij), np.linalg.norm(vik), np.linalg.norm(xyz[j] - xyz[k])
        eta, mu, lam, zeta = (ang[z[i], :, c] for c in range(4))
        cos = vij @ vik / rij / rik
        want_a[i] += (z[j] * z[k] * (1 + lam * cos) ** zeta * fc(rij) * fc(rik) * fc(rjk)
                      * np.exp(-eta * ((rij - mu) ** 2 + (rik - mu) ** 2 + (rjk - mu) ** 2)))
    want_a *= 2.0 ** (1 - ang[z, :, 3])
    got_a = OC.wacsf_ang(zr, xr, oracle_ragged(ijk, [len(ijk)]), ang).values.numpy()
    assert np.allclose(got_a, want_a, rtol=1e-12, atol=1e-14)

    x = torch.tensor(xyz, dtype=torch.float64, requires_grad=True)
    f_r = lambda t: OC.wacsf_rad(zr, xr.with_values(t), oracle_ragged(eij, [len(eij)]), rad).values
    f_a = lambda t: OC.wacsf_ang(zr, xr.with_values(t), oracle_ragged(ijk, [len(ijk)]), ang).values
    assert torch.autograd.gradcheck(f_r, (x,))
    assert torch.autograd.gradcheck(f_a, (x,))


def test_attention_pooling_oracle_properties():
    """segment_softmax (segment.py:4-24): coefficients sum to 1 per non-empty receiver, are invariant under a
    per-segment shift, and attention pooling of constant edges returns that constant; gradcheck in fp64."""
    rng = np.random.default_rng(4)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=5, sorted_edges=False)
    nodes = oracle_ragged(np.zeros((int(node_len.sum()), 1)), node_len)
    E = int(edge_len.sum())
    att = oracle_ragged(rng.standard_normal((E, 1)), edge_len, torch.float64)
    ones = oracle_ragged(np.ones((E, 3)), edge_len, torch.float64)
    o_idx = oracle_ragged(index, edge_len)
    out = OL.pooling_local_edges_attention(nodes, ones, att, o_idx).values
    assert bool(((out - 1.0).abs().max(dim=1).values < 1e-12).logical_or(out.abs().max(dim=1).values == 0).all())
    shifted = OL.pooling_local_edges_attention(nodes, ones, att.with_values(att.values + 5.0), o_idx).values
    assert float((shifted - out).abs().max()) < 1e-12
    e = torch.as_tensor(rng.standard_normal((E, 3)), dtype=torch.float64).requires_grad_(True)
    a = att.values.clone().requires_grad_(True)
    fn = lambda e_, a_: OL.pooling_local_edges_attention(nodes, ones.with_values(e_), att.with_values(a_), o_idx).values
    assert torch.autograd.gradcheck(fn, (e, a))


def test_oracle_lstm_pooling_agrees_with_an_independent_lstm():
    """The LSTM restatement (oracle/layers.py, Keras gate order i, f, c, o) against torch.nn.LSTM (same gate order,
    weights transposed) on every node's edge sequence."""
    import numpy as np
    import torch
    from oracle import R
    from oracle import layers as OL
    rng = np.random.default_rng(0)
    n, feat, u = 5, 6, 4
    recv = np.array([0, 0, 0, 2, 2, 3, 4, 4, 4, 4])
    idx = np.stack([recv, rng.integers(0, n, size=recv.shape)], 1)
    edges = torch.tensor(rng.standard_normal((len(recv), feat)))
    w, uu, b = (torch.tensor(rng.standard_normal(s)) for s in ((feat, 4 * u), (u, 4 * u), (4 * u,)))
    nodes = R.from_row_lengths(torch.zeros((n, 1), dtype=torch.float64), [n])
    got = OL.pooling_local_edges_lstm(nodes, R.from_row_lengths(edges, [len(recv)]),
                                      R.from_row_lengths(torch.tensor(idx), [len(recv)]), w, uu, b).values
    lstm = torch.nn.LSTM(feat, u, batch_first=True).double()
    with torch.no_grad():
        lstm.weight_ih_l0.copy_(w.t()), lstm.weight_hh_l0.copy_(uu.t()), lstm.bias_ih_l0.copy_(b), lstm.bias_hh_l0.zero_()
    for i in range(n):
        seq = edges[torch.tensor(recv == i)]
        want = torch.zeros(u, dtype=torch.float64) if len(seq) == 0 else lstm(seq[None])[1][0][0, 0]
        assert torch.allclose(got[i], want.detach(), atol=1e-12)
