"""Fused PaiNN message kernels (csrc/painn.cu) vs the op-for-op oracle of painn_conv.py:104-114 in fp64:
forward, first derivatives and the second derivatives force training needs."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, random_graph_batch, rel_err
from oracle import layers as OL

pytestmark = pytest.mark.gpu

NAMES = ("phi", "v", "w", "r")


def _inputs(seed, units, sorted_edges, n_graphs=5, max_nodes=14, max_deg=6):
    rng = np.random.default_rng(seed)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs, max_nodes, max_deg, sorted_edges=sorted_edges)
    n, e = int(node_len.sum()), int(edge_len.sum())
    data = {"phi": rng.standard_normal((n, 3 * units)), "v": rng.standard_normal((n, 3, units)),
            "w": rng.standard_normal((e, 3 * units)), "r": rng.standard_normal((e, 3))}
    probe = {"cs": rng.standard_normal((n, units)), "cv": rng.standard_normal((n, 3, units)),
             "h": {k: rng.standard_normal(data[k].shape) for k in NAMES}}
    return node_len, edge_len, index, data, probe


def _oracle_message(t, node_len, edge_len, index):
    """painn_conv.py:100,104-114 with the oracle's gather / pooling restatements."""
    from oracle import R
    node = R.from_row_lengths(t["phi"], node_len)
    idx = oracle_ragged(index, edge_len)
    s_j = OL.gather_nodes_outgoing(node, idx)
    sw = s_j.values * t["w"]
    sw1, sw2, sw3 = torch.chunk(sw, 3, dim=-1)
    ds = OL.pooling_local_edges(node, s_j.with_values(sw1), idx, pooling_method="sum")
    vj = OL.gather_nodes_outgoing(node.with_values(t["v"]), idx)
    dv = sw2.unsqueeze(-2) * vj.values + sw3.unsqueeze(-2) * t["r"].unsqueeze(-1)
    dv = OL.pooling_local_edges(node, s_j.with_values(dv), idx, pooling_method="sum")
    return ds.values, dv.values


def _fused_message(t, node_len, edge_len, index):
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import get_plan
    node = device_ragged(np.zeros((int(node_len.sum()), 1), np.float32), node_len)
    idx = device_ragged(index, edge_len)
    plan = get_plan(idx, node)
    return F.painn_message(t["phi"], t["v"], t["w"], t["r"], plan.relation(0), plan.relation(1))


def _derivatives(message, t, probe, conv):
    """outputs, first derivatives of <cs,ds>+<cv,dv>, and derivatives of sum_k <h_k, first_k> (second order)."""
    ds, dv = message(t)
    scalar = (ds * conv(probe["cs"])).sum() + (dv * conv(probe["cv"])).sum()
    first = torch.autograd.grad(scalar, [t[k] for k in NAMES], create_graph=True)
    second_scalar = sum((g * conv(probe["h"][k])).sum() for k, g in zip(NAMES, first))
    second = torch.autograd.grad(second_scalar, [t[k] for k in NAMES], allow_unused=True)
    return (ds, dv), first, second


@pytest.mark.parametrize("units,sorted_edges", [(128, True), (32, False), (160, False), (256, True)])
def test_painn_message_all_orders(units, sorted_edges):
    node_len, edge_len, index, data, probe = _inputs(units, units, sorted_edges)
    t64 = {k: torch.tensor(x, dtype=torch.float64, requires_grad=True) for k, x in data.items()}
    t32 = {k: torch.tensor(x, dtype=torch.float32, device="cuda", requires_grad=True) for k, x in data.items()}
    ref = _derivatives(lambda t: _oracle_message(t, node_len, edge_len, index), t64, probe,
                       lambda x: torch.tensor(x, dtype=torch.float64))
    got = _derivatives(lambda t: _fused_message(t, node_len, edge_len, index), t32, probe,
                       lambda x: torch.tensor(x, dtype=torch.float32, device="cuda"))
    for level, (a, b) in enumerate(zip(got, ref)):
        for k, (x, y) in enumerate(zip(a, b)):
            if y is None:
                assert x is None or float(x.abs().max()) == 0.0
                continue
            assert rel_err(x.detach().cpu().numpy(), y.detach().numpy()) < RTOL, (level, k)


def test_painn_message_is_deterministic_and_matches_composed_layer():
    from gcnn_keras_b200.layers.conv import painn_conv as PC
    node_len, edge_len, index, data, _ = _inputs(3, 64, False)
    t32 = {k: torch.tensor(x, dtype=torch.float32, device="cuda") for k, x in data.items()}
    a = _fused_message(t32, node_len, edge_len, index)
    b = _fused_message(t32, node_len, edge_len, index)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert PC.FUSED
