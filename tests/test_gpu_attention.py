"""Segment softmax + attention pooling through the C-ABI against the fp64 oracle: values,
gradients and double gradients, 1e-5 relative."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, random_graph_batch, rel_err
from oracle import layers as OL
from oracle import ops as OO

pytestmark = pytest.mark.gpu


def _case(seed, heads, sorted_edges, feat=16, n_graphs=7):
    rng = np.random.default_rng(seed)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=n_graphs, sorted_edges=sorted_edges)
    x = np.zeros((int(node_len.sum()), 2), dtype=np.float32)
    e = rng.standard_normal((int(edge_len.sum()), feat)).astype(np.float32)
    a = (3.0 * rng.standard_normal((int(edge_len.sum()), heads))).astype(np.float32)
    return node_len, edge_len, index, x, e, a


@pytest.mark.parametrize("sorted_edges", [True, False])
@pytest.mark.parametrize("heads", [1, 16])
def test_pooling_local_edges_attention(sorted_edges, heads):
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdgesAttention
    node_len, edge_len, index, x, e, a = _case(3, heads, sorted_edges)
    oe = oracle_ragged(e, edge_len, torch.float64)
    oa = oracle_ragged(a, edge_len, torch.float64)
    oe.values.requires_grad_(True), oa.values.requires_grad_(True)
    want = OL.pooling_local_edges_attention(oracle_ragged(x, node_len), oe, oa, oracle_ragged(index, edge_len)).values
    proj = torch.as_tensor(np.random.default_rng(0).standard_normal(tuple(want.shape)))
    wge, wga = torch.autograd.grad((want * proj).sum(), [oe.values, oa.values], create_graph=True)
    wgg = torch.autograd.grad((wga ** 2).sum() + (wge ** 2).sum(), [oe.values, oa.values])

    de, da = device_ragged(e, edge_len), device_ragged(a, edge_len)
    de.values.requires_grad_(True), da.values.requires_grad_(True)
    got = PoolingLocalEdgesAttention()([device_ragged(x, node_len), de, da, device_ragged(index, edge_len)]).values
    ge, ga = torch.autograd.grad((got * proj.to(got)).sum(), [de.values, da.values], create_graph=True)
    gg = torch.autograd.grad((ga ** 2).sum() + (ge ** 2).sum(), [de.values, da.values])
    assert rel_err(got.detach().cpu().numpy(), want.detach().numpy()) < RTOL
    assert rel_err(ge.detach().cpu().numpy(), wge.detach().numpy()) < RTOL
    assert rel_err(ga.detach().cpu().numpy(), wga.detach().numpy()) < RTOL
    assert rel_err(gg[0].cpu().numpy(), wgg[0].numpy()) < RTOL
    assert rel_err(gg[1].cpu().numpy(), wgg[1].numpy()) < RTOL


def test_pooling_nodes_attention():
    from gcnn_keras_b200.layers.pooling import PoolingNodesAttention
    rng = np.random.default_rng(11)
    node_len = np.array([5, 0, 9, 1, 40], dtype=np.int64)
    n = rng.standard_normal((int(node_len.sum()), 32)).astype(np.float32)
    a = (2.0 * rng.standard_normal((int(node_len.sum()), 1))).astype(np.float32)
    want = OL.pooling_embedding_attention(oracle_ragged(n, node_len, torch.float64), oracle_ragged(a, node_len, torch.float64))
    got = PoolingNodesAttention()([device_ragged(n, node_len), device_ragged(a, node_len)])
    assert tuple(got.shape) == (5, 32)
    assert rel_err(got.cpu().numpy(), want.numpy()) < RTOL
    assert float(got[1].abs().max()) == 0.0                       # empty graph -> 0


def test_segment_softmax_properties_large():
    """C5-like degree distribution at a size the oracle does not need: coefficients of every non-empty segment
    sum to 1, are shift-invariant, and repeat bit-for-bit run to run; large logits do not overflow."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation
    g = torch.Generator(device="cuda").manual_seed(0)
    n, e = 200_000, 6_000_000
    ids = torch.randint(0, n, (e,), device="cuda", generator=g, dtype=torch.int32)
    a = 30.0 * torch.randn((e, 1), device="cuda", generator=g)
    rel = Relation(ids, n)
    alpha = F.segment_softmax(a, rel)
    sums = F.segment_reduce(alpha, rel, "sum")
    nonempty = (rel.rowptr[1:] > rel.rowptr[:-1]).unsqueeze(1)
    assert bool(torch.isfinite(alpha).all())
    assert float((sums[nonempty] - 1.0).abs().max()) < 1e-5
    assert torch.equal(alpha, F.segment_softmax(a, rel))
    shifted = F.segment_softmax(a + 7.0, rel)
    assert float((shifted - alpha).abs().max()) < 1e-5
    # normalize=False (segment.py:16) agrees where exp does not overflow
    small = 0.1 * a
    assert rel_err(F.segment_softmax(small, rel, normalize=False).cpu().numpy(),
                   F.segment_softmax(small, rel).cpu().numpy()) < RTOL


def test_segment_softmax_oracle_unsorted_heads():
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation
    rng = np.random.default_rng(2)
    ids = rng.integers(0, 50, size=2000)
    a = rng.standard_normal((2000, 4)).astype(np.float32)
    order = np.argsort(ids, kind="stable")
    want = np.empty((2000, 4))
    want[order] = OO.segment_softmax(torch.as_tensor(a[order], dtype=torch.float64), torch.as_tensor(ids[order])).numpy()
    got = F.segment_softmax(torch.as_tensor(a, device="cuda"), Relation(torch.as_tensor(ids, device="cuda", dtype=torch.int32), 50))
    assert rel_err(got.cpu().numpy(), want) < RTOL
