"""GPU parity tests (through the C-ABI) of plan / gather / segment primitives against the oracle."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from helpers import RTOL, c1_batch, device_ragged, oracle_ragged, random_graph_batch, rel_err
from oracle import layers as OL
from oracle import ops as OO

pytestmark = pytest.mark.gpu


def _plan_for(node_len, edge_len, index):
    from gcnn_keras_b200.plan import get_plan
    nodes = device_ragged(np.zeros((int(node_len.sum()), 1), dtype=np.float32), node_len)
    idx = device_ragged(index, edge_len)
    return get_plan(idx, nodes), nodes, idx


@pytest.mark.parametrize("sorted_edges", [True, False])
def test_shift_index_bit_exact(sorted_edges):
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(0)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=9, sorted_edges=sorted_edges)
    nodes = oracle_ragged(np.zeros((int(node_len.sum()), 1)), node_len)
    idx = oracle_ragged(index, edge_len)
    want = OO.partition_row_indexing(idx.values, nodes.row_splits, idx.row_lengths(), "row_splits", "row_length")
    out32, out64 = F.raw_shift_index(torch.as_tensor(index).cuda(), nodes.row_splits.cuda(), idx.row_splits.cuda(),
                                     want_i64=True)
    assert torch.equal(out64.cpu(), want)                         # bit-exact, same layout
    assert torch.equal(out32.cpu().to(torch.int64).t(), want)     # int32 copy is column-major [K, E]


def test_shift_index_triples_and_value_rowids():
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(1)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=5, k=3)
    nodes = oracle_ragged(np.zeros((int(node_len.sum()), 1)), node_len)
    idx = oracle_ragged(index, edge_len)
    want = OO.partition_row_indexing(idx.values, nodes.row_splits, idx.row_lengths(), "row_splits", "row_length")
    out32 = F.raw_shift_index(torch.as_tensor(index).cuda(), nodes.row_splits.cuda(), idx.row_splits.cuda())
    assert torch.equal(out32.cpu().to(torch.int64).t(), want)
    ids = F.raw_value_rowids(nodes.row_splits.cuda(), int(node_len.sum()))
    assert torch.equal(ids.cpu().to(torch.int64), nodes.value_rowids())


@pytest.mark.parametrize("sorted_edges", [True, False])
def test_plan_csr_and_stable_perm(sorted_edges):
    rng = np.random.default_rng(2)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=8, max_nodes=15, max_deg=6,
                                                   sorted_edges=sorted_edges)
    plan, nodes, idx = _plan_for(node_len, edge_len, index)
    for col in (0, 1):
        rel = plan.relation(col)
        key = plan.idx32[col].cpu().numpy()
        counts = np.bincount(key, minlength=plan.N)
        want_rowptr = np.concatenate([[0], np.cumsum(counts)])
        assert np.array_equal(rel.rowptr.cpu().numpy(), want_rowptr)
        want_perm = np.argsort(key, kind="stable")                # tf.argsort(stable=True), pooling.py:66
        if rel.perm is None:
            assert np.array_equal(want_perm, np.arange(len(key)))
        else:
            assert np.array_equal(rel.perm.cpu().numpy(), want_perm)


def test_plan_out_of_range_index_raises():
    node_len = np.array([3, 2], dtype=np.int64)
    edge_len = np.array([2, 1], dtype=np.int64)
    index = np.array([[0, 1], [2, 0], [5, 0]], dtype=np.int64)     # 5 is outside graph 1 AND the batch
    plan, _, _ = _plan_for(node_len, edge_len, index)
    with pytest.raises(IndexError):
        plan.relation(0).rowptr


def test_gather_nodes_reference_fixture():
    # the reference's own gather test (test/test_gather.py:28-44) through the CUDA path
    from gcnn_keras_b200.layers.gather import GatherNodes
    d = np.load(os.path.join(GOLDEN, "gather.npz"))
    nodes = device_ragged(d["nodes"], d["node_lengths"])
    index = device_ragged(d["index"], d["index_lengths"])
    out = GatherNodes()([nodes, index])
    assert np.sum(np.abs(out[1].cpu().numpy() - d["expected_graph1_concat"])) < 1e-6
    out = GatherNodes(concat_axis=None)([nodes, index])
    assert np.sum(np.abs(out[1].cpu().numpy() - d["expected_graph1"])) < 1e-6


@pytest.mark.parametrize("feat", [1, 3, 35, 64, 128, 384, 1000])
def test_gather_selection_bitwise(feat):
    from gcnn_keras_b200.layers.gather import GatherNodesIngoing, GatherNodesOutgoing, GatherNodesSelection
    rng = np.random.default_rng(3)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=7, sorted_edges=False)
    x = rng.standard_normal((int(node_len.sum()), feat)).astype(np.float32)
    o_nodes, o_idx = oracle_ragged(x, node_len), oracle_ragged(index, edge_len)
    nodes, idx = device_ragged(x, node_len), device_ragged(index, edge_len)
    want = OL.gather_embedding_selection(o_nodes, o_idx, [0, 1])
    got = GatherNodesSelection([0, 1])([nodes, idx])
    for w, g in zip(want, got):
        assert torch.equal(g.values.cpu(), w.values)              # a copy: bit-exact
    assert torch.equal(GatherNodesIngoing()([nodes, idx]).values.cpu(), want[0].values)
    assert torch.equal(GatherNodesOutgoing()([nodes, idx]).values.cpu(), want[1].values)


def test_gather_equivariant_and_int64():
    from gcnn_keras_b200.layers.gather import GatherNodesOutgoing
    rng = np.random.default_rng(4)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=4)
    v = rng.standard_normal((int(node_len.sum()), 3, 16)).astype(np.float32)
    z = rng.integers(1, 9, size=(int(node_len.sum()),)).astype(np.int64)
    idx = device_ragged(index, edge_len)
    o_idx = oracle_ragged(index, edge_len)
    got = GatherNodesOutgoing()([device_ragged(v, node_len), idx])
    assert torch.equal(got.values.cpu(), OL.gather_nodes_outgoing(oracle_ragged(v, node_len), o_idx).values)
    gotz = GatherNodesOutgoing()([device_ragged(z, node_len), idx])
    assert torch.equal(gotz.values.cpu(), OL.gather_nodes_outgoing(oracle_ragged(z, node_len), o_idx).values)


@pytest.mark.parametrize("method", ["sum", "mean", "max", "min", "segment_sum", "reduce_mean"])
@pytest.mark.parametrize("sorted_edges", [True, False])
@pytest.mark.parametrize("feat", [1, 3, 50, 64, 128, 384])
def test_pooling_local_edges(method, sorted_edges, feat):
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdges
    rng = np.random.default_rng(5)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=8, sorted_edges=sorted_edges)
    x = rng.standard_normal((int(node_len.sum()), 2)).astype(np.float32)
    e = rng.standard_normal((int(edge_len.sum()), feat)).astype(np.float32)
    want = OL.pooling_local_edges(oracle_ragged(x, node_len), oracle_ragged(e, edge_len, torch.float64),
                                  oracle_ragged(index, edge_len), pooling_method=method)
    got = PoolingLocalEdges(pooling_method=method)([device_ragged(x, node_len), device_ragged(e, edge_len),
                                                    device_ragged(index, edge_len)])
    assert got.values.shape == want.values.shape
    assert rel_err(got.values.cpu().numpy(), want.values.numpy()) < RTOL
    if method in ("max", "min"):
        assert torch.equal(got.values.cpu().double(), want.values)   # selection: exact


def test_pooling_unknown_method_raises_type_error():
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdges
    rng = np.random.default_rng(6)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=3)
    x = np.zeros((int(node_len.sum()), 2), dtype=np.float32)
    e = np.zeros((int(edge_len.sum()), 2), dtype=np.float32)
    with pytest.raises(TypeError):
        PoolingLocalEdges(pooling_method="softmax")([device_ragged(x, node_len), device_ragged(e, edge_len),
                                                     device_ragged(index, edge_len)])


@pytest.mark.parametrize("method", ["sum", "mean", "max"])
def test_pooling_nodes(method):
    from gcnn_keras_b200.layers.pooling import PoolingNodes
    rng = np.random.default_rng(7)
    node_len = np.array([4, 0, 7, 1, 3], dtype=np.int64)
    for feat in (1, 64):
        x = rng.standard_normal((int(node_len.sum()), feat)).astype(np.float32)
        want = OL.pooling_embedding(oracle_ragged(x, node_len, torch.float64), method)
        got = PoolingNodes(pooling_method=method)(device_ragged(x, node_len))
        assert got.shape == (5, feat)
        assert rel_err(got.cpu().numpy(), want.numpy()) < RTOL


@pytest.mark.parametrize("normalize", [False, True])
@pytest.mark.parametrize("wcols", [1, 16])
def test_pooling_weighted_local_edges(normalize, wcols):
    from gcnn_keras_b200.layers.pooling import PoolingWeightedLocalEdges
    rng = np.random.default_rng(8)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=6, sorted_edges=False)
    x = np.zeros((int(node_len.sum()), 2), dtype=np.float32)
    e = rng.standard_normal((int(edge_len.sum()), 16)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, (int(edge_len.sum()), wcols)).astype(np.float32)
    want = OL.pooling_weighted_local_edges(oracle_ragged(x, node_len), oracle_ragged(e, edge_len, torch.float64),
                                           oracle_ragged(index, edge_len), oracle_ragged(w, edge_len, torch.float64),
                                           pooling_method="sum", normalize_by_weights=normalize)
    got = PoolingWeightedLocalEdges(pooling_method="sum", normalize_by_weights=normalize)(
        [device_ragged(x, node_len), device_ragged(e, edge_len), device_ragged(index, edge_len),
         device_ragged(w, edge_len)])
    assert rel_err(got.values.cpu().numpy(), want.values.numpy()) < RTOL


def test_relational_pooling():
    from gcnn_keras_b200.layers.pooling import RelationalPoolingLocalEdges
    rng = np.random.default_rng(9)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=6, sorted_edges=False)
    x = np.zeros((int(node_len.sum()), 2), dtype=np.float32)
    e = rng.standard_normal((int(edge_len.sum()), 35)).astype(np.float32)
    r = rng.integers(0, 4, size=(int(edge_len.sum()),)).astype(np.int64)
    for method in ("sum", "max", "min"):
        want = OL.relational_pooling_local_edges(oracle_ragged(x, node_len), oracle_ragged(e, edge_len, torch.float64),
                                                 oracle_ragged(index, edge_len), oracle_ragged(r, edge_len), 4, method)
        got = RelationalPoolingLocalEdges(4, pooling_method=method)(
            [device_ragged(x, node_len), device_ragged(e, edge_len), device_ragged(index, edge_len),
             device_ragged(r, edge_len)])
        assert got.values.shape == (int(node_len.sum()), 4, 35)
        assert rel_err(got.values.cpu().numpy(), want.values.numpy()) < RTOL


def _fused_case(rng, feat, wcols, sorted_edges):
    node_len, edge_len, index = c1_batch(rng, n_graphs=4, n_nodes=9, deg=6, shuffled=not sorted_edges)
    x = rng.standard_normal((int(node_len.sum()), feat)).astype(np.float32)
    w = rng.standard_normal((int(edge_len.sum()), wcols)).astype(np.float32) if wcols else None
    return node_len, edge_len, index, x, w


@pytest.mark.parametrize("sorted_edges", [True, False])
@pytest.mark.parametrize("feat,wcols", [(64, 0), (64, 64), (128, 128), (128, 1), (3, 3), (384, 384), (20, 1)])
def test_fused_gather_mul_segment_sum_forward_backward_double_backward(sorted_edges, feat, wcols):
    """y_i = sum_{e->i} x[s_e] * w_e: value, dx, dw and the double backward against fp64 autograd of the
    reference op chain (gather -> multiply -> PoolingLocalEdges(sum))."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import get_plan
    rng = np.random.default_rng(10)
    node_len, edge_len, index, x, w = _fused_case(rng, feat, wcols, sorted_edges)
    # --- oracle (fp64) ---
    ox = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    ow = torch.tensor(w, dtype=torch.float64, requires_grad=True) if w is not None else None
    o_idx = oracle_ragged(index, edge_len)
    o_nodes = oracle_ragged(np.zeros((len(x), 1)), node_len).with_values(ox)
    msg = OL.gather_nodes_outgoing(o_nodes, o_idx).values
    if ow is not None:
        msg = msg * ow
    oy = OL.pooling_local_edges(o_nodes, o_idx.with_values(msg), o_idx, "sum").values
    cot = torch.tensor(rng.standard_normal(oy.shape))
    o_in = [ox] + ([ow] if ow is not None else [])
    o_g = torch.autograd.grad(oy, o_in, cot, create_graph=True)
    cot2 = [torch.tensor(rng.standard_normal(g.shape)) for g in o_g]
    o_gg = torch.autograd.grad(o_g, o_in + [], cot2, allow_unused=True) if ow is not None else None
    # --- device (fp32) ---
    nodes = device_ragged(x, node_len)
    idx = device_ragged(index, edge_len)
    plan = get_plan(idx, nodes)
    dx = nodes.values.clone().requires_grad_(True)
    dw = torch.tensor(w).cuda().requires_grad_(True) if w is not None else None
    y = F.gather_mul_segment_sum(dx, dw, plan.relation(0), plan.relation(1))
    assert rel_err(y.detach().cpu().numpy(), oy.detach().numpy()) < RTOL
    d_in = [dx] + ([dw] if dw is not None else [])
    g = torch.autograd.grad(y, d_in, cot.float().cuda(), create_graph=True)
    for a, b in zip(g, o_g):
        assert rel_err(a.detach().cpu().numpy(), b.detach().numpy()) < RTOL
    if ow is not None:
        gg = torch.autograd.grad(g, d_in, [c.float().cuda() for c in cot2], allow_unused=True)
        for a, b in zip(gg, o_gg):
            assert rel_err(a.cpu().numpy(), b.numpy()) < RTOL
    # --- determinism: bitwise equal run to run ---
    y2 = F.gather_mul_segment_sum(dx, dw, plan.relation(0), plan.relation(1))
    assert torch.equal(y, y2)


def test_c1_chain_gather_dense_pool():
    """BASELINE.json configs[0]: GatherNodes -> Dense -> PoolingLocalMessages(sum), 32 graphs x 20 nodes x 400
    edges, F=64, forward + gradients vs the fp64 oracle, receiver-sorted and shuffled."""
    from gcnn_keras_b200.layers.gather import GatherNodes
    from gcnn_keras_b200.layers.pooling import PoolingLocalMessages
    for shuffled in (False, True):
        rng = np.random.default_rng(11)
        node_len, edge_len, index = c1_batch(rng, shuffled=shuffled)
        x = rng.standard_normal((640, 64)).astype(np.float32)
        kernel = (rng.standard_normal((128, 64)) / np.sqrt(128)).astype(np.float32)
        bias = rng.standard_normal((64,)).astype(np.float32)
        ox = torch.tensor(x, dtype=torch.float64, requires_grad=True)
        ok = torch.tensor(kernel, dtype=torch.float64, requires_grad=True)
        ob = torch.tensor(bias, dtype=torch.float64)
        o_nodes = oracle_ragged(x, node_len).with_values(ox)
        o_idx = oracle_ragged(index, edge_len)
        msg = OL.dense(OL.gather_embedding(o_nodes, o_idx), ok, ob, "relu")
        oy = OL.pooling_local_edges(o_nodes, msg, o_idx, "sum").values
        ogx, ogk = torch.autograd.grad(oy.sum(), [ox, ok])

        nodes = device_ragged(x, node_len)
        dx = nodes.values.requires_grad_(True)
        dk = torch.tensor(kernel).cuda().requires_grad_(True)
        idx = device_ragged(index, edge_len)
        g = GatherNodes()([nodes, idx])
        m = g.with_values(torch.relu(g.values @ dk + torch.tensor(bias).cuda()))
        y = PoolingLocalMessages(pooling_method="sum")([nodes, m, idx]).values
        gx, gk = torch.autograd.grad(y.sum(), [dx, dk])
        assert rel_err(y.detach().cpu().numpy(), oy.detach().numpy()) < RTOL
        assert rel_err(gx.cpu().numpy(), ogx.numpy()) < RTOL
        assert rel_err(gk.cpu().numpy(), ogk.numpy()) < RTOL


def test_segment_max_gradient_splits_ties():
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation
    v = torch.tensor([[1.0], [3.0], [3.0], [2.0], [5.0]], device="cuda", requires_grad=True)
    ids = torch.tensor([0, 0, 0, 2, 2], dtype=torch.int32, device="cuda")
    y = F.segment_reduce(v, Relation(ids, 4), "max")
    assert y.detach().cpu().flatten().tolist() == [3.0, 0.0, 5.0, 0.0]
    y.sum().backward()
    assert v.grad.cpu().flatten().tolist() == [0.0, 0.5, 0.5, 0.0, 1.0]
    ov = torch.tensor([[1.0], [3.0], [3.0], [2.0], [5.0]], dtype=torch.float64, requires_grad=True)
    OO.segment_ops_by_name("max", ov, torch.tensor([0, 0, 0, 2, 2])).sum().backward()
    assert ov.grad.flatten().tolist() == [0.0, 0.5, 0.5, 0.0, 1.0]


def test_empty_inputs():
    from gcnn_keras_b200.layers.gather import GatherNodes
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdges
    node_len = np.array([3, 2], dtype=np.int64)
    edge_len = np.array([0, 0], dtype=np.int64)
    x = np.ones((5, 8), dtype=np.float32)
    nodes, idx = device_ragged(x, node_len), device_ragged(np.zeros((0, 2), dtype=np.int64), edge_len)
    g = GatherNodes()([nodes, idx])
    assert tuple(g.values.shape) == (0, 16)
    e = device_ragged(np.zeros((0, 8), dtype=np.float32), edge_len)
    p = PoolingLocalEdges(pooling_method="sum")([nodes, e, idx])
    assert tuple(p.values.shape) == (5, 8) and float(p.values.abs().sum()) == 0.0


@pytest.mark.parametrize("feat", [1, 3, 64, 128, 200])
def test_long_segments_warp_per_row_path(feat):
    """Average segment length > 64 selects the warp-per-row mapping (hub nodes / per-graph pooling)."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.layers.pooling import PoolingNodes
    from gcnn_keras_b200.plan import Relation
    rng = np.random.default_rng(12)
    node_len = np.array([300, 0, 150, 97], dtype=np.int64)
    x = rng.standard_normal((int(node_len.sum()), feat)).astype(np.float32)
    for method in ("sum", "mean", "max", "min"):
        want = OL.pooling_embedding(oracle_ragged(x, node_len, torch.float64), method)
        got = PoolingNodes(pooling_method=method)(device_ragged(x, node_len))
        assert rel_err(got.cpu().numpy(), want.numpy()) < RTOL
    # hub graph: 3 receivers, 400 unsorted edges, gather from 50 sources with a per-edge weight row
    recv = rng.integers(0, 3, size=400).astype(np.int32)
    send = rng.integers(0, 50, size=400).astype(np.int32)
    src = rng.standard_normal((50, feat)).astype(np.float32)
    w = rng.standard_normal((400, feat)).astype(np.float32)
    r_out, r_src = Relation(torch.tensor(recv).cuda(), 3), Relation(torch.tensor(send).cuda(), 50)
    got = F.gather_mul_segment_sum(torch.tensor(src).cuda(), torch.tensor(w).cuda(), r_out, r_src)
    want = np.zeros((3, feat))
    np.add.at(want, recv, src[send].astype(np.float64) * w.astype(np.float64))
    assert rel_err(got.cpu().numpy(), want) < RTOL


@pytest.mark.parametrize("sorted_edges", [True, False])
@pytest.mark.parametrize("feat,wcols", [(64, 0), (64, 64), (128, 128), (128, 1), (3, 3), (35, 35), (384, 384), (1, 1)])
def test_fused_many_rows_group_per_row_mapping(sorted_edges, feat, wcols):
    """n_out > 1024 selects the lane-group-per-row kernel (short segments): value + dx + dw vs fp64 NumPy."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import get_plan
    rng = np.random.default_rng(13)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=220, max_nodes=14, max_deg=6,
                                                   sorted_edges=sorted_edges)
    N, E = int(node_len.sum()), int(edge_len.sum())
    assert N > 1024
    x = rng.standard_normal((N, feat)).astype(np.float32)
    w = rng.standard_normal((E, wcols)).astype(np.float32) if wcols else None
    nodes, idx = device_ragged(x, node_len), device_ragged(index, edge_len)
    plan = get_plan(idx, nodes)
    b_idx = plan.idx32.cpu().numpy().T.astype(np.int64)
    msg = x[b_idx[:, 1]].astype(np.float64) * (w.astype(np.float64) if w is not None else 1.0)
    want = np.zeros((N, feat))
    np.add.at(want, b_idx[:, 0], msg)
    dx = nodes.values.clone().requires_grad_(True)
    dw = torch.tensor(w).cuda().requires_grad_(True) if w is not None else None
    y = F.gather_mul_segment_sum(dx, dw, plan.relation(0), plan.relation(1))
    assert rel_err(y.detach().cpu().numpy(), want) < RTOL
    cot = rng.standard_normal((N, feat))
    grads = torch.autograd.grad(y, [dx] + ([dw] if dw is not None else []), torch.tensor(cot).float().cuda())
    want_dx = np.zeros((N, feat))
    np.add.at(want_dx, b_idx[:, 1], cot[b_idx[:, 0]] * (w.astype(np.float64) if w is not None else 1.0))
    assert rel_err(grads[0].cpu().numpy(), want_dx) < RTOL
    if w is not None:
        want_dw = cot[b_idx[:, 0]] * x[b_idx[:, 1]].astype(np.float64)
        if wcols == 1 and feat > 1:
            want_dw = want_dw.sum(1, keepdims=True)
        assert rel_err(grads[1].cpu().numpy(), want_dw) < RTOL
    for method in ("mean", "max", "min"):
        e = rng.standard_normal((E, feat)).astype(np.float32)
        got = F.segment_reduce(torch.tensor(e).cuda(), plan.relation(0), method)
        want_m = OL.pooling_local_edges(oracle_ragged(x, node_len), oracle_ragged(e, edge_len, torch.float64),
                                        oracle_ragged(index, edge_len), method).values
        assert rel_err(got.cpu().numpy(), want_m.numpy()) < RTOL


@pytest.mark.parametrize("feat", [1, 64, 128])
def test_many_long_rows_warp_per_row_mapping(feat):
    """n_out > 1024 with average segment length > 64 selects the warp-per-row kernel."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation
    rng = np.random.default_rng(14)
    n_out, E = 1100, 1100 * 80
    ids = np.sort(rng.integers(0, n_out, size=E)).astype(np.int32)
    v = rng.standard_normal((E, feat)).astype(np.float32)
    rel = Relation(torch.tensor(ids).cuda(), n_out)
    for method in ("sum", "mean", "max"):
        got = F.segment_reduce(torch.tensor(v).cuda(), rel, method)
        want = OO.segment_ops_by_name(method, torch.tensor(v, dtype=torch.float64), torch.tensor(ids, dtype=torch.int64))
        want = torch.cat([want, torch.zeros(n_out - want.shape[0], feat, dtype=torch.float64)])
        assert rel_err(got.cpu().numpy(), want.numpy()) < RTOL


@pytest.mark.parametrize("M,K1,K2", [(18432, 128, 128), (5000, 25, 128), (700, 3, 1), (4097, 35, 50), (100, 64, 64),
                                       (40000, 384, 20)])
def test_weight_gradient_tn_and_colsum(M, K1, K2):
    """(dW, db) = (x^T g, colsum g): tcgen05 TN contraction + deterministic column sum."""
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(15)
    x = rng.standard_normal((M, K1)).astype(np.float32)
    g = rng.standard_normal((M, K2)).astype(np.float32)
    dw = F.raw_dense_tc(2, torch.tensor(x).cuda(), torch.tensor(g).cuda())
    assert rel_err(dw.cpu().numpy(), x.astype(np.float64).T @ g.astype(np.float64)) < RTOL
    dw2 = F.raw_dense_tc(2, torch.tensor(x).cuda(), torch.tensor(g).cuda())
    assert torch.equal(dw, dw2)                      # deterministic split-M reduction
    cs = F.raw_colsum(torch.tensor(g).cuda())
    assert rel_err(cs.cpu().numpy(), g.astype(np.float64).sum(0)) < RTOL


@pytest.mark.parametrize("act", ["kgcnn>shifted_softplus", "swish", "tanh", "linear", "relu", "sigmoid",
                                 "kgcnn>softplus2"])
def test_dense_forward_backward_double_backward(act):
    """Dense (library GEMM + our bias/activation family + our weight-gradient GEMM) to second order."""
    from gcnn_keras_b200 import functional as F
    from oracle import ops as OOps
    rng = np.random.default_rng(16)
    M, K, U = 1500, 25, 32
    x = rng.standard_normal((M, K)).astype(np.float32)
    w = (rng.standard_normal((K, U)) / 5).astype(np.float32)
    b = rng.standard_normal((U,)).astype(np.float32)

    def run(dtype, dev, fn):
        tx = torch.tensor(x, dtype=dtype, device=dev, requires_grad=True)
        tw = torch.tensor(w, dtype=dtype, device=dev, requires_grad=True)
        tb = torch.tensor(b, dtype=dtype, device=dev, requires_grad=True)
        y = fn(tx, tw, tb)
        gx, gw, gb = torch.autograd.grad((y ** 2).sum(), [tx, tw, tb], create_graph=True)
        ggw, ggb = torch.autograd.grad((gx ** 2).sum(), [tw, tb], allow_unused=True)
        return [t.detach().cpu().numpy() if t is not None else None for t in (y, gx, gw, gb, ggw, ggb)]

    got = run(torch.float32, "cuda", lambda a, k, c: F.dense(a, k, c, act))
    want = run(torch.float64, "cpu", lambda a, k, c: OOps.activation(act)(a @ k + c))
    for a, o in zip(got, want):
        if o is None:
            assert a is None or float(np.abs(a).max()) == 0.0
            continue
        if a is None:
            assert float(np.abs(o).max()) < 1e-12
            continue
        assert rel_err(a, o) < RTOL


def test_hub_rows_are_reduced_by_the_long_row_pass():
    """A few receivers with tens of thousands of edges among many ordinary rows (power-law graphs)."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation
    rng = np.random.default_rng(17)
    n, feat = 3000, 64
    recv = np.concatenate([rng.integers(0, n, size=20000), np.full(30000, 7), np.full(5000, 2999),
                           np.full(1500, 1234)]).astype(np.int32)
    send = rng.integers(0, n, size=len(recv)).astype(np.int32)
    perm = rng.permutation(len(recv))
    recv, send = recv[perm], send[perm]
    x = rng.standard_normal((n, feat)).astype(np.float32)
    w = rng.standard_normal((len(recv), 1)).astype(np.float32)
    r_out, r_src = Relation(torch.tensor(recv).cuda(), n), Relation(torch.tensor(send).cuda(), n)
    assert r_out.rowptr is not None and r_out.long_cap > 0 and int(r_out.n_long.item()) == 3
    got = F.gather_mul_segment_sum(torch.tensor(x).cuda(), torch.tensor(w).cuda(), r_out, r_src)
    want = np.zeros((n, feat))
    np.add.at(want, recv, x[send].astype(np.float64) * w.astype(np.float64))
    assert rel_err(got.cpu().numpy(), want) < RTOL
    got2 = F.gather_mul_segment_sum(torch.tensor(x).cuda(), torch.tensor(w).cuda(), r_out, r_src)
    assert torch.equal(got, got2)
    for method in ("mean", "max"):
        v = rng.standard_normal((len(recv), feat)).astype(np.float32)
        g = F.segment_reduce(torch.tensor(v).cuda(), r_out, method)
        order = np.argsort(recv, kind="stable")
        wm = OO.segment_ops_by_name(method, torch.tensor(v[order], dtype=torch.float64),
                                    torch.tensor(recv[order], dtype=torch.int64))
        assert rel_err(g.cpu().numpy()[:wm.shape[0]], wm.numpy()) < RTOL
