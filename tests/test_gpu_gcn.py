"""GCN layer / model on the GPU vs the fp64 oracle, and the partitioned (config 5) path emulated on one GPU."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, random_graph_batch, rel_err
from oracle import conv as OC

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("normalize", [False, True])
def test_gcn_layer_forward_and_gradients(normalize):
    from gcnn_keras_b200.layers.conv.gcn_conv import GCN
    rng = np.random.default_rng(0)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=7, sorted_edges=False)
    x = rng.standard_normal((int(node_len.sum()), 16)).astype(np.float32)
    a = rng.uniform(0.2, 1.0, (int(edge_len.sum()), 1)).astype(np.float32)
    layer = GCN(units=24, activation="kgcnn>leaky_relu", normalize_by_weights=normalize)
    nodes = device_ragged(x, node_len)
    nodes.values.requires_grad_(True)
    out = layer([nodes, device_ragged(a, edge_len), device_ragged(index, edge_len)])
    params = {"lay_dense.kernel": layer.lay_dense.kernel.detach().cpu().double().requires_grad_(True),
              "lay_dense.bias": layer.lay_dense.bias.detach().cpu().double().requires_grad_(True)}
    ox = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    o_nodes = oracle_ragged(x, node_len).with_values(ox)
    want = OC.gcn(OC.P(params), o_nodes, oracle_ragged(a, edge_len, torch.float64), oracle_ragged(index, edge_len),
                  activation="kgcnn>leaky_relu", normalize_by_weights=normalize)
    assert rel_err(out.values.detach().cpu().numpy(), want.values.detach().numpy()) < RTOL
    g = torch.autograd.grad((out.values ** 2).sum(), [nodes.values, layer.lay_dense.kernel])
    og = torch.autograd.grad((want.values ** 2).sum(), [ox, params["lay_dense.kernel"]])
    for u, v in zip(g, og):
        assert rel_err(u.cpu().numpy(), v.numpy()) < RTOL


def test_gcn_model_runs():
    from gcnn_keras_b200.literature.GCN import make_model
    rng = np.random.default_rng(1)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=5, allow_empty=False)
    z = rng.integers(1, 9, size=int(node_len.sum()))
    a = rng.uniform(0.2, 1.0, (int(edge_len.sum()), 1)).astype(np.float32)
    model = make_model(depth=2, gcn_args={"units": 32, "use_bias": True, "activation": "relu",
                                          "pooling_method": "sum"})
    out = model([device_ragged(z, node_len), device_ragged(a, edge_len), device_ragged(index, edge_len)])
    assert out.shape == (5, 1) and bool(torch.isfinite(out).all())


def test_partitioned_aggregate_matches_single_shard():
    """Config 5 emulated on one GPU: 4 receiver-range shards, each reading the full (all-gathered) feature
    block, concatenated == the unpartitioned fused op; and the transposed op summed over shards == d/dh."""
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.partition import PartitionedGraph
    from gcnn_keras_b200.plan import Relation
    rng = np.random.default_rng(2)
    n, e, feat, world = 1003, 20000, 32, 4
    recv = np.sort(rng.integers(0, n, size=e))
    send = rng.integers(0, n, size=e)
    w = rng.uniform(0.1, 1.0, size=e).astype(np.float32)
    shards = [PartitionedGraph(n, recv, send, w, r, world, "cuda") for r in range(world)]
    per = shards[0].per
    h = torch.randn(per * world, feat, device="cuda")
    h[n:] = 0
    full = F.gather_mul_segment_sum(h, torch.tensor(w).cuda().reshape(-1, 1),
                                    Relation(torch.tensor(recv, dtype=torch.int32).cuda(), per * world),
                                    Relation(torch.tensor(send, dtype=torch.int32).cuda(), per * world))
    parts = [F.gather_mul_segment_sum(h, s.weights, s.rel_recv, s.rel_send) for s in shards]
    assert torch.equal(torch.cat(parts, 0), full)       # same in-segment order -> bitwise identical
    cot = torch.randn_like(full)
    hg = h.clone().requires_grad_(True)
    (want,) = torch.autograd.grad(F.gather_mul_segment_sum(
        hg, torch.tensor(w).cuda().reshape(-1, 1),
        Relation(torch.tensor(recv, dtype=torch.int32).cuda(), per * world),
        Relation(torch.tensor(send, dtype=torch.int32).cuda(), per * world)), hg, cot)
    got = sum(F.raw_gms(cot[r * per:(r + 1) * per].contiguous(), s.weights, s.rel_send, s.rel_recv)
              for r, s in enumerate(shards))           # = what reduce-scatter(sum) adds up across ranks
    assert rel_err(got.cpu().numpy(), want.cpu().numpy()) < RTOL
