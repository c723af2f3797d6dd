"""PoolingLocalEdgesLSTM (kgcnn/layers/pooling.py:288-460) against the fp64 restatement of the
reference call + Keras' LSTMCell: last hidden state per node, gradients of the edges and of the three weights."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, random_graph_batch, rel_err
from oracle import layers as OL

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sorted_edges", [True, False])
@pytest.mark.parametrize("go_backwards", [False, True])
def test_pooling_local_edges_lstm_matches_fp64_oracle(sorted_edges, go_backwards):
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdgesLSTM, PoolingLocalMessagesLSTM
    assert PoolingLocalMessagesLSTM is PoolingLocalEdgesLSTM
    rng = np.random.default_rng(5)
    node_len, edge_len, index = random_graph_batch(rng, n_graphs=6, sorted_edges=sorted_edges)
    feat, units = 16, 32
    x = np.zeros((int(node_len.sum()), 2), dtype=np.float32)
    e = rng.standard_normal((int(edge_len.sum()), feat)).astype(np.float32)
    torch.manual_seed(0)
    lay = PoolingLocalEdgesLSTM(units=units, go_backwards=go_backwards)
    de = device_ragged(e, edge_len)
    de.values.requires_grad_(True)
    got = lay([device_ragged(x, node_len), de, device_ragged(index, edge_len)]).values
    assert lay.bias.detach()[units:2 * units].eq(1).all() and lay.bias.detach()[:units].eq(0).all()   # unit_forget_bias
    with torch.no_grad():
        lay.bias.add_(0.1 * torch.randn_like(lay.bias))
    got = lay([device_ragged(x, node_len), de, device_ragged(index, edge_len)]).values
    proj = torch.as_tensor(rng.standard_normal(tuple(got.shape)))
    params = [lay.kernel, lay.recurrent_kernel, lay.bias]
    grads = torch.autograd.grad((got * proj.to(got)).sum(), [de.values] + params)

    oe = oracle_ragged(e, edge_len, torch.float64)
    oe.values.requires_grad_(True)
    op = [p.detach().cpu().double().requires_grad_(True) for p in params]
    want = OL.pooling_local_edges_lstm(oracle_ragged(x, node_len), oe, oracle_ragged(index, edge_len), op[0], op[1], op[2],
                                       go_backwards=go_backwards).values
    o_grads = torch.autograd.grad((want * proj).sum(), [oe.values] + op)
    assert tuple(got.shape) == (int(node_len.sum()), units)
    assert rel_err(got.detach().cpu().numpy(), want.detach().numpy()) < RTOL
    for g, o in zip(grads, o_grads):
        assert rel_err(g.cpu().numpy(), o.numpy()) < RTOL
    deg = np.bincount(index[:, 0] + np.repeat(np.concatenate([[0], np.cumsum(node_len)[:-1]]), edge_len),
                      minlength=int(node_len.sum()))
    assert float(got.detach()[torch.as_tensor(deg == 0).cuda()].abs().max()) == 0.0       # no edges: initial state


def test_pooling_local_edges_lstm_config_and_errors():
    from gcnn_keras_b200.layers.base import get_registered
    from gcnn_keras_b200.layers.pooling import PoolingLocalEdgesLSTM
    assert get_registered("kgcnn>PoolingLocalEdgesLSTM") is PoolingLocalEdgesLSTM
    cfg = PoolingLocalEdgesLSTM(units=8).get_config()
    for key in ("pooling_method", "pooling_index", "activation", "recurrent_activation", "use_bias", "unit_forget_bias",
                "return_sequences", "go_backwards", "unroll"):
        assert key in cfg
    with pytest.raises(ValueError):
        PoolingLocalEdgesLSTM(units=8, return_sequences=True)
    # initializers / constraints other than the ones the layer implements are refused, not echoed and ignored
    PoolingLocalEdgesLSTM(units=8, kernel_initializer="GlorotUniform", bias_initializer={"class_name": "Zeros", "config": {}})
    with pytest.raises(ValueError, match="kernel_initializer"):
        PoolingLocalEdgesLSTM(units=8, kernel_initializer="he_normal")
    with pytest.raises(ValueError, match="recurrent_initializer"):
        PoolingLocalEdgesLSTM(units=8, recurrent_initializer="glorot_uniform")
    with pytest.raises(ValueError, match="bias_constraint"):
        PoolingLocalEdgesLSTM(units=8, bias_constraint="non_neg")


def test_relation_max_degree_is_cached_with_the_plan():
    from gcnn_keras_b200.plan import Relation
    ids = torch.tensor([0, 0, 0, 2, 2, 5], dtype=torch.int32).cuda()
    rel = Relation(ids, 7)
    assert rel.max_degree() == 3
    rel._rowptr = None                      # a second call must not touch the device plan again
    assert rel.max_degree() == 3
    assert Relation(ids[:0], 4).max_degree() == 0
