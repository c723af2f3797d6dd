"""RelationalDense / RelationalMLP (kgcnn/layers/relational.py:219-238, mlp.py:329-416) through the grouped tcgen05
launch, against the reference formula (per-row kernel gather + batch_dot) in float64."""
import numpy as np
import pytest
import torch

from helpers import RTOL
from helpers import rel_err as _rel_err

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return _rel_err(a.detach().cpu().numpy(), b.detach().cpu().numpy())


def _reference(x, rel, kernel, bias, act):
    """relational.py:233-237 in torch float64."""
    y = torch.einsum("nk,nku->nu", x, kernel[rel]) + (0 if bias is None else bias)
    if act == "swish":
        return y * torch.sigmoid(y)
    return y


@pytest.mark.parametrize("n,k,u,r", [(1000, 40, 64, 7), (4096, 64, 1, 96), (777, 33, 25, 4), (130, 640, 64, 3)])
def test_relational_dense_values_and_two_derivative_orders(n, k, u, r):
    from gcnn_keras_b200.layers.relational import RelationalDense
    from gcnn_keras_b200.ragged import RaggedTensor as R
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(n + k)
    rel = torch.randint(0, r, (n,), generator=gen)
    rel[rel == 1] = 0                                          # relation 1 has no rows
    x = torch.randn(n, k, generator=gen)
    lens = torch.tensor([n // 3, n - n // 3])
    layer = RelationalDense(units=u, num_relations=r, activation="swish", use_bias=True)
    xr = R.from_row_lengths(x.to(dev).requires_grad_(True), lens.to(dev))
    rr = R.from_row_lengths(rel.to(dev), lens.to(dev))
    out = layer([xr, rr])
    with torch.no_grad():
        layer.bias.copy_(torch.randn(u, generator=gen).to(dev) * 0.1)
    out = layer([xr, rr])
    x64 = x.double().requires_grad_(True)
    k64 = layer.kernel.detach().double().cpu().requires_grad_(True)
    b64 = layer.bias.detach().double().cpu().requires_grad_(True)
    ref = _reference(x64, rel, k64, b64, "swish")
    assert rel_err(out.values, ref) < RTOL
    # first derivative w.r.t. the rows (kept differentiable), then a loss on it: exercises NT, TN and their VJPs
    cot = torch.randn(n, u, generator=gen)
    gx, = torch.autograd.grad(out.values, xr.values, cot.to(dev), create_graph=True)
    gx64, = torch.autograd.grad(ref, x64, cot.double(), create_graph=True)
    assert rel_err(gx, gx64) < RTOL
    w = torch.randn(n, k, generator=gen)
    loss = (gx * w.to(dev)).sum() + (out.values * out.values).sum() * 0.5
    loss64 = (gx64 * w.double()).sum() + (ref * ref).sum() * 0.5
    gk, gb, gxx = torch.autograd.grad(loss, [layer.kernel, layer.bias, xr.values])
    gk64, gb64, gxx64 = torch.autograd.grad(loss64, [k64, b64, x64])
    assert rel_err(gk, gk64) < RTOL
    assert rel_err(gb, gb64) < RTOL
    assert rel_err(gxx, gxx64) < RTOL
    assert float(gk[1].abs().max()) == 0.0                     # the empty relation's kernel gets an exact zero


def test_relational_mlp_sorts_once_and_matches_per_layer_formula():
    from gcnn_keras_b200.layers.relational import RelationalMLP
    from gcnn_keras_b200.ragged import RaggedTensor as R
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(5)
    n, k, r = 3000, 48, 20
    rel = torch.randint(0, r, (n,), generator=gen)
    x = torch.randn(n, k, generator=gen)
    lens = torch.full((30,), 100)
    mlp = RelationalMLP(units=[64, 64, 1], num_relations=r, activation=["swish", "swish", "linear"])
    out = mlp([R.from_row_lengths(x.to(dev), lens.to(dev)), R.from_row_lengths(rel.to(dev), lens.to(dev))])
    h = x.double()
    for layer, act in zip(mlp.mlp_dense_layer_list, ["swish", "swish", "linear"]):
        h = _reference(h, rel, layer.kernel.detach().double().cpu(), layer.bias.detach().double().cpu(), act)
    assert rel_err(out.values, h) < RTOL
    assert out.values.shape == (n, 1)
    # already sorted relations take the no-permutation path and give the same rows
    order = torch.argsort(rel, stable=True)
    out_s = mlp([R.from_row_lengths(x[order].to(dev), lens.to(dev)),
                 R.from_row_lengths(rel[order].to(dev), lens.to(dev))])
    assert torch.equal(out_s.values, out.values[order.to(dev)])


def test_relation_out_of_range_raises():
    from gcnn_keras_b200.layers.relational import RelationalDense
    from gcnn_keras_b200.ragged import RaggedTensor as R
    dev = torch.device("cuda:0")
    lens = torch.tensor([4, 4], device=dev)
    x = R.from_row_lengths(torch.randn(8, 16, device=dev), lens)
    rel = R.from_row_lengths(torch.tensor([0, 1, 2, 3, 4, 1, 0, 2], device=dev), lens)
    with pytest.raises(IndexError):
        RelationalDense(units=8, num_relations=4)([x, rel])
