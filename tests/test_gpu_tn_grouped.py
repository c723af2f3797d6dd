"""Grouped weight-gradient kernel (kgcnn_dense_tn_grouped) and the deferred sink against fp64, and the training
step with / without the sink."""
import numpy as np
import pytest
import torch

from helpers import RTOL, rel_err

pytestmark = pytest.mark.gpu


def _flat_params(shapes, device="cuda"):
    from gcnn_keras_b200.training import FlatAdam
    params = [torch.nn.Parameter(torch.randn(*s, device=device)) for s in shapes]
    opt = FlatAdam(params)
    return params, opt


def test_grouped_tn_matches_fp64_and_accumulates_shared_outputs():
    from gcnn_keras_b200.functional import WeightGradSink
    torch.manual_seed(0)
    shapes = [(128, 128), (128,), (25, 128), (128,), (384, 20), (7, 3), (130, 260), (260,)]
    params, opt = _flat_params(shapes)
    sink = WeightGradSink(params, opt.gflat, opt.offsets)
    cases = [  # (w index, bias index | None, M)
        (0, 1, 18432), (0, None, 18432), (2, 3, 5000), (4, None, 777), (5, None, 1), (6, 7, 3001), (0, None, 64),
        (5, None, 0)]
    want = [torch.zeros(s, dtype=torch.float64) for s in shapes]
    for wi, bi, M in cases:
        K, U = shapes[wi]
        x, g = torch.randn(M, K, device="cuda"), torch.randn(M, U, device="cuda")
        assert sink.add(params[wi], x, g, None if bi is None else params[bi])
        want[wi] += x.double().cpu().t() @ g.double().cpu()
        if bi is not None:
            want[bi] += g.double().cpu().sum(0)
    sink.flush()
    torch.cuda.synchronize()
    for view, w in zip(opt.grad_views, want):
        assert rel_err(view.cpu().numpy(), w.numpy()) < RTOL
    # second flush accumulates on top; identical list -> exactly twice (bitwise repeatable partial sums)
    first = opt.gflat.clone()
    torch.manual_seed(0)
    _ = [torch.nn.Parameter(torch.randn(*s, device="cuda")) for s in shapes]     # replay the RNG stream
    for wi, bi, M in cases:
        K, U = shapes[wi]
        x, g = torch.randn(M, K, device="cuda"), torch.randn(M, U, device="cuda")
        sink.add(params[wi], x, g, None if bi is None else params[bi])
    sink.flush()
    assert rel_err(opt.gflat.cpu().numpy(), (2 * first).cpu().numpy()) < 1e-6


def test_sink_rejects_unregistered_or_reshaped_weights():
    from gcnn_keras_b200.functional import WeightGradSink
    params, opt = _flat_params([(16, 8)])
    sink = WeightGradSink(params, opt.gflat, opt.offsets)
    x, g = torch.randn(5, 16, device="cuda"), torch.randn(5, 8, device="cuda")
    assert not sink.add(torch.randn(16, 8, device="cuda"), x, g)
    assert not sink.add(params[0].view(8, 16), torch.randn(5, 8, device="cuda"), torch.randn(5, 16, device="cuda"))
    assert sink.items == []


def test_linear_backward_through_sink_equals_autograd_path():
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.functional import WeightGradSink, weight_grad_sink
    torch.manual_seed(1)
    params, opt = _flat_params([(64, 32), (32,), (32, 48)])
    w1, b1, w2 = params
    x = torch.randn(3000, 64, device="cuda", requires_grad=True)

    def loss_fn():
        h = F.dense(x, w1, b1, "swish")
        y = F.matmul(h, w2)
        (gx,) = torch.autograd.grad(y.sum() + (y ** 2).sum(), x, create_graph=True)    # force-like inner derivative
        return (gx ** 2).sum() + y.sum()

    ref = torch.autograd.grad(loss_fn(), params)
    opt.zero_grad()
    with weight_grad_sink(WeightGradSink(params, opt.gflat, opt.offsets)):
        got = torch.autograd.grad(loss_fn(), params, allow_unused=True)
    assert all(g is None for g in got)                 # everything went through the sink
    for view, r in zip(opt.grad_views, ref):
        assert rel_err(view.cpu().numpy(), r.cpu().numpy()) < RTOL


def test_grouped_tn_large_edge_problems():
    """PaiNN C3 shapes (405 k edge rows, 20 -> 384 filter Dense, twelve times, plus node-level problems): the
    launch spans several waves of CTAs with many row splits per problem; results against fp64."""
    from gcnn_keras_b200.functional import WeightGradSink
    torch.manual_seed(3)
    shapes = [(20, 384), (384,), (128, 128), (128,), (128, 384), (384,)]
    params, opt = _flat_params(shapes)
    sink = WeightGradSink(params, opt.gflat, opt.offsets)
    want = [torch.zeros(s, dtype=torch.float64) for s in shapes]
    cases = [(0, 1, 405180)] * 6 + [(2, 3, 15360)] * 4 + [(4, 5, 15360)] * 2 + [(0, None, 405180)] * 6
    for wi, bi, M in cases:
        K, U = shapes[wi]
        x, g = torch.randn(M, K, device="cuda"), torch.randn(M, U, device="cuda")
        assert sink.add(params[wi], x, g, None if bi is None else params[bi])
        want[wi] += (x.double().t() @ g.double()).cpu()
        if bi is not None:
            want[bi] += g.double().sum(0).cpu()
    sink.flush()
    torch.cuda.synchronize()
    for view, w in zip(opt.grad_views, want):
        assert rel_err(view.cpu().numpy(), w.numpy()) < RTOL
