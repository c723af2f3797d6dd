"""tcgen05 (3xTF32) Dense contractions NN / NT / TN against float64 matmul."""
import numpy as np
import pytest
import torch

from helpers import RTOL, rel_err

pytestmark = pytest.mark.gpu

SHAPES = [(18432, 128, 128), (18432, 25, 128), (2304, 128, 128), (300, 64, 1), (1000, 35, 35), (5000, 20, 384),
          (129, 8, 16), (4100, 640, 35), (70, 3, 5)]


@pytest.mark.parametrize("M,K,U", SHAPES)
def test_dense_tc_nn_nt_tn(M, K, U):
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(M + K + U)
    x = rng.standard_normal((M, K)).astype(np.float32)
    w = (rng.standard_normal((K, U)) / np.sqrt(K)).astype(np.float32)
    b = rng.standard_normal((U,)).astype(np.float32)
    g = rng.standard_normal((M, U)).astype(np.float32)
    xd, wd, bd, gd = (torch.tensor(t).cuda() for t in (x, w, b, g))
    x64, w64, g64 = x.astype(np.float64), w.astype(np.float64), g.astype(np.float64)
    y = F.raw_dense_tc(0, xd, wd, bd)
    assert rel_err(y.cpu().numpy(), x64 @ w64 + b) < RTOL
    y0 = F.raw_dense_tc(0, xd, wd, None)
    assert rel_err(y0.cpu().numpy(), x64 @ w64) < RTOL
    dx = F.raw_dense_tc(1, gd, wd)
    assert rel_err(dx.cpu().numpy(), g64 @ w64.T) < RTOL
    dw = F.raw_dense_tc(2, xd, gd)
    assert rel_err(dw.cpu().numpy(), x64.T @ g64) < RTOL
    assert torch.equal(dw, F.raw_dense_tc(2, xd, gd))        # deterministic


ACTS = ["kgcnn>shifted_softplus", "swish", "tanh", "relu", "sigmoid", "kgcnn>leaky_relu"]


@pytest.mark.parametrize("act", ACTS)
@pytest.mark.parametrize("M,K,U", [(18432, 128, 128), (1000, 35, 35), (129, 8, 16), (70, 3, 5), (5000, 20, 384)])
def test_dense_act_epilogue_matches_two_launch_path(M, K, U, act):
    """kgcnn_dense_act (activation in the GEMM epilogue) == GEMM launch + bias/activation launch (pre bit for bit:
    same accumulator; the activation to 1e-6), and within RTOL of float64."""
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(M + K + U)
    x = torch.tensor(rng.standard_normal((M, K)).astype(np.float32)).cuda()
    w = torch.tensor((rng.standard_normal((K, U)) / np.sqrt(K)).astype(np.float32)).cuda()
    b = torch.tensor(rng.standard_normal((U,)).astype(np.float32)).cuda()
    code = F.act_code(act)
    pre, y = F.raw_dense_act(x, w, b, code)
    pre_ref = F.raw_dense_tc(0, x, w, b)
    assert torch.equal(pre, pre_ref)
    assert rel_err(y.cpu().numpy(), F.raw_bias_act(pre_ref, None, None, None, code, 0).cpu().numpy()) < 1e-6
    p64 = x.double() @ w.double() + b.double()
    f64 = {"kgcnn>shifted_softplus": lambda t: torch.nn.functional.softplus(t) - np.log(2.0),
           "swish": lambda t: t * torch.sigmoid(t), "tanh": torch.tanh, "relu": torch.relu, "sigmoid": torch.sigmoid,
           "kgcnn>leaky_relu": lambda t: torch.where(t > 0, t, 0.05 * t)}[act]
    assert rel_err(y.cpu().numpy(), f64(p64).cpu().numpy()) < RTOL


@pytest.mark.parametrize("act", ["kgcnn>shifted_softplus", "swish", "tanh"])
def test_dense_fused_activation_first_and_second_derivatives(act):
    """F.dense with the fused epilogue gives the same value, gradients and double-backward gradients (the force
    training pattern: d/dW of a function of dy/dx) as the two-launch path."""
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(7)
    M, K, U = 777, 64, 128
    x0 = torch.tensor(rng.standard_normal((M, K)).astype(np.float32)).cuda()
    w0 = torch.tensor((rng.standard_normal((K, U)) / np.sqrt(K)).astype(np.float32)).cuda()
    w1 = torch.tensor((rng.standard_normal((U, 1)) / np.sqrt(U)).astype(np.float32)).cuda()
    b0 = torch.tensor(rng.standard_normal((U,)).astype(np.float32)).cuda()
    c = torch.tensor(rng.standard_normal((M, K)).astype(np.float32)).cuda()

    def run(fuse):
        F.DENSE_FUSE_ACT = fuse
        try:
            x, w, b, v = (t.clone().requires_grad_(True) for t in (x0, w0, b0, w1))
            e = F.dense(F.dense(x, w, b, act), v, None, "linear").sum()
            (dx,) = torch.autograd.grad(e, x, create_graph=True)
            loss = (dx * c).sum() + e
            gw, gb, gv = torch.autograd.grad(loss, (w, b, v))
            return e.detach(), dx.detach(), gw, gb, gv
        finally:
            F.DENSE_FUSE_ACT = True

    for a, r in zip(run(True), run(False)):
        assert rel_err(a.cpu().numpy(), r.cpu().numpy()) < 1e-6


@pytest.mark.parametrize("act", ACTS)
@pytest.mark.parametrize("order", [0, 1, 2])
@pytest.mark.parametrize("M,U,with_bias", [(18432, 128, False), (777, 35, True), (70, 5, True)])
def test_bias_act_pair_matches_the_two_single_launches(act, order, M, U, with_bias):
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(M + U + order)
    pre, gy, g1 = (torch.tensor(rng.standard_normal((M, U)).astype(np.float32)).cuda() for _ in range(3))
    b = torch.tensor(rng.standard_normal((U,)).astype(np.float32)).cuda() if with_bias else None
    code = F.act_code(act)
    d_pre, d_g1 = F.raw_bias_act_pair(pre, b, gy, g1, code, order)
    assert rel_err(d_pre.cpu().numpy(), F.raw_bias_act(pre, b, gy, g1, code, order + 1).cpu().numpy(), ) < 1e-6
    assert rel_err(d_g1.cpu().numpy(), F.raw_bias_act(pre, b, gy, None, code, order).cpu().numpy()) < 1e-6


def test_bias_act_pair_in_the_last_sweep_equals_two_launch_backward():
    from gcnn_keras_b200 import functional as F
    rng = np.random.default_rng(11)
    M, K, U = 500, 32, 64
    x0 = torch.tensor(rng.standard_normal((M, K)).astype(np.float32)).cuda()
    w0 = torch.tensor((rng.standard_normal((K, U)) / np.sqrt(K)).astype(np.float32)).cuda()
    b0 = torch.tensor(rng.standard_normal((U,)).astype(np.float32)).cuda()
    c = torch.tensor(rng.standard_normal((M, K)).astype(np.float32)).cuda()

    def run(pair):
        F.BIAS_ACT_PAIR = pair
        try:
            x, w, b = (t.clone().requires_grad_(True) for t in (x0, w0, b0))
            e = F.dense(x, w, b, "swish").sum()
            (dx,) = torch.autograd.grad(e, x, create_graph=True)
            return torch.autograd.grad((dx * c).sum(), (w, b))
        finally:
            F.BIAS_ACT_PAIR = True

    for a, r in zip(run(True), run(False)):
        assert rel_err(a.cpu().numpy(), r.cpu().numpy()) < 1e-6
