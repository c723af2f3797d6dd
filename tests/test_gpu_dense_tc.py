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
