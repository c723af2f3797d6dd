"""Fused Keras-Adam + per-variable clipnorm kernel against a float64 restatement of the Keras update rule."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu


def test_flat_adam_clipnorm_matches_keras_rule():
    from gcnn_keras_b200.training import FlatAdam
    rng = np.random.default_rng(0)
    shapes = [(25, 128), (128,), (128, 128), (1,), (95, 16)]
    params = [torch.nn.Parameter(torch.tensor(rng.standard_normal(s).astype(np.float32), device="cuda"))
              for s in shapes]
    ref = [p.detach().cpu().double().numpy().copy() for p in params]
    m = [np.zeros_like(r) for r in ref]
    v = [np.zeros_like(r) for r in ref]
    opt = FlatAdam(params, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, clipnorm=1.0)
    for t in range(1, 4):
        grads = [rng.standard_normal(s).astype(np.float32) * (10.0 if i % 2 == 0 else 0.01)
                 for i, s in enumerate(shapes)]
        opt.step([torch.tensor(g, device="cuda") for g in grads])
        for i, g in enumerate(grads):
            g = g.astype(np.float64)
            g = g * (1.0 / max(np.sqrt((g * g).sum()), 1.0))              # tf.clip_by_norm(g, 1.0)
            m[i] = 0.9 * m[i] + 0.1 * g
            v[i] = 0.999 * v[i] + 0.001 * g * g
            lr_t = 1e-3 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)          # keras Adam
            ref[i] = ref[i] - lr_t * m[i] / (np.sqrt(v[i]) + 1e-7)
    assert int(opt.step_count.item()) == 3
    for p, r in zip(params, ref):
        assert rel_err(p.detach().cpu().numpy(), r) < 1e-6
