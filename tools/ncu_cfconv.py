"""Workload for the ncu capture of the fused cfconv at an HBM-sized graph (N = 262 144, 32 edges per node, F = 128):
forward (mode 0) and forward + input gradients (mode 1) a few times.  Not a bench."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from gcnn_keras_b200 import functional as F  # noqa: E402
import mb_cfconv  # noqa: E402

spec, t, E = mb_cfconv.case(262144, 32)
cot = torch.randn(262144, 128, device="cuda")
for _ in range(3):
    y = F.cfconv_fused(*t, spec)
    with F.input_gradients_only(t[0], t[1]):
        torch.autograd.grad(y, [t[0], t[1]], cot)
torch.cuda.synchronize()
print("E", E)
