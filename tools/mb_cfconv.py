"""Fused cfconv (csrc/cfconv.cu) against the unfused chain (basis -> Dense -> Dense -> gather-multiply-pool) on the same
inputs: forward, and forward + input gradients (the inner dE/dx sweep).  CUDA events, back to back, after warm-up.
    python tools/mb_cfconv.py [--big]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gcnn_keras_b200 import RaggedTensor, functional as F  # noqa: E402
from gcnn_keras_b200.plan import get_plan  # noqa: E402


def case(n_nodes, deg, seed=0):
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda", 0)
    recv = np.repeat(np.arange(n_nodes), deg)
    send = (recv + rng.integers(1, min(n_nodes, 4096), size=recv.shape)) % n_nodes
    idx = torch.tensor(np.stack([recv, send], 1), device=dev)
    E = idx.shape[0]
    nodes = RaggedTensor.from_row_lengths(torch.randn(n_nodes, 128, device=dev), torch.tensor([n_nodes], device=dev))
    index = RaggedTensor.from_row_lengths(idx, torch.tensor([E], device=dev))
    plan = get_plan(index, nodes)
    spec = F.CfconvSpec(plan.relation(0), plan.relation(1), 25, 3.125, 0.0, 0.2, F.act_code("kgcnn>shifted_softplus"))
    t = [nodes.values.clone().requires_grad_(True), (torch.rand(E, device=dev) * 4 + 0.8).requires_grad_(True),
         (torch.randn(25, 128, device=dev) * 0.4).requires_grad_(True), torch.zeros(128, device=dev, requires_grad=True),
         (torch.randn(128, 128, device=dev) * 0.15).requires_grad_(True), torch.zeros(128, device=dev, requires_grad=True)]
    return spec, t, E


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def main():
    big = "--big" in sys.argv
    shapes = [(2304, 8), (15360, 29)] + ([(262144, 32), (1 << 20, 16)] if big else [])
    print("%-22s %12s %12s %12s %12s   (us; fwd / fwd + input gradients)" % ("N x deg", "fused fwd", "unfused fwd", "fused f+b", "unfused f+b"))
    for n, deg in shapes:
        spec, t, E = case(n, deg)
        cot = torch.randn(n, 128, device="cuda")

        def fwd(fused):
            with torch.no_grad():
                return F.cfconv_fused(*t, spec) if fused else F._cfconv_unfused(*t, spec)

        def fb(fused):
            y = F.cfconv_fused(*t, spec) if fused else F._cfconv_unfused(*t, spec)
            with F.input_gradients_only(t[0], t[1]):
                return torch.autograd.grad(y, [t[0], t[1]], cot)

        reps = 20 if E > 1e6 else 100
        r = [timed(lambda: fwd(True), reps), timed(lambda: fwd(False), reps), timed(lambda: fb(True), reps),
             timed(lambda: fb(False), reps)]
        print("%-22s %12.1f %12.1f %12.1f %12.1f   E=%d  fused fwd %.2f G edges/s" % ("%d x %d" % (n, deg), *r, E, E / r[0] / 1e3))


if __name__ == "__main__":
    main()
