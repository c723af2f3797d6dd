"""Microbenchmark + accuracy of the tcgen05 Dense kernels at the shapes of the BASELINE configs (diagnostics)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gcnn_keras_b200.functional as F

torch.backends.cuda.matmul.allow_tf32 = False
flush = torch.empty(64 * 1024 * 1024, device="cuda")


def t(fn, iters=50, cold=False):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    if cold:
        for _ in range(iters):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / iters * 1e3
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters // 10):
        g.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / (iters // 10 * 20) * 1e3


def err(a, b):
    return float((a.double() - b).abs().max() / b.abs().max())


print("%-22s %9s %9s %9s   %9s %9s %9s   (us per launch in a graph, back to back; rel err vs fp64)" %
      ("shape M,K,U", "NN", "NT", "TN", "errNN", "errNT", "errTN"))
for M, K, U in [(18432, 128, 128), (2304, 128, 128), (18432, 25, 128), (2304, 128, 64), (256, 64, 1), (404544, 128, 128),
                (404544, 20, 384), (404544, 64, 128), (404544, 32, 128), (2000000, 128, 128), (50000, 20, 384), (15360, 128, 384), (16384, 640, 35), (16384, 35, 35)]:
    x = torch.randn(M, K, device="cuda"); w = torch.randn(K, U, device="cuda") / K ** 0.5
    g = torch.randn(M, U, device="cuda"); bias = torch.randn(U, device="cuda")
    r = [t(lambda: F.raw_dense_tc(0, x, w, bias)), t(lambda: F.raw_dense_tc(1, g, w)), t(lambda: F.raw_dense_tc(2, x, g))]
    e = [err(F.raw_dense_tc(0, x, w, bias), x.double() @ w.double() + bias.double()),
         err(F.raw_dense_tc(1, g, w), g.double() @ w.double().t()),
         err(F.raw_dense_tc(2, x, g), x.double().t() @ g.double())]
    print("%-22s %9.2f %9.2f %9.2f   %9.1e %9.1e %9.1e" % ((M, K, U), r[0], r[1], r[2], e[0], e[1], e[2]))
x = torch.randn(18432, 128, device="cuda"); w = torch.randn(128, 128, device="cuda") / 11
print("dense_act [18432,128]x[128,128]: %.2f us" % t(lambda: F.raw_dense_act(x, w, None, 1)))
print("fp32 matmul (torch) same shape: %.2f us" % t(lambda: x @ w))
