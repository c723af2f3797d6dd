import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from gcnn_keras_b200 import functional as F
torch.backends.cuda.matmul.allow_tf32 = False
rng = np.random.default_rng(0)
def rel(a, b): return float(np.abs(a - b).max() / np.abs(b).max())
for (M, K, U, scale) in [(36, 640, 1050, -1.0), (18432, 128, 128, -1.0), (36, 640, 1050, 1.0), (36, 640, 1050, 100.0), (18432, 128, 128, 1.0), (36, 35, 1050, 1.0), (4096, 640, 35, 30.0)]:
    x = (rng.standard_normal((M, K)) * scale).astype(np.float32); w = (rng.standard_normal((K, U)) / np.sqrt(K)).astype(np.float32)
    g = rng.standard_normal((M, U)).astype(np.float32)
    if scale < 0: x, w, g = np.abs(x), np.abs(w), np.abs(g)   # all-positive data exposes one-sided rounding
    xd, wd, gd = (torch.tensor(t).cuda() for t in (x, w, g))
    x64, w64, g64 = x.astype(np.float64), w.astype(np.float64), g.astype(np.float64)
    print(M, K, U, scale, "NN tc %.2e fp32 %.2e | NT tc %.2e fp32 %.2e | TN tc %.2e fp32 %.2e" % (
        rel(F.raw_dense_tc(0, xd, wd).cpu().numpy(), x64 @ w64), rel((xd @ wd).cpu().numpy(), x64 @ w64),
        rel(F.raw_dense_tc(1, gd, wd).cpu().numpy(), g64 @ w64.T), rel((gd @ wd.t()).cpu().numpy(), g64 @ w64.T),
        rel(F.raw_dense_tc(2, xd, gd).cpu().numpy(), x64.T @ g64), rel((xd.t() @ gd).cpu().numpy(), x64.T @ g64)))
