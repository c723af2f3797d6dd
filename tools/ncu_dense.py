"""A few representative Dense shapes through raw_dense_tc for an ncu capture (profiling helper)."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gcnn_keras_b200 import functional as F
for (M, K, U) in [(404544, 128, 128), (404544, 128, 384), (18432, 128, 128)]:
    x = torch.randn(M, K, device="cuda"); w = torch.randn(K, U, device="cuda"); g = torch.randn(M, U, device="cuda")
    for _ in range(2):
        F.raw_dense_tc(0, x, w); F.raw_dense_tc(1, g, w); F.raw_dense_tc(2, x, g)
torch.cuda.synchronize()
