"""The two Dense shapes of the SchNet C2 step (edge rows 18 432, atom rows 2 304; K = U = 128) through the NN / NT
kernels and the fused Dense+activation entry, for an ncu capture (profiling helper, never a bench number).
    ncu --set full --clock-control none --import-source on -k regex:dense_tc_kernel -s 12 -c 6 -f -o gpurun_out/dense_small \
        python tools/ncu_dense_small.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gcnn_keras_b200 import functional as F  # noqa: E402

if __name__ == "__main__":
    for rep in range(3):                       # launches 0-11 warm up, 12-17 are captured
        for M in (18432, 2304):
            x = torch.randn(M, 128, device="cuda")
            w = torch.randn(128, 128, device="cuda")
            b = torch.randn(128, device="cuda")
            F.raw_dense_tc(0, x, w, b)
            F.raw_dense_tc(1, x, w)
            F.raw_dense_act(x, w, b, F.act_code("kgcnn>shifted_softplus"))
    torch.cuda.synchronize()
