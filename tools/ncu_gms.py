"""Runs the north-star kernel (fused gather -> CSR segment-sum) a few times on GSS-R and GSS-L for an
`ncu --set full -k regex:gms` capture (profiling helper, never a bench number).
    ncu --set full --clock-control none --import-source on -k regex:gms -c 4 -o gpurun_out/gms python tools/ncu_gms.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

if __name__ == "__main__":
    print(bench.gss_roofline(torch.device("cuda", 0), iters=1))
