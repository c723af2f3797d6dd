"""One launch of every kernel of the roofline block (bench.roofline_shapes: plain GSS-R, cfconv variant, GSS-G, GSS-L), in
that order, for an `ncu --set full` capture (profiling helper, never a bench number):
    ncu --set full --clock-control none --import-source on -k regex:"gms_kernel|gather2_mul" -f -o gpurun_out/gms_full \\
        python tools/ncu_gms.py
    python profiles/summarize.py gpurun_out/gms_full.ncu-rep profiles/r2_gms_full.csv      # read by bench.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

if __name__ == "__main__":
    for key, workload, kernel, alg_bytes, edges, fns in bench.roofline_shapes(torch.device("cuda", 0)):
        for f in fns:
            f()
        torch.cuda.synchronize()
        print(key, workload, alg_bytes)
