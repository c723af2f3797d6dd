"""Kernel-time breakdown of one captured training step (torch.profiler / CUPTI; not a bench number).
    python tools/profile_step.py [schnet] [--batch 256] > gpurun_out/step_kernels.txt
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", nargs="?", default="schnet")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--rows", type=int, default=45)
    args = ap.parse_args()
    device = torch.device("cuda", 0)
    torch.backends.cuda.matmul.allow_tf32 = False
    from gcnn_keras_b200.training import CapturedForceTrainStep
    build = {"schnet": bench.build_schnet, "painn": getattr(bench, "build_painn", None),
             "hdnnp": getattr(bench, "build_hdnnp", None)}[args.workload]
    model, step, data = build(device, args.batch, 0)
    cap = CapturedForceTrainStep(step, data, device)
    for _ in range(3):
        cap.run(load=False)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            cap.run(load=False)
        torch.cuda.synchronize()
    ev = prof.key_averages()
    rows = sorted(ev, key=lambda e: -e.device_time_total)
    total = sum(e.device_time_total for e in rows)
    n = sum(e.count for e in rows)
    print("total device time per step: %.3f ms, kernels per step: %d" % (total / 5 / 1e3, n // 5))
    for e in rows[:args.rows]:
        print("%8.1f us/step  %5.1f%%  x%-5d %s" % (e.device_time_total / 5, 100 * e.device_time_total / total,
                                                   e.count // 5, e.key[:110]))


if __name__ == "__main__":
    main()
