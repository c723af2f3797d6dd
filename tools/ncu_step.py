"""One eager training step of a force workload for an `ncu -k regex:<kernel>` capture (profiling helper, never a
bench number).   python tools/ncu_step.py painn|hdnnp|schnet [batch]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gcnn_keras_b200.training import StaticBatch, default_model_inputs, molecule_batch_to_arrays  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda", 0)
    kind = sys.argv[1] if len(sys.argv) > 1 else "painn"
    build, default_batch = bench.FORCE_WORKLOADS[kind][0], bench.FORCE_WORKLOADS[kind][1]
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else default_batch
    model, step, data = build(dev, batch, 0)
    static = StaticBatch(molecule_batch_to_arrays(data), dev)
    static.load()
    inputs, e_t, f_t = default_model_inputs(static.dev, int(data["node_len"].max()))
    loss = step.step(inputs, e_t, f_t)
    torch.cuda.synchronize()
    print("loss", float(loss.detach()))
