"""One eager HDNNP C4 training step (batch 256) for an `ncu --set full -k regex:acsf_g4` capture (profiling helper,
never a bench number).
    ncu --set full --clock-control none --import-source on -k regex:acsf_g4 -c 3 -o gpurun_out/g4 python tools/ncu_g4.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gcnn_keras_b200.training import StaticBatch, default_model_inputs, molecule_batch_to_arrays  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda", 0)
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    model, step, data = bench.build_hdnnp(dev, batch, 0)
    static = StaticBatch(molecule_batch_to_arrays(data), dev)
    static.load()
    inputs, e_t, f_t = default_model_inputs(static.dev, int(data["node_len"].max()))
    loss = step.step(inputs, e_t, f_t)
    torch.cuda.synchronize()
    print("loss", float(loss))
