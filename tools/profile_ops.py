"""Which eager ATen ops still run inside a training step: device time grouped by (op, input shapes, innermost
repo frame).  Not a bench number.   python tools/profile_ops.py painn --batch 512
"""
import argparse
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", nargs="?", default="painn")
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--rows", type=int, default=60)
    args = ap.parse_args()
    device = torch.device("cuda", 0)
    build = {"schnet": bench.build_schnet, "painn": bench.build_painn, "hdnnp": bench.build_hdnnp}[args.workload]
    model, step, data = build(device, args.batch, 0)
    from gcnn_keras_b200.training import StaticBatch, default_model_inputs, molecule_batch_to_arrays
    static = StaticBatch(molecule_batch_to_arrays(data), device,
                         device_angles="kj" if args.workload == "hdnnp" else None)
    static.load()
    max_nodes = int(data["node_len"].max())

    def one():
        inputs, e_t, f_t = default_model_inputs(static.dev, max_nodes)
        step.step(inputs, e_t, f_t)

    for _ in range(2):
        one()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=True) as prof:
        one()
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0.0, 0])
    for e in prof.events():
        if not e.name.startswith("aten::") or e.self_device_time_total <= 0:
            continue
        frame = ""
        for s in e.stack:
            if "gcnn_keras_b200" in s and "torch/" not in s:
                frame = s.split("gcnn_keras_b200/")[-1]
                break
        key = (e.name, str(e.input_shapes)[:70], frame[:70])
        agg[key][0] += e.self_device_time_total
        agg[key][1] += 1
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    total = sum(v[0] for v in agg.values())
    print("eager ATen device time per step: %.3f ms in %d ops" % (total / 1e3, sum(v[1] for v in agg.values())))
    for (name, shapes, frame), (t, n) in rows[:args.rows]:
        print("%8.1f us x%-4d %-22s %-70s %s" % (t, n, name, shapes, frame))


if __name__ == "__main__":
    main()
