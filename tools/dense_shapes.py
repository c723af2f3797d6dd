"""Diagnostics: time of every Dense contraction of one eager training step, aggregated by (op, M, K, U)."""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gcnn_keras_b200.functional as F  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "painn"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    device = torch.device("cuda", 0)
    build = {"schnet": bench.build_schnet, "painn": bench.build_painn, "hdnnp": bench.build_hdnnp}[workload]
    model, step, data = build(device, batch or {"schnet": 256, "painn": 512, "hdnnp": 256}[workload], 0)
    from gcnn_keras_b200.training import CapturedForceTrainStep
    cap = CapturedForceTrainStep(step, data, device)      # warm-up happens inside
    records = []
    raw = F.raw_dense_tc

    def timed(op, a, b, bias=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = raw(op, a, b, bias)
        e1.record()
        M = int(a.shape[0])
        K, U = (int(b.shape[0]), int(b.shape[1])) if op < 2 else (int(a.shape[1]), int(b.shape[1]))
        records.append(((op, M, K, U), e0, e1))
        return out

    F.raw_dense_tc = timed
    from gcnn_keras_b200.plan import clear_plan_cache
    clear_plan_cache()
    inputs, e_t, f_t = cap.make_inputs(cap.static.dev, cap.max_nodes)
    step.step(inputs, e_t, f_t)
    torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0, 0.0])
    for key, e0, e1 in records:
        agg[key][0] += 1
        agg[key][1] += e0.elapsed_time(e1) * 1e3
    total = sum(v[1] for v in agg.values())
    print("dense calls %d, total %.1f us" % (len(records), total))
    for key, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        op, M, K, U = key
        print("%s M=%-8d K=%-4d U=%-4d x%-4d %9.1f us  (%.1f us each)" % (("NN", "NT", "TN")[op], M, K, U, n, t, t / n))


if __name__ == "__main__":
    main()
