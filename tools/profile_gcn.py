"""Kernel table + phase timings of the partitioned GCN step on N GPUs (rank 0 prints).  Not a bench number.
   torchrun --nproc-per-node N tools/profile_gcn.py [--exchange halo|allgather]"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--exchange", default="halo")
    ap.add_argument("--nodes", type=int, default=2000000)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.partition import PartitionedGraph, gcn_layer_partitioned, partition_bounds, _all_to_all_rows
    feat = 128
    per, bounds = partition_bounds(a.nodes, world)
    lo, hi = bounds[rank]
    recv, send, w = bench.gcn_c5_shard(a.nodes, lo, hi, seed=0)
    graph = PartitionedGraph.from_shard(a.nodes, recv - rank * per, send, w, rank, world, device, exchange=a.exchange)
    graph.warm()
    x = torch.randn(per, feat, device=device, requires_grad=True)
    kernel = (torch.randn(feat, feat, device=device) / feat ** 0.5).requires_grad_(True)
    bias = torch.zeros(feat, device=device, requires_grad=True)
    dy = torch.randn(per, feat, device=device)

    def step():
        y = gcn_layer_partitioned(graph, x, kernel, bias, "relu")
        return torch.autograd.grad(y, [x, kernel, bias], dy)

    def timed(fn, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / n

    res = {"step": timed(step)}
    if graph.halo is not None and world > 1:
        p = graph.halo
        h = x.detach()
        res["gather_rows"] = timed(lambda: F.raw_gather_rows(h, p.send_ids))
        sb = F.raw_gather_rows(h, p.send_ids)

        def a2a():
            wk, out = _all_to_all_rows(sb, p.send_counts, p.need_counts)
            wk.wait()
        res["a2a_fwd"] = timed(a2a)
        res["gms_local"] = timed(lambda: F.raw_gms(h, graph.w_loc, graph.recv_loc, graph.send_loc))
        halo = torch.randn(p.n_halo, feat, device=device)
        res["gms_remote"] = timed(lambda: F.raw_gms(halo, graph.w_rem, graph.recv_rem, graph.send_rem))
        res["gmsT_local"] = timed(lambda: F.raw_gms(dy, graph.w_loc, graph.send_loc, graph.recv_loc))
        res["gmsT_remote"] = timed(lambda: F.raw_gms(dy, graph.w_rem, graph.send_rem, graph.recv_rem))
        back = torch.randn(p.n_send, feat, device=device)
        res["segsum_back"] = timed(lambda: F.raw_gms(back, None, p.rel_back, None))
        res["counts"] = {"n_halo": p.n_halo, "n_send": p.n_send, "E_loc": graph.recv_loc.n_elements,
                         "E_rem": graph.recv_rem.n_elements}
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            step()
        torch.cuda.synchronize()
    if rank == 0:
        print(res)
        rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
        for e in rows[:14]:
            print("%9.1f us/step x%-3d %s" % (e.device_time_total / 3, e.count // 3, e.key[:100]))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
