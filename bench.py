#!/usr/bin/env python
"""bench.py -- driver-facing benchmark of the kgcnn hot path on B200 (contract: see DESIGN.md 'Measurement').

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload schnet|gss] [--impl reference]

Prints ONE JSON line on rank 0.  Workloads:
  gss     gather + segment-sum forward + backward on GSS-L (C1 pattern scaled to 65 536 graphs,
          N=1.31 M nodes, E=26.2 M edges, F=64) -> edges/s + HBM roofline
  schnet  SchNet energy+force training step, C2 (256 ethanol-shaped molecules) -> molecules/s
`--impl reference` times the oracle (CPU restatement of the reference's TF op chain) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# ----------------------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons DURING the timed region."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index=0, period=0.2):
        self.device_index, self.period = device_index, period
        self.samples, self._stop, self._thread = [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device_index), "--query-gpu=" + self.QUERY,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                parts = [p.strip() for p in out.stdout.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=10)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_tensor_peak():
    """(dense bf16 TFLOP/s burst peak, where it comes from).  Measured file with the key -> measured; no file at all ->
    the recipe's stated fallback, labelled so; a file WITHOUT the key -> (None, ...): fractions are then reported as
    null instead of resting on an unmeasured number."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if not os.path.exists(path):
        return 1590.0, "fallback (B200_PROFILING.md, 1.59 PFLOP/s)"
    try:
        with open(path) as fh:
            return float(json.load(fh)["bf16_tflops"]), "measured (MEASURED_PEAKS.json bf16_tflops, burst)"
    except Exception:
        return None, "unavailable (MEASURED_PEAKS.json has no bf16_tflops)"


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(value, world, device):
    if world == 1:
        return value
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, world, device):
    if world == 1:
        return value
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


PROFILE = False


def kernel_table(step_fn, iters=3, rows=30):
    """Kernel-time breakdown through CUPTI (diagnostics only; never a bench number)."""
    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(iters):
            step_fn()
        torch.cuda.synchronize()
    ev = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
    total = sum(e.device_time_total for e in ev)
    print("[profile] device time per step %.3f ms, %d kernels/step" % (total / iters / 1e3,
                                                                       sum(e.count for e in ev) // iters), file=sys.stderr)
    for e in ev[:rows]:
        print("[profile] %9.1f us/step %5.1f%% x%-4d %s" % (e.device_time_total / iters, 100 * e.device_time_total / total,
                                                            e.count // iters, e.key[:120]), file=sys.stderr)


def timed_region(step_fn, steps, warmup, world, device):
    """W warm-ups, then K steps bracketed by barrier + synchronize, CUDA events on the launching stream,
    max over ranks.  Returns (ms_per_step, clocks)."""
    for _ in range(warmup):
        step_fn()
    if PROFILE:
        kernel_table(step_fn)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(device.index or 0) as clocks:
        start.record()
        for _ in range(steps):
            step_fn()
        stop.record()
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
    ms = start.elapsed_time(stop)
    ms = max_over_ranks(ms, world, device)
    return ms / steps, clocks.summary()


# ----------------------------------------------------------------------------------------------------
# synthetic inputs (seeded, no dataset files)
# ----------------------------------------------------------------------------------------------------
def gss_inputs(n_graphs, n_nodes=20, deg=20, feat=64, seed=0):
    """C1 pattern scaled: every receiver gets `deg` senders from its own graph (19 distinct others in random
    order + 1 repeat, no self loops), receiver-sorted.  Returns host (numpy) arrays in SAMPLE indexing."""
    rng = np.random.default_rng(seed)
    n_total = n_graphs * n_nodes
    recv_local = np.tile(np.arange(n_nodes, dtype=np.int64), n_graphs)            # [N]
    offs = rng.permuted(np.tile(np.arange(1, n_nodes, dtype=np.int64), (n_total, 1)), axis=1)   # [N, 19]
    if deg <= n_nodes - 1:
        offs = offs[:, :deg]
    else:
        extra = rng.integers(1, n_nodes, size=(n_total, deg - (n_nodes - 1)))
        offs = np.concatenate([offs, extra], axis=1)
    send_local = (recv_local[:, None] + offs) % n_nodes                            # [N, deg]
    index = np.stack([np.repeat(recv_local, deg), send_local.reshape(-1)], axis=1)  # [E, 2] int64
    x = rng.standard_normal((n_total, feat), dtype=np.float32)
    node_splits = np.arange(n_graphs + 1, dtype=np.int64) * n_nodes
    edge_splits = np.arange(n_graphs + 1, dtype=np.int64) * (n_nodes * deg)
    return x, index, node_splits, edge_splits


def gss_bytes(E, N, F):
    """B_plain of SURVEY.md 8d: fwd (4EF + 4E + 4N + 4NF) + bwd (same op on the transposed graph)."""
    return 8 * E * F + 8 * E + 8 * N * (F + 1)


# ----------------------------------------------------------------------------------------------------
# workload: gather + segment-sum forward + backward
# ----------------------------------------------------------------------------------------------------
def run_gss(args, rank, world, device):
    from gcnn_keras_b200 import RaggedTensor, _lib
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import get_plan

    n_graphs, feat = args.gss_graphs, args.gss_feat
    x_h, idx_h, ns_h, es_h = gss_inputs(n_graphs, feat=feat, seed=rank)
    N, E = x_h.shape[0], idx_h.shape[0]
    pin = lambda a: torch.from_numpy(a).pin_memory()
    x_p, idx_p, ns_p, es_p = pin(x_h), pin(idx_h), pin(ns_h), pin(es_h)
    dy = torch.randn(N, feat, device=device)

    # ---- resident inputs: plan built once per batch, fwd + bwd through the public autograd API ----
    x_d = x_p.to(device).requires_grad_(True)
    nodes = RaggedTensor.from_row_splits(x_d, ns_p.to(device))
    index = RaggedTensor.from_row_splits(idx_p.to(device), es_p.to(device))
    plan = get_plan(index, nodes)
    recv, send = plan.relation(0), plan.relation(1)
    recv.rowptr, send.rowptr, send.permuted_ids(recv)   # build CSR + CSC before timing

    def step():
        y = F.gather_mul_segment_sum(x_d, None, recv, send)
        (gx,) = torch.autograd.grad(y, x_d, dy)
        return y, gx

    _lib.launch_count_reset()
    step()
    launches_per_step = _lib.launch_count()
    ms_step, clocks = timed_region(step, args.steps, args.warmup, world, device)

    # ---- end to end: host buffers -> H2D -> plan -> fwd -> bwd -> D2H of the results ----
    y_host = torch.empty((N, feat), dtype=torch.float32).pin_memory()
    gx_host = torch.empty((N, feat), dtype=torch.float32).pin_memory()

    def e2e_step():
        xd = x_p.to(device, non_blocking=True).requires_grad_(True)
        nd = RaggedTensor.from_row_splits(xd, ns_p.to(device, non_blocking=True))
        ix = RaggedTensor.from_row_splits(idx_p.to(device, non_blocking=True), es_p.to(device, non_blocking=True))
        pl = get_plan(ix, nd)
        y = F.gather_mul_segment_sum(xd, None, pl.relation(0), pl.relation(1))
        (gx,) = torch.autograd.grad(y, xd, dy)
        y_host.copy_(y.detach(), non_blocking=True)
        gx_host.copy_(gx, non_blocking=True)

    e2e_steps = max(2, min(args.steps, 10))
    ms_e2e, _ = timed_region(e2e_step, e2e_steps, 2, world, device)

    total_edges = sum_over_ranks(float(E), world, device)
    peak, peak_src = measured_peaks()
    alg_bytes = gss_bytes(E, N, feat)
    n_kernel_launches = 2   # fwd + bwd are the same kernel on the receiver CSR / sender CSC
    achieved = (alg_bytes / n_kernel_launches) / (ms_step / n_kernel_launches * 1e-3) / 1e9
    h2d = x_p.numel() * 4 + idx_p.numel() * 8 + ns_p.numel() * 8 + es_p.numel() * 8
    d2h = 2 * N * feat * 4
    line = {
        "metric": "edges/sec (gather+segment-sum fwd+bwd)", "value": total_edges / (ms_step * 1e-3),
        "unit": "edges/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "GSS-L: gather+segment-sum fwd+bwd, C1 pattern x %d graphs" % n_graphs,
                   "graphs_per_gpu": n_graphs, "nodes_per_gpu": N, "edges_per_gpu": E, "features": feat,
                   "l2_policy": "inputs larger than L2 (%.0f MB working set)" % ((2 * N * feat * 4 + 2 * E * 4) / 1e6),
                   "parallelism": "graphs sharded across %d rank(s), no data-path collective" % world},
        "hbm_gbs_algorithmic": alg_bytes / (ms_step * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": profile_traffic(GMS_PROFILE, "gms_kernel", 6, 2),
                     "kernel": "gms_kernel<float4,8,2,SumW0,gather>",
                     "algorithmic_bytes_per_launch": alg_bytes / n_kernel_launches, "peak_source": peak_src,
                     "note": "algorithmic bytes count every gathered row (no cache credit, SURVEY 8d); E/N = 20, so "
                             "19 of 20 row reads hit L1/L2: frac > 1 is cache reuse (on-chip gather throughput), not an "
                             "HBM number. The HBM-bound shapes are in the default bench line's roofline block"},
        "e2e": {"value": total_edges / (ms_e2e * 1e-3), "unit": "edges/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
    }
    return line


# ----------------------------------------------------------------------------------------------------
# CPU baseline: the oracle (op-for-op restatement of the reference's TF chain) on the host cores
# ----------------------------------------------------------------------------------------------------
def cpu_gss(n_graphs, feat, steps, warmup, threads):
    """gather -> PoolingLocalEdges(sum) forward + backward with the reference's op chain
    (partition_row_indexing in BOTH layers, stable argsort, permutation gathers, sorted segment_sum,
    zero padding), torch CPU autograd, `threads` host threads.  Returns (edges_per_s, ms_per_step, E)."""
    from oracle import R
    from oracle import layers as OL
    torch.set_num_threads(threads)
    x_h, idx_h, ns_h, es_h = gss_inputs(n_graphs, feat=feat, seed=0)
    x = torch.from_numpy(x_h).requires_grad_(True)
    nodes = R(x, ns_h)
    index = R(torch.from_numpy(idx_h), es_h)
    dy = torch.randn(x.shape)

    def step():
        msg = OL.gather_nodes_outgoing(nodes, index)
        y = OL.pooling_local_edges(nodes, msg, index, pooling_method="sum").values
        (gx,) = torch.autograd.grad(y, x, dy)
        return gx

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return idx_h.shape[0] / dt, dt * 1e3, idx_h.shape[0]


def cpu_baseline_gss(args):
    cores = os.cpu_count() or 1
    n_graphs = 8192    # bounded sample: 1/8 of GSS-L, ~10-30 s of CPU work
    eps, ms, E = cpu_gss(n_graphs, args.gss_feat, steps=3, warmup=1, threads=cores)
    return {"value": eps, "unit": "edges/s", "cores": cores, "kind": "port",
            "sample": "same op on %d graphs (E=%d, F=%d), oracle torch-CPU fwd+bwd, 3 steps, %.0f ms/step"
                      % (n_graphs, E, args.gss_feat, ms)}


def run_reference_gss(args):
    cores = os.cpu_count() or 1
    n_graphs = 8192
    eps, ms, E = cpu_gss(n_graphs, args.gss_feat, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 2)),
                         threads=cores)
    return {"impl": "reference", "metric": "edges/sec (gather+segment-sum fwd+bwd)", "value": eps,
            "unit": "edges/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "GSS-L: gather+segment-sum fwd+bwd, C1 pattern (CPU sample: %d graphs)" % n_graphs,
                       "features": args.gss_feat, "edges_per_step": E},
            "cpu_baseline": {"value": eps, "unit": "edges/s", "cores": cores, "kind": "port",
                             "sample": "oracle torch-CPU restatement of the reference TF op chain, %d graphs" % n_graphs},
            "e2e": {"value": eps, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "TensorFlow is not installable here; the reference arm is the oracle port on the host cores"}


# ----------------------------------------------------------------------------------------------------
# workload: SchNet energy + force training step (BASELINE.json configs[1], C2)
# ----------------------------------------------------------------------------------------------------
def build_schnet(device, batch, seed):
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    from gcnn_keras_b200.training import ForceTrainStep
    torch.manual_seed(1234)                      # identical initial weights on every rank (DP replicas)
    model = EnergyForceModel(model_energy=make_model(**SCHNET_C2_CONFIG), energy_output=0, output_to_tensor=True,
                             output_as_dict=False, output_squeeze_states=True, is_physical_force=False)
    return model, ForceTrainStep(model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0), schnet_c2_batch(batch, seed)


def timed_steps_with_l2_flush(step_fn, steps, warmup, world, device):
    """Per-step CUDA events with a 256 MB L2 flush between steps (outside the events); max over ranks."""
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=device)
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    pairs = []
    with ClockSampler(device.index or 0) as clocks:
        for _ in range(steps):
            flush.fill_(0.0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step_fn()
            b.record()
            pairs.append((a, b))
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in pairs)
    return max_over_ranks(ms, world, device) / steps, clocks.summary()


def _time(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def profile_traffic(csv_name, kernel_substr, skip=0, count=None):
    """DRAM bytes per launch (read + write, mean over the matching rows) of a COMMITTED ncu summary under profiles/
    (written by profiles/summarize.py from an `ncu --set full` capture of tools/ncu_gms.py).  None if absent."""
    import csv
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        return None
    with open(path) as fh:
        rows = list(csv.reader(fh))
    hdr = rows[0]
    try:
        i_r = next(i for i, h in enumerate(hdr) if h.startswith("dram__bytes_read.sum"))
        i_w = next(i for i, h in enumerate(hdr) if h.startswith("dram__bytes_write.sum"))
    except StopIteration:
        return None
    scale = lambda h: {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(h.split("[")[-1].rstrip("]"), 1.0)
    sel = [r for r in rows[1:] if kernel_substr in r[0]][skip:]
    sel = sel[:count] if count else sel
    if not sel:
        return None
    return sum(float(r[i_r]) * scale(hdr[i_r]) + float(r[i_w]) * scale(hdr[i_w]) for r in sel) / len(sel)


def profile_traffic_named(csv_name, kernel_substr, grids, count):
    """As profile_traffic, but the rows are chosen by kernel-name substring AND launch__grid_size; None unless exactly
    `count` rows match (a summary regenerated from a different build or shape list then yields null, not a wrong row)."""
    import csv
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        return None
    with open(path) as fh:
        rows = list(csv.reader(fh))
    hdr = rows[0]
    try:
        i_g = next(i for i, h in enumerate(hdr) if h.startswith("launch__grid_size"))
        i_r = next(i for i, h in enumerate(hdr) if h.startswith("dram__bytes_read.sum"))
        i_w = next(i for i, h in enumerate(hdr) if h.startswith("dram__bytes_write.sum"))
    except StopIteration:
        return None
    scale = lambda h: {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(h.split("[")[-1].rstrip("]"), 1.0)
    sel = [r for r in rows[1:] if kernel_substr in r[0] and int(float(r[i_g])) in grids]
    if len(sel) != count:
        return None
    return sum(float(r[i_r]) * scale(hdr[i_r]) + float(r[i_w]) * scale(hdr[i_w]) for r in sel) / len(sel)


def roofline_shapes(device):
    """Generator over the measured shapes of the north-star kernel family: (key, workload text, kernel text,
    algorithmic bytes, edges, [launch callables]).  Used by gss_roofline (timing) and tools/ncu_gms.py (one launch of
    each callable, in this order, under ncu -> profiles/r2_gms_full.csv)."""
    from gcnn_keras_b200 import RaggedTensor
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import Relation, get_plan
    rng = np.random.default_rng(0)
    # ---- plain, GSS-R ----
    n, deg, feat = 2_000_000, 16, 128
    recv_r = torch.from_numpy(np.repeat(np.arange(n, dtype=np.int32), deg)).to(device)
    send_r = torch.from_numpy(rng.integers(0, n, size=n * deg, dtype=np.int32)).to(device)
    r_recv, r_send = Relation(recv_r, n, is_sorted=True), Relation(send_r, n)
    r_recv.rowptr, r_send.rowptr, r_send.permuted_ids(r_recv)
    x = torch.randn(n, feat, device=device)
    dy = torch.randn(n, feat, device=device)
    yield ("plain", "GSS-R (N=%d, E=%d uniform random senders, F=%d): gathered rows miss L2" % (n, n * deg, feat),
           "gms_kernel (gather -> CSR segment-sum), forward + backward launch", gss_bytes(n * deg, n, feat), n * deg,
           [lambda: F.raw_gms(x, None, r_recv, r_send), lambda: F.raw_gms(dy, None, r_send, r_recv)])
    del x, dy, r_recv, r_send, recv_r, send_r
    # ---- weighted (cfconv) ----
    n, deg, feat = 1_000_000, 16, 128
    e_w = n * deg
    recv_w = torch.from_numpy(np.repeat(np.arange(n, dtype=np.int32), deg)).to(device)
    send_w = torch.from_numpy(rng.integers(0, n, size=e_w, dtype=np.int32)).to(device)
    w_recv, w_send = Relation(recv_w, n, is_sorted=True), Relation(send_w, n)
    w_recv.rowptr, w_send.rowptr, w_send.permuted_ids(w_recv)
    x = torch.randn(n, feat, device=device)
    dy = torch.randn(n, feat, device=device)
    w = torch.randn(e_w, feat, device=device)
    yield ("weighted", "cfconv variant (N=%d, E=%d, F=%d, w [E,F]): y = sum w*x[s], dx, dw" % (n, e_w, feat),
           "gms_kernel (w [E,F]) forward + transposed, gather2_mul_kernel (dw)",
           20 * e_w * feat + 12 * n * feat + 12 * e_w + 8 * n, e_w,
           [lambda: F.raw_gms(x, w, w_recv, w_send), lambda: F.raw_gms(dy, w, w_send, w_recv),
            lambda: F.raw_g2m(x, w_send.ids, dy, w_recv.ids, e_w, False)])
    del x, dy, w, w_recv, w_send, recv_w, send_w
    # ---- GSS-G: one C5 shard ----
    n_src, n_dst, e_g, feat = 2_000_000, 250_000, 7_500_000, 128
    deg_g = e_g // n_dst
    recv_g = torch.from_numpy(np.repeat(np.arange(n_dst, dtype=np.int32), deg_g)).to(device)
    base = np.repeat(np.arange(n_dst, dtype=np.int64), deg_g)
    local = rng.random(e_g) < 0.8
    send_np = np.where(local, np.mod(base + rng.integers(-50000, 50001, size=e_g), n_src),
                       rng.integers(0, n_src, size=e_g))
    send_g = torch.from_numpy(send_np.astype(np.int32)).to(device)
    g_recv, g_send = Relation(recv_g, n_dst, is_sorted=True), Relation(send_g, n_src)
    g_recv.rowptr, g_send.rowptr, g_send.permuted_ids(g_recv)
    x = torch.randn(n_src, feat, device=device)
    dy = torch.randn(n_dst, feat, device=device)
    w1 = torch.rand(e_g, 1, device=device)
    yield ("gss_g", "GSS-G, one C5 shard (N_src=%d, N_dst=%d, E=%d, F=%d, w [E,1])" % (n_src, n_dst, e_g, feat),
           "gms_kernel (w [E,1]) forward on the receiver CSR + backward on the sender CSC",
           8 * e_g * feat + 4 * (n_src + n_dst) * feat + 16 * e_g + 4 * (n_src + n_dst), e_g,
           [lambda: F.raw_gms(x, w1, g_recv, g_send), lambda: F.raw_gms(dy, w1, g_send, g_recv)])
    del x, dy, w1, g_recv, g_send, recv_g, send_g
    # ---- GSS-L: served by L1/L2 ----
    x_h, idx_h, ns_h, es_h = gss_inputs(65536, feat=64, seed=0)
    N, E = x_h.shape[0], idx_h.shape[0]
    x = torch.from_numpy(x_h).to(device)
    nodes = RaggedTensor.from_row_splits(x, torch.from_numpy(ns_h).to(device))
    index = RaggedTensor.from_row_splits(torch.from_numpy(idx_h).to(device), torch.from_numpy(es_h).to(device))
    plan = get_plan(index, nodes)
    recv, send = plan.relation(0), plan.relation(1)
    recv.rowptr, send.rowptr, send.permuted_ids(recv)
    dyl = torch.randn(N, 64, device=device)
    yield ("cached_shape", "GSS-L (N=%d, E=%d, F=64, C1 pattern x 65536 graphs)" % (N, E),
           "gms_kernel, forward + backward launch", gss_bytes(E, N, 64), E,
           [lambda: F.raw_gms(x, None, recv, send), lambda: F.raw_gms(dyl, None, send, recv)])


CFCONV_PROFILE = "r2_cfconv_full.csv"   # ncu --set full of tools/ncu_cfconv.py (mode 0 and mode 1 launch, same shape)


def cfconv_roofline(device, iters=5):
    """The fused SchNet cfconv (csrc/cfconv.cu: basis -> Dense(ssp) -> Dense -> gather * filter -> segment sum in ONE
    kernel, kgcnn/layers/conv/schnet_conv.py:73-79) at an HBM-sized graph, forward (mode 0) and input-gradient (mode 1)
    launch, beside the unfused chain on the same inputs.  The kernel is TENSOR-bound, not HBM-bound: its DRAM traffic
    is the node rows and the indices only (ncu: profiles/r2_cfconv_full.csv).  `achieved` counts the ALGORITHMIC
    flops of the two Dense layers (2 (R + 128) 128 per edge and derivative order); `issued` the bf16 MMA flops the
    three-piece split really executes (6 MMAs per product, R padded to 32)."""
    from gcnn_keras_b200 import RaggedTensor
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.plan import get_plan
    rng = np.random.default_rng(0)
    n, deg, R = 262144, 32, 25
    recv = np.repeat(np.arange(n), deg)
    send = (recv + rng.integers(1, 4096, size=recv.shape)) % n
    E = recv.shape[0]
    nodes = RaggedTensor.from_row_lengths(torch.randn(n, 128, device=device), torch.tensor([n], device=device))
    index = RaggedTensor.from_row_lengths(torch.tensor(np.stack([recv, send], 1), device=device),
                                          torch.tensor([E], device=device))
    plan = get_plan(index, nodes)
    spec = F.CfconvSpec(plan.relation(0), plan.relation(1), R, 3.125, 0.0, 0.2, F.act_code("kgcnn>shifted_softplus"))
    x, d = nodes.values, torch.rand(E, device=device) * 4 + 0.8
    w1, b1 = torch.randn(R, 128, device=device) * 0.4, torch.zeros(128, device=device)
    w2, b2 = torch.randn(128, 128, device=device) * 0.15, torch.zeros(128, device=device)
    gy = torch.randn(n, 128, device=device)
    spec.recv.rowptr, spec.send.rowptr, spec.send.permuted_ids(spec.recv), spec.send.permuted_ids(spec.send)
    fwd = lambda: F.raw_cfconv(0, spec, d, w1, b1, w2, b2, spec.recv, spec.send, x)
    vjp = lambda: F.raw_cfconv(1, spec, d, w1, b1, w2, b2, spec.send, spec.recv, gy, b=x)
    with torch.no_grad():
        ms_f, ms_b = _time(fwd, iters), _time(vjp, iters)
        ms_unf = _time(lambda: F._cfconv_unfused(x, d, w1, b1, w2, b2, spec), iters)
        # the training step's reverse sweep also asks for the filter network's weight-gradient operands (want_wg):
        # 4 planes [E, R | 128] of fp32 written to HBM in mode 1 (8 in mode 3) for the grouped TN launch
        ms_b_wg = _time(lambda: F.raw_cfconv(1, spec, d, w1, b1, w2, b2, spec.send, spec.recv, gy, b=x, want_wg=True),
                        iters)
    tensor_peak, tensor_src = measured_tensor_peak()
    frac_of = lambda tflops: None if tensor_peak is None else tflops / tensor_peak
    wg_bytes_mode1 = 4 * E * (R + 3 * 128)
    alg = 2.0 * (R + 128) * 128 * E                      # forward; the VJP evaluates phi and phi': twice that
    issued = 60 * 2.0 * 128 * 128 * 16 * ((E + 127) // 128)
    t_f = profile_traffic(CFCONV_PROFILE, "cfconv_kernel<0", 0, 1)
    t_b = profile_traffic(CFCONV_PROFILE, "cfconv_kernel<1", 0, 1)
    alg_bytes_f = 8 * n * 128 + 12 * E                    # node rows in, node rows out, (row, other, distance) per edge
    hbm_peak, _ = measured_peaks()
    return {"bound": "tensor", "workload": "fused cfconv, N=%d, E=%d, F=128, %d Gaussians" % (n, E, R),
            "kernel": "cf::cfconv_kernel<0> (value) and <1> (input gradients), csrc/cfconv.cu",
            "unit": "TFLOP/s", "peak": tensor_peak, "peak_source": tensor_src,
            "achieved": alg / (ms_f * 1e-3) / 1e12, "frac": frac_of(alg / (ms_f * 1e-3) / 1e12),
            "issued_TFLOPs": issued / (ms_f * 1e-3) / 1e12, "frac_issued": frac_of(issued / (ms_f * 1e-3) / 1e12),
            "ms_forward": ms_f, "ms_input_gradients": ms_b, "edges_per_s_forward": E / (ms_f * 1e-3),
            "achieved_input_gradients": 2 * alg / (ms_b * 1e-3) / 1e12,
            "frac_issued_input_gradients": frac_of(2 * issued / (ms_b * 1e-3) / 1e12),
            "weight_gradient_sweep": {
                "ms_input_gradients_with_operands": ms_b_wg,
                "operand_plane_bytes_mode1": wg_bytes_mode1, "operand_plane_bytes_mode3": 2 * wg_bytes_mode1,
                "operand_write_GBps": wg_bytes_mode1 / (ms_b_wg * 1e-3) / 1e9,
                "note": "in the TRAINING step the per-edge tensors do reach HBM once: modes 1 and 3 with want_wg write "
                        "4 / 8 fp32 planes [E, R | 128] for kgcnn_dense_tn_grouped; the 0.35 GB traffic figure and "
                        "'the filter never reaches HBM' cover the forward and the input-gradient launches only"},
            "unfused_forward_ms": ms_unf, "speedup_over_unfused_forward": ms_unf / ms_f,
            "algorithmic_bytes_forward": alg_bytes_f, "traffic": t_f, "traffic_input_gradients": t_b,
            "traffic_over_algorithmic_bytes": None if t_f is None else t_f / alg_bytes_f,
            "unfused_algorithmic_bytes_forward": 4 * E * (2 * R + 5 * 128) + 8 * n * 128 + 8 * E,
            "hbm_floor_ms_of_unfused_forward": (4 * E * (2 * R + 5 * 128) + 8 * n * 128 + 8 * E) / hbm_peak / 1e6,
            "traffic_source": "profiles/" + CFCONV_PROFILE,
            "note": "the fused kernel moves 0.35 GB where the unfused chain moves >= 23 GB (rbf, two [E,128] "
                    "activations written and re-read, the filter read by the pooling kernel); what bounds it is the "
                    "serial per-tile chain basis -> MMA -> activation -> MMA -> epilogue (ncu: tensor pipe 25-31 % "
                    "active, issue slots 32-34 %)"}


GMS_PROFILE = "r2_gms_full.csv"     # ncu --set full of tools/ncu_gms.py on the final build (rows in launch order)
# rows of the ncu summary by kernel instantiation AND grid size (not by position): (name substring, grid sizes, launches)
_PROFILE_ROWS = {"plain": [("gms_kernel<float4, 8, 4, 0,", (62500,), 2)],
                 "weighted": [("gms_kernel<float4, 8, 4, 2,", (31250,), 2), ("gather2_mul_kernel<float4, 32, 0>", (500000,), 1)],
                 "gss_g": [("gms_kernel<float4, 8, 4, 1,", (7813, 62500), 2)],
                 "cached_shape": [("gms_kernel<float4, 8, 2, 0,", (40960,), 2)]}


def gss_roofline(device, iters=10):
    """The north-star kernel family (fused gather [* w] -> CSR segment-sum and its two VJPs) timed alone against the
    measured HBM copy peak, on the three HBM-bound shapes of SURVEY.md 8(d) (plain GSS-R; the cfconv variant with
    w [E, F]: forward, dx, dw = B_cfconv; GSS-G = one C5 shard, B_gcn) and on GSS-L, whose gathered rows are served by
    L1/L2 and which is therefore reported as on-chip throughput, not as an HBM fraction.  `traffic` = DRAM bytes of
    the same launches from the committed ncu capture (profiles/r2_gms_full.csv), summed over the launches."""
    peak, peak_src = measured_peaks()
    out = {}
    for key, workload, kernel, alg_bytes, edges, fns in roofline_shapes(device):
        ms = _time(lambda: [f() for f in fns], iters if key != "weighted" else max(3, iters // 2))
        parts = [profile_traffic_named(GMS_PROFILE, k, grids, cnt) for k, grids, cnt in _PROFILE_ROWS[key]]
        traffic = None if any(p is None for p in parts) else sum(
            p * cnt for p, (_, _, cnt) in zip(parts, _PROFILE_ROWS[key]))
        ach = alg_bytes / (ms * 1e-3) / 1e9
        if key == "cached_shape":
            out[key] = {"bound": "l2", "workload": workload, "kernel": kernel, "on_chip_throughput_GBps": ach,
                        "ms": ms, "launches": len(fns), "edges_per_s": edges / (ms * 1e-3),
                        "algorithmic_bytes": alg_bytes, "traffic": traffic,
                        "note": "E/N = 20 and the 20 senders of a receiver sit in its own graph: 19 of 20 row reads "
                                "hit L1/L2, so this shape measures on-chip gather throughput, not an HBM fraction"}
            continue
        e = {"bound": "hbm", "workload": workload, "kernel": kernel, "achieved": ach, "peak": peak, "unit": "GB/s",
             "frac": ach / peak, "algorithmic_bytes": alg_bytes, "ms": ms, "launches": len(fns),
             "edges_per_s": edges / (ms * 1e-3), "traffic": traffic,
             "frac_on_traffic": None if traffic is None else traffic / (ms * 1e-3) / 1e9 / peak,
             "traffic_source": ("profiles/" + GMS_PROFILE) if traffic is not None else None, "peak_source": peak_src,
             "algorithmic_bytes_per_launch": alg_bytes / len(fns), "ms_per_launch": ms / len(fns),
             "traffic_per_launch": None if traffic is None else traffic / len(fns)}
        if key == "plain":
            out.update(e)
        else:
            out[key] = e
    out["cfconv_fused"] = cfconv_roofline(device, max(3, iters // 2))
    out["kernel_family"] = "gms_kernel / gather2_mul_kernel (csrc/segment_impl.cuh, csrc/gather.cu)"
    out["note"] = ("achieved = algorithmic bytes of SURVEY 8(d) (every gathered row counted, no cache credit) / time of "
                   "the launches; traffic = measured DRAM bytes of the same launches (ncu), so frac_on_traffic is the "
                   "fraction of the HBM peak the kernels actually drive")
    return out


def flatten_roofline(r):
    """Scalar copies of the nested entries (weighted, gss_g, cached_shape, cfconv_fused) beside the flat keys of the
    plain kernel, so that a reader that keeps only the top level of `roofline` still sees every fraction."""
    get = lambda key, field: (r.get(key) or {}).get(field)
    r.update({
        "weighted_frac": get("weighted", "frac"), "weighted_frac_on_traffic": get("weighted", "frac_on_traffic"),
        "weighted_ms": get("weighted", "ms"),
        "gss_g_frac": get("gss_g", "frac"), "gss_g_frac_on_traffic": get("gss_g", "frac_on_traffic"),
        "gss_g_ms": get("gss_g", "ms"),
        "cached_shape_on_chip_GBps": get("cached_shape", "on_chip_throughput_GBps"),
        "cfconv_tensor_frac": get("cfconv_fused", "frac"), "cfconv_tensor_frac_issued": get("cfconv_fused", "frac_issued"),
        "cfconv_ms_forward": get("cfconv_fused", "ms_forward"),
        "cfconv_ms_input_gradients": get("cfconv_fused", "ms_input_gradients"),
        "cfconv_traffic_over_algorithmic_bytes": get("cfconv_fused", "traffic_over_algorithmic_bytes"),
        "cfconv_tensor_peak_source": get("cfconv_fused", "peak_source")})
    return r


def build_painn(device, batch, seed):
    from gcnn_keras_b200.literature.PAiNN import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import PAINN_C3_CONFIG, painn_c3_batch
    from gcnn_keras_b200.training import ForceTrainStep
    torch.manual_seed(1234)
    model = EnergyForceModel(model_energy=make_model(**PAINN_C3_CONFIG), energy_output=0, output_to_tensor=True,
                             output_as_dict=False, output_squeeze_states=True, is_physical_force=False)
    return model, ForceTrainStep(model, lr=1e-3, force_loss_factor=95.0, clipnorm=1.0), painn_c3_batch(batch, seed)


def build_hdnnp(device, batch, seed):
    from gcnn_keras_b200.literature.HDNNP2nd import make_model_behler
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import HDNNP_C4_CONFIG, hdnnp_c4_batch
    from gcnn_keras_b200.training import ForceTrainStep
    torch.manual_seed(1234)
    model = EnergyForceModel(model_energy=make_model_behler(**HDNNP_C4_CONFIG), energy_output=0,
                             output_to_tensor=True, output_as_dict=False, output_squeeze_states=True,
                             is_physical_force=False)
    return model, ForceTrainStep(model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0), hdnnp_c4_batch(batch, seed)


FORCE_WORKLOADS = {
    "schnet": (build_schnet, 256, "SchNet force-train molecules/sec",
               "C2: SchNet energy+force training step (force_schnet.py), synthetic MD17-ethanol-shaped data: 9 atoms, "
               "cutoff 5 A (72 edges)"),
    "painn": (build_painn, 512, "PaiNN force-train molecules/sec",
              "C3: PaiNN energy+force training step (force_painn.py), synthetic 30-atom molecules, all pairs < 5 A"),
    "hdnnp": (build_hdnnp, 256, "HDNNP2nd force-train molecules/sec",
              "C4: HDNNP2nd ACSF G2/G4 energy+force training step (force_hdnnp2nd.py), synthetic 64-atom systems, "
              "25 neighbours, 'kj' angle triples"),
}


def run_schnet(args, rank, world, device):
    return run_force(args, rank, world, device, "schnet")


def force_config(workload, batch, data, n_params, world):
    """The `config` block of a force-training line: ONE function for the B200 arm and the reference arm, so that both
    lines carry the identical dict for the same workload."""
    return {"workload": workload + ", batch %d per GPU" % batch,
            "batch_per_gpu": batch, "atoms": int(data["node_len"].sum()), "edges": int(data["edge_len"].sum()),
            "triples": int(data["angle_len"].sum()) if "angle_len" in data else 0,
            "parameters": int(n_params), "optimizer": "Adam(1e-3), clipnorm 1.0",
            "step": "fwd + dE/dx + loss + d loss/d weights through the force + Adam (B200 arm: one CUDA graph)",
            "l2_policy": "256 MB L2 flush between timed steps (per-step CUDA events) on the B200 arm; "
                         "the CPU reference arm has no device cache to flush",
            "parallelism": "graphs sharded across %d rank(s); weight-gradient all-reduce only "
                           "(the CPU reference arm runs on rank 0 alone)" % world}


def run_force(args, rank, world, device, kind):
    from gcnn_keras_b200.training import CapturedForceTrainStep
    build, default_batch, metric, workload = FORCE_WORKLOADS[kind]
    batch = args.batch or default_batch
    model, train_step, data = build(device, batch, seed=rank)
    # C4: the 'kj' angle triples (278 MB per batch) are rebuilt on the device from the pair list in the e2e leg
    device_angles = "kj" if (kind == "hdnnp" and not args.host_angles) else None
    captured = CapturedForceTrainStep(train_step, data, device, device_angles=device_angles)
    if world > 1:       # replicas start from rank 0's weights (the seed already makes them equal; this makes it a fact)
        from gcnn_keras_b200.distributed import broadcast_parameters
        broadcast_parameters(model)
    n_params = sum(p.numel() for p in model.parameters())

    ms_step, clocks = timed_steps_with_l2_flush(lambda: captured.run(load=False), args.steps, args.warmup, world,
                                                device)
    assert captured.check(), "plan assumptions of the captured graph were violated"
    ms_e2e, _ = timed_steps_with_l2_flush(lambda: captured.run(load=True), args.steps, args.warmup, world, device)
    loss_value = float(captured.loss_host)
    total_mol = sum_over_ranks(float(batch), world, device)
    line = {
        "metric": metric, "value": total_mol / (ms_step * 1e-3), "unit": "molecules/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": force_config(workload, batch, data, n_params, world),
        "loss": loss_value,
        "e2e": {"value": total_mol / (ms_e2e * 1e-3), "unit": "molecules/s",
                "h2d_bytes_per_step": captured.static.h2d_bytes, "d2h_bytes_per_step": 4 + (
                    0 if captured.flags is None else captured.flags.numel() * 4), "ms_per_step": ms_e2e,
                **({"angle_triples": "built on the device from the pair list (kgcnn_angle_count/emit), not copied"}
                   if device_angles else {})},
        "gpu_launches": captured.kernels_per_step * args.steps,
        "our_kernels_per_step": captured.kernels_per_step,
        "clocks": clocks,
    }
    opt = getattr(train_step, "optimizer", None)
    line["allreduce"] = "none (1 GPU)" if world == 1 else (
        "peer (kgcnn_allreduce_peer over NVLink peer memory)" if getattr(opt, "peer", None) is not None
        else "nccl (ncclAllReduce)")
    if rank == 0 and not args.no_roofline:
        line["roofline"] = flatten_roofline(gss_roofline(device))
    return line


def other_workloads(args, rank, world, device):
    """The other BASELINE configurations at this GPU count, a few steps each, so that the driver's BENCH / SCALE
    records carry them too: C3 PaiNN and C4 HDNNP2nd training steps (molecules/s) and the C5 GCN layer (ms/step)."""
    import argparse as _ap
    import gc
    out = {}
    for kind in ("painn", "hdnnp", "gcn"):
        sub = _ap.Namespace(**vars(args))
        sub.steps, sub.warmup, sub.batch, sub.no_roofline, sub.workload = 3, 3, 0, True, kind
        try:
            res = run_gcn(sub, rank, world, device) if kind == "gcn" else run_force(sub, rank, world, device, kind)
            keep = {k: res[k] for k in ("metric", "value", "unit", "ms_per_step", "scaling", "steps", "warmup") if k in res}
            keep["workload"] = res["config"]["workload"]
            keep["e2e"] = {k: res["e2e"][k] for k in ("value", "unit", "ms_per_step") if k in res["e2e"]}
            if kind == "gcn":
                keep["roofline_frac_algorithmic"] = res["roofline"]["frac"]
            out[kind] = keep
        except Exception as exc:                                    # a failure here must not lose the headline line
            out[kind] = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:200])}
        gc.collect()
        torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------------------------------
# workload: GCN layer on one large graph (BASELINE.json configs[4], C5), receiver-range partition
# ----------------------------------------------------------------------------------------------------
def gcn_c5_shard(n_nodes, lo, hi, mean_deg=30, window=50000, seed=0):
    """In-edges of receivers [lo, hi): Poisson(30) senders each, 80 % from a +-50 000 index window, 20 % uniform;
    weights 1/sqrt(deg_i deg_j) with the expected degree of the sender.  Receiver-sorted."""
    rng = np.random.default_rng(seed * 1000003 + lo)
    deg = rng.poisson(mean_deg, size=hi - lo).astype(np.int64)
    recv = np.repeat(np.arange(lo, hi, dtype=np.int64), deg)
    e = int(deg.sum())
    local = rng.random(e) < 0.8
    near = recv + rng.integers(-window, window + 1, size=e)
    far = rng.integers(0, n_nodes, size=e)
    send = np.where(local, np.mod(near, n_nodes), far)      # the index window wraps around (no boundary hubs)
    w = (1.0 / np.sqrt(np.maximum(np.repeat(deg, deg), 1) * float(mean_deg))).astype(np.float32)
    return recv, send, w


def gcn_bytes(E, N, F):
    """B_gcn of SURVEY.md 8d (forward + backward, edge weights [E,1] without gradient)."""
    return 8 * E * F + 8 * N * F + 16 * E + 8 * N


def run_gcn(args, rank, world, device):
    from gcnn_keras_b200 import _lib
    from gcnn_keras_b200 import functional as F
    from gcnn_keras_b200.partition import PartitionedGraph, gcn_layer_partitioned, partition_bounds
    n_nodes, feat = args.gcn_nodes, args.gcn_feat
    per, bounds = partition_bounds(n_nodes, world)
    lo, hi = bounds[rank]
    recv, send, w = gcn_c5_shard(n_nodes, lo, hi, seed=0)
    # every rank generates only its own in-edges (receiver ids local to its block, sender ids global); the halo plan
    # (which rows cross the cut) is host-side preprocessing done once per graph, like the reference's adjacency set-up
    graph = PartitionedGraph.from_shard(n_nodes, recv - rank * per, send, w, rank, world, device,
                                        exchange=args.gcn_exchange)
    from gcnn_keras_b200.plan import Relation
    x_p = torch.randn(per, feat).pin_memory()
    E = len(recv)
    # e2e: the index columns, edge weights and features of the step come from pinned host memory
    if graph.halo is None:
        names = [("rel_recv", per, True), ("rel_send", per * world, False)]
        w_names = ["weights"]
    else:
        names = [("recv_loc", per, True), ("send_loc", per, False), ("recv_rem", per, True),
                 ("send_rem", max(graph.halo.n_halo, 1), False)]
        w_names = ["w_loc", "w_rem"]
    pinned_ids = {n: getattr(graph, n).ids.cpu().pin_memory() for n, _, _ in names}
    pinned_w = {n: getattr(graph, n).cpu().pin_memory() for n in w_names}
    h2d_bytes = sum(t.numel() * t.element_size() for t in list(pinned_ids.values()) + list(pinned_w.values())) \
        + x_p.numel() * 4

    def load():
        for n, targets, is_sorted in names:
            setattr(graph, n, Relation(pinned_ids[n].to(device, non_blocking=True), targets, is_sorted=is_sorted))
        for n in w_names:
            setattr(graph, n, pinned_w[n].to(device, non_blocking=True))
        return x_p.to(device, non_blocking=True).requires_grad_(True)

    torch.manual_seed(0)
    kernel = (torch.randn(feat, feat, device=device) / feat ** 0.5).requires_grad_(True)
    bias = torch.zeros(feat, device=device, requires_grad=True)
    x_d = load()
    graph.warm()
    dy = torch.randn(per, feat, device=device)

    def step(x=None):
        x = x_d if x is None else x
        y = gcn_layer_partitioned(graph, x, kernel, bias, "relu")
        return torch.autograd.grad(y, [x, kernel, bias], dy)

    _lib.launch_count_reset()
    step()
    launches = _lib.launch_count()
    ms_step, clocks = timed_region(step, args.steps, args.warmup, world, device)
    g_host = torch.empty(per, feat).pin_memory()

    def e2e_step():
        x = load()
        gx, _, _ = step(x)
        g_host.copy_(gx, non_blocking=True)

    ms_e2e, _ = timed_region(e2e_step, max(2, min(args.steps, 5)), 2, world, device)
    total_edges = sum_over_ranks(float(E), world, device)
    peak, peak_src = measured_peaks()
    alg = gcn_bytes(E, per, feat)
    achieved = alg / (ms_step * 1e-3) / 1e9
    return {
        "metric": "edges/sec (GCN layer fwd+bwd, one large graph)", "value": total_edges / (ms_step * 1e-3),
        "unit": "edges/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C5: GCN layer (Dense -> gather*a_e -> segment-sum -> relu) fwd+bwd on one graph, "
                               "%d nodes, Poisson(30) in-edges, F=%d, receiver-range partition over %d GPU(s)"
                               % (n_nodes, feat, world),
                   "nodes": n_nodes, "edges_this_rank": E, "features": feat,
                   "l2_policy": "inputs larger than L2 (node features %.0f MB)" % (n_nodes * feat * 4 / 1e6),
                   "parallelism": ("halo exchange: only the rows referenced across the cut travel (row gather -> NCCL "
                                   "all-to-all, %d of %d remote rows per rank), overlapped with the local edges; "
                                   "transpose on the way back; weights replicated"
                                   % (graph.halo.n_halo, per * (world - 1))) if graph.halo is not None else
                                  "all-gather of the whole feature matrix fwd / reduce-scatter bwd (NCCL)",
                   "exchange": args.gcn_exchange},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "kernel": "whole layer step (gms fwd + gms^T bwd dominate)",
                     "algorithmic_bytes_per_step": alg, "peak_source": peak_src},
        "e2e": {"value": total_edges / (ms_e2e * 1e-3), "unit": "edges/s",
                "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": per * feat * 4, "ms_per_step": ms_e2e},
        "gpu_launches": launches * args.steps, "clocks": clocks,
    }


def cpu_schnet(batch, steps, warmup, threads):
    """Oracle SchNet train step (reference op chain, torch-CPU fp32 autograd incl. the double backward + Adam)."""
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    from oracle import R
    from oracle import models as OM
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import schnet_params
    torch.set_num_threads(threads)
    cfg = SCHNET_C2_CONFIG
    b = schnet_c2_batch(batch, 0)
    params = schnet_params(cfg, np.random.default_rng(0), dtype=torch.float32)
    for v in params.values():
        v.requires_grad_(True)
    plist = list(params.values())
    opt = torch.optim.Adam(plist, lr=1e-3, eps=1e-7)
    z = R.from_row_lengths(torch.from_numpy(b["z"]), b["node_len"])
    xyz = R.from_row_lengths(torch.from_numpy(b["xyz"]), b["node_len"])
    edi = R.from_row_lengths(torch.from_numpy(b["index"]), b["edge_len"])
    e_t, f_t = torch.from_numpy(b["energy"]), torch.from_numpy(b["force"])

    def energy_fn(x):
        return OM.schnet_energy(params, z, x, edi, cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                cfg["last_mlp"]["activation"], "sum", None)

    def step():
        eng, force = OM.energy_force(energy_fn, xyz)
        loss = OM.force_training_loss(eng, force.squeeze(-1), e_t, f_t)
        grads = torch.autograd.grad(loss, plist)
        for p, g in zip(plist, grads):
            p.grad = g * (1.0 / torch.clamp(g.norm(), min=1.0))
        opt.step()
        return loss

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    cpu_schnet.last = (b, sum(int(p.numel()) for p in plist))      # for the reference arm's config block
    return batch / dt, dt * 1e3


def cpu_baseline_schnet(args):
    cores = os.cpu_count() or 1
    steps = 20                                                   # full C2 batch; ~5-10 s of CPU work on the box
    mps, ms = cpu_schnet(256, steps=steps, warmup=2, threads=cores)
    return {"value": mps, "unit": "molecules/s", "cores": cores, "kind": "port",
            "sample": "same SchNet C2 train step at the full batch 256 (oracle: reference op chain on torch-CPU fp32, "
                      "autograd double backward + Adam), %d steps, %.0f ms/step" % (steps, ms)}


def run_reference_schnet(args):
    cores = os.cpu_count() or 1
    batch = 256                                                  # the arm's own config (C2), every step the full batch
    mps, ms = cpu_schnet(batch, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 2)), threads=cores)
    return {"impl": "reference", "metric": "SchNet force-train molecules/sec", "value": mps, "unit": "molecules/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": force_config(FORCE_WORKLOADS["schnet"][3], batch, cpu_schnet.last[0], cpu_schnet.last[1],
                                   int(os.environ.get("WORLD_SIZE", "1"))),
            "cpu_baseline": {"value": mps, "unit": "molecules/s", "cores": cores, "kind": "port",
                             "sample": "oracle torch-CPU restatement of the reference TF op chain, batch %d" % batch},
            "e2e": {"value": mps, "unit": "molecules/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "TensorFlow is not installable here; the reference arm is the oracle port on the host cores"}


# ----------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("KGCNN_BENCH_WORKLOAD", "schnet"), choices=["gss", "schnet", "painn", "hdnnp", "gcn"])
    ap.add_argument("--gss-graphs", type=int, default=65536)
    ap.add_argument("--gss-feat", type=int, default=64)
    ap.add_argument("--gcn-nodes", type=int, default=2000000)
    ap.add_argument("--gcn-exchange", default="halo", choices=["halo", "allgather"],
                    help="C5 multi-GPU exchange: referenced rows only + overlap (default) or the whole feature matrix")
    ap.add_argument("--gcn-feat", type=int, default=128)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-other-workloads", action="store_true",
                    help="default schnet line only: skip the short C3 / C4 / C5 runs reported under other_workloads")
    ap.add_argument("--batch", type=int, default=0, help="molecules per GPU (0 = the config's batch size)")
    ap.add_argument("--host-angles", action="store_true",
                    help="hdnnp e2e leg: copy the angle triples from the host instead of building them on the device")
    ap.add_argument("--profile", action="store_true", help="print a CUPTI kernel table to stderr (diagnostics)")
    args = ap.parse_args()
    global PROFILE
    PROFILE = args.profile
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        fn = {"gss": run_reference_gss}.get(args.workload, run_reference_schnet)
        print(json.dumps(fn(args)))
        return 0

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU oracle)")
    rank, world, local = dist_setup(args.gpus)
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    if args.workload in FORCE_WORKLOADS:
        line = run_force(args, rank, world, device, args.workload)
    else:
        line = {"gss": run_gss, "gcn": run_gcn}[args.workload](args, rank, world, device)
    if args.workload == "schnet" and not args.no_other_workloads:
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        line["other_workloads"] = other_workloads(args, rank, world, device)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            cb = {"gss": cpu_baseline_gss, "schnet": cpu_baseline_schnet}.get(args.workload)
            if cb is not None:
                line["cpu_baseline"] = cb(args)
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
