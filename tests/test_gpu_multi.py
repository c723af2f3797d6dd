"""Two-GPU NCCL checks of the partitioned-graph path (skipped on a one-GPU box; run with `gpurun --gpus 2`)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.partition import PartitionedGraph, gcn_layer_partitioned, partition_bounds
    dev = torch.device("cuda", rank)
    n_nodes, feat = 5000, 64
    rng = np.random.default_rng(3)                         # the same graph on every rank
    recv = np.sort(rng.integers(0, n_nodes, size=60000))
    send = rng.integers(0, n_nodes, size=60000)
    w = rng.random(60000).astype(np.float32)
    per, bounds = partition_bounds(n_nodes, world)
    torch.manual_seed(0)
    x_full = torch.randn(per * world, feat, device=dev)
    kernel = (torch.randn(feat, feat, device=dev) / 8).requires_grad_(True)
    bias = torch.zeros(feat, device=dev, requires_grad=True)
    dy = torch.randn(per * world, feat, device=dev)[rank * per:(rank + 1) * per]
    outs = {}
    for mode in ("halo", "allgather"):
        graph = PartitionedGraph(n_nodes, recv, send, w, rank, world, dev, exchange=mode)
        x = x_full[rank * per:(rank + 1) * per].clone().requires_grad_(True)
        y = gcn_layer_partitioned(graph, x, kernel, bias, "relu")
        gx, gk, gb = torch.autograd.grad(y, [x, kernel, bias], dy)
        outs[mode] = (y.detach(), gx, gk, gb)
        if mode == "halo":
            n_halo = graph.halo.n_halo
    ok = n_halo < per * (world - 1)                        # fewer rows travel than the all-gather moves
    # both schemes against the fp64 oracle of the UNPARTITIONED layer (oracle.conv.gcn): this rank's rows of y and dx,
    # and the weight gradients summed over the ranks
    from oracle import R
    from oracle import conv as OC
    ox = x_full[:n_nodes].detach().cpu().double().requires_grad_(True)
    prm = {"lay_dense.kernel": kernel.detach().cpu().double().requires_grad_(True),
           "lay_dense.bias": bias.detach().cpu().double().requires_grad_(True)}
    want = OC.gcn(OC.P(prm), R.from_row_lengths(ox, np.array([n_nodes])),
                  R.from_row_lengths(torch.tensor(w, dtype=torch.float64).reshape(-1, 1), np.array([len(recv)])),
                  R.from_row_lengths(torch.tensor(np.stack([recv, send], 1)), np.array([len(recv)])),
                  "relu", "sum", False).values
    dy_all = [torch.empty_like(dy) for _ in range(world)]
    dist.all_gather(dy_all, dy.contiguous())
    cot = torch.cat(dy_all, 0)[:n_nodes].cpu().double()
    ogx, ogk, ogb = torch.autograd.grad(want, [ox, prm["lay_dense.kernel"], prm["lay_dense.bias"]], cot)
    lo, hi = bounds[rank]
    for mode in ("halo", "allgather"):
        y, gx, gk, gb = outs[mode]
        gk, gb = gk.clone(), gb.clone()
        dist.all_reduce(gk)
        dist.all_reduce(gb)
        for a, b in ((y[:hi - lo], want[lo:hi]), (gx[:hi - lo], ogx[lo:hi]), (gk, ogk), (gb, ogb)):
            b = b.detach()
            ok = ok and float((a.cpu().double() - b).abs().max()) <= 1e-5 * (float(b.abs().max()) + 1e-30)
    for a, b in zip(outs["halo"], outs["allgather"]):
        scale = float(b.abs().max()) + 1e-30
        ok = ok and float((a - b).abs().max()) <= 2e-6 * scale
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_halo_exchange_equals_all_gather_scheme_on_two_gpus():
    world, port = 2, 35500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))


def _peer_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.distributed import PeerAllReduce
    dev = torch.device("cuda", rank)
    n = 469_123                                             # not a multiple of 4: the tail path too
    peer = PeerAllReduce.create(n, dev)
    ok = peer is not None
    if ok:
        for step in range(3):                               # reuse across steps: the two barriers order the reuse
            torch.manual_seed(100 * step + rank)
            g = torch.randn(n, device=dev)
            peer.buffer.copy_(g)
            out = peer.reduce(average=True).clone()
            want = g.double().clone()
            dist.all_reduce(want, op=dist.ReduceOp.SUM)
            want = want / world
            ok = ok and float((out.double() - want).abs().max()) <= 1e-6 * float(want.abs().max())
            gathered = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(gathered, out)
            ok = ok and all(bool(torch.equal(gathered[0], t)) for t in gathered)   # bitwise identical on every rank
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_memory_all_reduce_matches_nccl_and_is_identical_on_every_rank():
    world, port = 2, 37500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert all(ret.get(r) for r in range(world)), dict(ret)


def _group_guard_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from test_gpu_guard import small_schnet_step, unsorted_copy
    from gcnn_keras_b200.training import CapturedForceTrainStep
    dev = torch.device("cuda", rank)
    model, step, good = small_schnet_step(dev, seed=rank)
    captured = CapturedForceTrainStep(step, good, dev, warmup=2)
    opt = step.optimizer

    def same_on_all_ranks(t):
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous())
        return all(bool(torch.equal(parts[0], q)) for q in parts)

    captured.run(good)
    torch.cuda.synchronize()
    ok = captured.check() and same_on_all_ranks(opt.flat)
    before = [t.clone() for t in (opt.flat, opt.m, opt.v, opt.step_count)]
    # ONLY rank 1 gets a batch that breaks its plan; rank 0's batch is fine
    captured.run(unsorted_copy(good) if rank == 1 else good)
    torch.cuda.synchronize()
    ok = ok and not captured.check()                               # reported on BOTH ranks
    ok = ok and all(bool(torch.equal(a, b)) for a, b in zip(before, (opt.flat, opt.m, opt.v, opt.step_count)))
    try:
        captured.run(good)
        ok = False
    except RuntimeError:
        pass                                                       # raised on both ranks, nobody is left in a barrier
    captured.run(good)
    torch.cuda.synchronize()
    ok = ok and captured.check() and int(opt.step_count) == int(before[3]) + 1
    ok = ok and same_on_all_ranks(opt.flat) and not torch.equal(before[0], opt.flat)
    ret[rank] = (bool(ok), "peer" if opt.peer is not None else "nccl")
    # a CUDA graph that captured an NCCL collective keeps a reference on the communicator: release it before the
    # process group goes away, or the communicator's destruction waits for it
    captured.graph = None
    del captured
    import gc
    gc.collect()
    torch.cuda.synchronize()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("peer", ["auto", "0"])
def test_one_rank_breaking_its_plan_makes_every_replica_skip_the_update(peer):
    """The guard word travels through the step's all-reduce (peer-memory kernel and NCCL path): replicas stay
    bitwise identical and every rank raises."""
    world, port = 2, 39500 + (os.getpid() % 2000) + (7 if peer == "0" else 0)
    os.environ["KGCNN_PEER_ALLREDUCE"] = peer
    try:
        ctx = mp.get_context("spawn")
        ret = ctx.Manager().dict()
        procs = [ctx.Process(target=_group_guard_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(150)
        hung = [p for p in procs if p.is_alive()]
        for p in hung:                                   # never leave a rank behind: pytest would wait for it at exit
            p.kill()
            p.join(10)
        assert not hung, "a rank did not finish: %s" % dict(ret)
        assert all(p.exitcode == 0 for p in procs)
    finally:
        os.environ.pop("KGCNN_PEER_ALLREDUCE", None)
    assert all(ret.get(r, (False,))[0] for r in range(world)), dict(ret)
    if peer == "0":
        assert all(ret[r][1] == "nccl" for r in range(world))
