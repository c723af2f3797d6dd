"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: graph sharding and the gradient bucket all-reduce."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.distributed import allreduce_gradients, broadcast_parameters, shard_range
    torch.manual_seed(rank)
    grads = [torch.randn(3, 4), torch.randn(5), torch.randn(2, 2, 2)]
    red = allreduce_gradients(grads)
    gathered = [torch.zeros(3, 4) for _ in range(world)]
    dist.all_gather(gathered, grads[0])
    ok = torch.allclose(red[0], sum(gathered) / world) and [r.shape for r in red] == [g.shape for g in grads]
    lin = torch.nn.Linear(3, 2)
    broadcast_parameters(lin)
    w = [torch.zeros_like(lin.weight) for _ in range(world)]
    dist.all_gather(w, lin.weight.data)
    ok = ok and torch.equal(w[0], w[1])
    lo, hi = shard_range(7, rank, world)
    ok = ok and (hi - lo) in (3, 4)
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_gradient_bucket_allreduce_and_broadcast_world2():
    world, port = 2, 29500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))


def test_shard_graphs_by_cost_is_contiguous_and_balanced():
    from gcnn_keras_b200.distributed import shard_graphs_by_cost, shard_range
    costs = np.random.default_rng(0).integers(10, 900, size=512)
    parts = shard_graphs_by_cost(costs, 8)
    assert parts[0][0] == 0 and parts[-1][1] == 512
    assert all(parts[i][1] == parts[i + 1][0] for i in range(7))
    loads = [costs[a:b].sum() for a, b in parts]
    assert max(loads) < 1.1 * (costs.sum() / 8) + costs.max()
    assert [shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]


def _halo_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.partition import halo_all_gather, partition_bounds, slice_edges_for_rank
    per, bounds = partition_bounds(7, world)
    local = torch.full((per, 3), float(rank + 1), requires_grad=True)
    full = halo_all_gather(local)
    ok = full.shape == (per * world, 3) and float(full[0, 0]) == 1.0 and float(full[-1, 0]) == float(world)
    weight = torch.arange(per * world, dtype=torch.float32).reshape(-1, 1) * (rank + 1)
    (full * weight).sum().backward()        # d/d local = reduce-scatter(sum) of every rank's cotangent
    want = sum(torch.arange(per * world, dtype=torch.float32).reshape(-1, 1) * (r + 1) for r in range(world))
    ok = ok and torch.allclose(local.grad, want[rank * per:(rank + 1) * per].expand(per, 3))
    recv = np.array([0, 0, 1, 3, 3, 4, 6])
    e0, e1 = slice_edges_for_rank(recv, rank, per)
    ok = ok and (e0, e1) == ((0, 5) if rank == 0 else (5, 7)) and bounds == [(0, 4), (4, 7)]
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_halo_all_gather_and_reduce_scatter_adjoint_world2():
    world, port = 2, 31500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_halo_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))
