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


def _halo_plan_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.partition import HaloPlan, _all_to_all_rows, partition_bounds
    n_nodes, feat = 23, 3
    per, bounds = partition_bounds(n_nodes, world)
    rng = np.random.default_rng(11)                       # the same graph on both ranks; each keeps its receivers
    recv_all = np.sort(rng.integers(0, n_nodes, size=90))
    send_all = rng.integers(0, n_nodes, size=90)
    mine = (recv_all // per) == rank
    send = send_all[mine]
    plan = HaloPlan(send, rank, world, per, torch.device("cpu"))
    ok = True
    # what the OTHER rank reads from this rank's block, from the global edge list
    for q in range(world):
        theirs = np.unique(send_all[((recv_all // per) == q) & ((send_all // per) == rank)]) if q != rank else []
        got = plan.send_ids.numpy()[sum(plan.send_counts[:q]):sum(plan.send_counts[:q + 1])]
        ok = ok and np.array_equal(got, np.asarray(theirs, dtype=np.int64) - rank * per)
    ok = ok and plan.n_halo == len(np.unique(send[(send // per) != rank])) and sum(plan.need_counts) == plan.n_halo
    # forward exchange: the halo buffer holds exactly the remote rows, in ascending global id order
    table = torch.arange(per * world * feat, dtype=torch.float32).reshape(per * world, feat)
    local = table[rank * per:(rank + 1) * per]
    _, halo = _all_to_all_rows(local[plan.send_ids.long()], plan.send_counts, plan.need_counts)
    ok = ok and torch.equal(halo, table[torch.as_tensor(plan.halo_ids)])
    # transpose: <halo(h), g> summed over the ranks == <h, back(g)> summed over the ranks
    g_halo = torch.from_numpy(np.random.default_rng(rank).normal(size=(plan.n_halo, feat)).astype(np.float32))
    _, back = _all_to_all_rows(g_halo, plan.need_counts, plan.send_counts)
    dlocal = torch.zeros(per, feat).index_add_(0, plan.send_ids.long(), back)
    sums = torch.tensor([float((halo * g_halo).sum()), float((local * dlocal).sum())], dtype=torch.float64)
    dist.all_reduce(sums)
    ok = ok and abs(float(sums[0] - sums[1])) <= 1e-6 * abs(float(sums[0]))
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_halo_plan_exchanges_only_referenced_rows_and_is_adjoint_world2():
    world, port = 2, 33500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_halo_plan_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))


def _isolated_shard_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.partition import HaloPlan, partition_bounds
    n_nodes, feat = 12, 2
    per, _ = partition_bounds(n_nodes, world)             # 4 rows per rank
    # ranks 0 and 1 read each other's rows; rank 2's edges stay inside its own block and nobody reads its rows
    send = {0: [0, 1, 5, 6], 1: [4, 2, 3, 7], 2: [8, 9, 10, 11, 8]}[rank]
    plan = HaloPlan(np.asarray(send), rank, world, per, torch.device("cpu"))
    ok = plan.exchange_needed                             # the same answer on all three ranks
    if rank == 2:
        ok = ok and plan.n_halo == 0 and plan.n_send == 0
    table = torch.arange(per * world * feat, dtype=torch.float32).reshape(per * world, feat)
    local = table[rank * per:(rank + 1) * per]
    for _ in range(2):                                    # two layers: the collective sequence stays aligned
        _, halo = plan.exchange_rows(local[plan.send_ids.long()])
        ok = ok and torch.equal(halo, table[torch.as_tensor(plan.halo_ids, dtype=torch.int64)])
        _, back = plan.exchange_rows(halo, reverse=True)
        ok = ok and torch.equal(back, local[plan.send_ids.long()])
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_isolated_shard_still_joins_the_halo_exchange_world3():
    """A rank with no row crossing the cut calls the all-to-all with zero counts instead of skipping it (the gloo
    emulation is built from true collectives: it hangs if one rank stays out)."""
    world, port = 3, 35500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_isolated_shard_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))


def _no_crossing_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from gcnn_keras_b200.partition import HaloPlan
    plan = HaloPlan(np.asarray([rank * 4, rank * 4 + 3]), rank, world, 4, torch.device("cpu"))   # block-diagonal graph
    ret[rank] = (not plan.exchange_needed) and plan.exchange_rows(torch.zeros(0, 2)) == (None, None)
    dist.destroy_process_group()


def test_block_diagonal_graph_needs_no_exchange_on_any_rank_world2():
    world, port = 2, 37500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_no_crossing_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))
