"""CPU: the N>1 host logic (frame sharding, result gather, gradient bucket all-reduce) on the gloo
backend with world_size 2 — the same code runs over NCCL on the GPU box."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from plr2_b200 import parallel


@pytest.mark.parametrize("n,world", [(0, 1), (1, 2), (7, 2), (512, 8), (513, 8), (5, 8)])
def test_shard_range_partitions_all_frames(n, world):
    spans = [parallel.shard_range(n, world, r) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (a0, b0), (a1, b1) in zip(spans, spans[1:]):
        assert b0 == a1
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        parallel.shard_range(n, world, world)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                      # identical replicas
        net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 1))
        bucket = parallel.GradBucket(net.parameters())
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        g = torch.Generator().manual_seed(1)
        x = torch.randn(n_frames, 6, generator=g)
        y = torch.randn(n_frames, 1, generator=g)
        mine = parallel.shard_frames({"x": x, "y": y}, world, rank)

        def step_fn(i):
            loss = ((net(mine["x"][i:i + 1]) - mine["y"][i:i + 1]) ** 2).sum()
            loss.backward()
            return loss.detach()

        # SUM semantics (tools/train.py) make the sharded step identical to the single-process one
        parallel.data_parallel_step(range(mine["x"].shape[0]), step_fn, bucket, opt, average=False)
        per_frame = (net(mine["x"]) - mine["y"]).detach().reshape(-1)
        full = parallel.gather_results(per_frame, n_frames)
        if rank == 0:
            out.put(([p.detach().tolist() for p in net.parameters()], full.tolist()))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_step_equals_single_process():
    n_frames, world = 7, 2                        # uneven shards: 4 + 3
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    params2, full = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # single-process reference of the same step
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 1))
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n_frames, 6, generator=g)
    y = torch.randn(n_frames, 1, generator=g)
    ((net(x) - y) ** 2).sum().backward()
    opt.step()
    for a, b in zip(params2, net.parameters()):
        assert torch.allclose(torch.tensor(a), b.detach(), rtol=1e-5, atol=1e-6)
    want = (net(x) - y).detach().reshape(-1)
    full = torch.tensor(full)
    assert full.shape == (n_frames,)
    assert torch.allclose(full, want, rtol=1e-5, atol=1e-6)


def test_bucket_aliases_param_grads_and_is_noop_without_group():
    net = torch.nn.Linear(3, 2)
    b = parallel.GradBucket(net.parameters())
    net(torch.ones(1, 3)).sum().backward()
    assert b.flat.abs().sum() > 0 and b.numel == 8
    assert all(p.grad.data_ptr() >= b.flat.data_ptr() for p in net.parameters())
    assert b.all_reduce() is None
    b.zero()
    assert all(float(p.grad.abs().sum()) == 0.0 for p in net.parameters())


def test_cpulist_parser_and_numa_binding_is_optional():
    from plr2_b200.parallel import _parse_cpulist, bind_to_gpu_numa
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert _parse_cpulist("") == set()
    # no CUDA device / no sysfs topology here: must be a silent no-op, never an exception
    assert bind_to_gpu_numa(0) is None


# ---------------------------------------------------------------------------------------------
# cost-aware assignment of the global training batch (config C5)
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("n_frames", [8, 16, 64, 61])
def test_deal_by_cost_is_a_balanced_partition(world, n_frames):
    sym = [12, 15, 18, 19, 20]
    g = torch.Generator().manual_seed(n_frames * 31 + world)
    objs = torch.randint(0, 21, (n_frames,), generator=g).tolist()
    cost = [parallel.frame_cost(o, sym, 1000, 2300) for o in objs]
    parts = parallel.deal_by_cost(cost, world)
    assert len(parts) == world
    flat = sorted(i for p in parts for i in p)
    assert flat == list(range(n_frames))                                  # a partition: every frame exactly once
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 1                                   # same batch size per rank (+-1)
    nsym = [sum(objs[i] in sym for i in p) for p in parts]
    assert max(nsym) - min(nsym) <= 1                                     # symmetric (expensive) frames level across ranks
    assert parts == parallel.deal_by_cost(cost, world)                    # deterministic: every rank computes the same table
    assert all(p == sorted(p) for p in parts)


def test_deal_by_cost_beats_contiguous_shards_on_a_skewed_batch():
    sym = [12, 15, 18, 19, 20]
    objs = [12, 15, 18, 19, 20, 12, 15, 18] + [0] * 56                    # all symmetric frames at the front
    cost = [parallel.frame_cost(o, sym, 1000, 2300) for o in objs]
    dealt = [sum(cost[i] for i in p) for p in parallel.deal_by_cost(cost, 8)]
    contig = [sum(cost[a:b]) for a, b in (parallel.shard_range(64, 8, r) for r in range(8))]
    assert max(dealt) < 0.2 * max(contig) and max(dealt) / min(dealt) < 1.01


def test_synthetic_global_batch_composition():
    from plr2_b200 import synth
    from plr2_b200.train import synthetic_global_batch
    shp = synth.Shape(32, 21, 20, 8, 8, tuple(synth.YCB_SYM))
    for world in (1, 2, 8):
        frames = synthetic_global_batch(shp, 8 * world)
        nsym = sum(int(f["idx"]) in shp.sym_list for f in frames)
        assert nsym == 2 * world                                          # one frame in four is symmetric
        cost = [parallel.frame_cost(int(f["idx"]), shp.sym_list, 1000, 2300) for f in frames]
        per_rank = [sum(int(frames[i]["idx"]) in shp.sym_list for i in p) for p in parallel.deal_by_cost(cost, world)]
        assert per_rank == [2] * world
