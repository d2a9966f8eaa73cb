"""N > 1 host logic on CPU (gloo, world_size 2): unit partitioning covers everything exactly once, shards reassemble
on rank 0, and the training all-reduce averages gradients."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_navigation_b200.sharding import shard_range, shard_tiles, gather_to_rank0, allreduce_mean_


def test_shard_range_partitions():
    for n in (0, 1, 7, 256, 4194304 + 3):
        for world in (1, 2, 3, 8):
            cuts = [shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_shard_tiles_cover_frames():
    for world in (1, 2, 4, 8):
        seen = sorted(t for r in range(world) for t in shard_tiles(3, 800, 100, r, world))
        assert seen == sorted((f, r, r + 100) for f in range(3) for r in range(0, 800, 100))


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 11
    full = torch.arange(n * 3, dtype=torch.float32).view(n, 3)
    lo, hi = shard_range(n, rank, world)
    got = gather_to_rank0(full[lo:hi] * 2.0, n, rank, world)       # each rank processes its own slice
    ok = True
    if rank == 0:
        ok = torch.equal(got, full * 2.0)
    g1 = torch.full((5,), float(rank + 1)); g2 = torch.full((2, 2), float(10 * (rank + 1)), dtype=torch.float64)
    allreduce_mean_([g1, g2], world)
    ok = ok and torch.allclose(g1, torch.full((5,), 1.5)) and torch.allclose(g2, torch.full((2, 2), 15.0, dtype=torch.float64))
    # the training step's gradient exchange (training.allreduce_grads_): SUM over ranks into .grad, in place
    from nerf_navigation_b200.training import allreduce_grads_
    ps = [torch.nn.Parameter(torch.zeros(6)), torch.nn.Parameter(torch.zeros(3, 2)), torch.nn.Parameter(torch.zeros(4))]
    ps[0].grad = torch.arange(6, dtype=torch.float32) * (rank + 1)
    ps[1].grad = torch.full((3, 2), 2.0 * (rank + 1))
    allreduce_grads_(ps, world)                                     # ps[2] has no gradient and is skipped
    ok = ok and torch.equal(ps[0].grad, torch.arange(6, dtype=torch.float32) * 3) and torch.equal(ps[1].grad, torch.full((3, 2), 6.0))
    ok = ok and ps[2].grad is None
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_gloo_world2():
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret[0] and ret[1]
