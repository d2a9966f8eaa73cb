"""Rank partitioning of the path's independent units: trajectory sample points, pose hypotheses (with
their rays) and image row-tiles are split across one process per GPU with replicated weights and NO data-path
collective; only the training configuration exchanges anything (all-reduce of table + MLP gradients).
Pure host logic: works with any torch.distributed backend (tested with gloo on CPU, world_size 2)."""
import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """Contiguous [lo, hi) slice of n units owned by `rank`; sizes differ by at most one, every unit owned once."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_tiles(n_frames, rows, tile_rows, rank, world):
    """(frame, row0, row1) image row-tiles dealt round-robin to the ranks (config 4)."""
    tiles = [(f, r, min(r + tile_rows, rows)) for f in range(n_frames) for r in range(0, rows, tile_rows)]
    return tiles[rank::world]


def gather_to_rank0(local, n_total, rank, world):
    """Parity-check helper (not on the timed path): concatenate the ranks' contiguous shards on rank 0."""
    if world == 1:
        return local
    sizes = [hi - lo for lo, hi in (shard_range(n_total, r, world) for r in range(world))]
    mx = max(sizes)  # all_gather wants equal shapes: pad every shard to the largest
    padded = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    bufs = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(bufs, padded)
    return torch.cat([b[:n] for b, n in zip(bufs, sizes)], 0) if rank == 0 else None


def allreduce_mean_(tensors, world):
    """Training configuration only: in-place mean of gradients over the ranks (one flat bucket per dtype)."""
    if world == 1:
        return
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault(t.dtype, []).append(t)
    for ts in by_dtype.values():
        flat = torch.cat([t.reshape(-1) for t in ts])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat /= world
        o = 0
        for t in ts:
            t.copy_(flat[o:o + t.numel()].view_as(t))
            o += t.numel()
