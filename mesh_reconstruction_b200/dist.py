"""Multi-GPU host logic: one process per GPU, torch.distributed for the plumbing.

The hot path shards two ways and neither needs a data-path collective for object-parallel work:
  * object-parallel: independent objects, a contiguous block of object ids per rank;
  * slab-parallel: one grid split along x (the slowest storage axis, so the reference's band /
    cell / triangle orders concatenate across ranks), each slab extended by a halo.
Collectives are only used for timing (max over ranks) and for gathering counts.
"""
import os


def env_world():
    """(rank, local_rank, world_size) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def shard_objects(n_objects, world, rank):
    """Contiguous, balanced block of object ids for `rank`: sizes differ by at most one."""
    base, rem = divmod(n_objects, world)
    start = rank * base + min(rank, rem)
    return range(start, start + base + (1 if rank < rem else 0))


def slab_bounds(nx, world, rank, halo=0):
    """Owned x-range [x0, x1) of `rank` and the halo-extended range [h0, h1) clipped to the grid."""
    r = shard_objects(nx, world, rank)
    x0, x1 = r.start, r.stop
    return (x0, x1), (max(0, x0 - halo), min(nx, x1 + halo))


def max_over_ranks(values, device=None):
    """Element-wise maximum of a list of floats over all ranks (identity when not initialised)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(v) for v in values]
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.cpu()]


def exclusive_prefix_over_ranks(value, device=None):
    """Exclusive prefix sum of one integer per rank (global ids = local id + prefix), plus the total."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 0, int(value)
    world, rank = dist.get_world_size(), dist.get_rank()
    t = torch.zeros(world, dtype=torch.int64, device=device)
    t[rank] = int(value)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t = t.cpu()
    return int(t[:rank].sum()), int(t.sum())
