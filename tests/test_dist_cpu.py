"""N > 1 host logic on CPU: world_size-2 gloo process group (no GPU, no compute calls)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mesh_reconstruction_b200 import dist as mdist


def test_shard_objects_partition():
    for n, w in ((1024, 8), (10, 3), (2, 4), (0, 2)):
        got = [i for r in range(w) for i in mdist.shard_objects(n, w, r)]
        assert got == list(range(n))
        sizes = [len(mdist.shard_objects(n, w, r)) for r in range(w)]
        assert max(sizes) - min(sizes) <= 1


def test_slab_bounds_cover_with_halo():
    nx, w = 256, 8
    owned = []
    for r in range(w):
        (x0, x1), (h0, h1) = mdist.slab_bounds(nx, w, r, halo=2)
        owned += list(range(x0, x1))
        assert h0 == max(0, x0 - 2) and h1 == min(nx, x1 + 2)
    assert owned == list(range(nx))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert mdist.env_world() == (rank, rank, world)
    mx = mdist.max_over_ranks([1.0 + rank, 10.0 - rank])
    pre, tot = mdist.exclusive_prefix_over_ranks(100 * (rank + 1))
    mine = list(mdist.shard_objects(5, world, rank))
    dist.barrier()
    q.put((rank, mx, pre, tot, mine))
    dist.destroy_process_group()


def test_gloo_world2_timing_and_prefix():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1] == [2.0, 10.0]          # max over ranks, identical on every rank
    assert (res[0][2], res[0][3]) == (0, 300) and (res[1][2], res[1][3]) == (100, 300)
    assert res[0][4] + res[1][4] == [0, 1, 2, 3, 4]
