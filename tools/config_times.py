#!/usr/bin/env python3
"""Step times of the other BASELINE configs on one GPU (tuning aid, not a bench number)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth

ctx = mrb.Context(0)
what = sys.argv[1:] or ["512", "batch"]
if "512" in what:
    v = torch.from_numpy(synth.sphere(512, 220.0)).cuda()
    for _ in range(2):
        ctx.reconstruct(v, return_stats=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        _, _, st = ctx.reconstruct(v, return_stats=True)
    torch.cuda.synchronize()
    print("512^3 step ms %.2f" % ((time.perf_counter() - t0) / 3 * 1e3), {k: round(st[k], 3) for k in st if k.startswith("ms_")}, st["n_band"])
    del v
if "batch" in what:
    n = 1024 if "1024" in what else 128
    vols = [synth.random_object(1000 + i) for i in range(n)]
    stack = np.stack(vols)
    ctx.reconstruct_batch(stack[:8])
    for label, env in (("stacked", None), ("one by one", "1")):
        if env:
            os.environ["MRB_BATCH_SEQUENTIAL"] = env
        else:
            os.environ.pop("MRB_BATCH_SEQUENTIAL", None)
        if env and n > 128:
            continue
        ctx.reconstruct_batch(stack)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = ctx.reconstruct_batch(stack)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print("batch of %d x 52^3, one context, %s: %.1f ms total, %.0f meshes/s" % (n, label, dt * 1e3, n / dt))
    os.environ.pop("MRB_BATCH_SEQUENTIAL", None)
    dstack = torch.from_numpy(stack).cuda()
    ctx.reconstruct_batch(dstack)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = ctx.reconstruct_batch(dstack)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("batch of %d x 52^3, volumes resident in HBM, stacked: %.1f ms total, %.0f meshes/s" % (n, dt * 1e3, n / dt))
    del dstack
    if n > 128:
        sys.exit(0)
    for workers in (2, 4, 8):
        pool = mrb.Pool(0, workers=workers)
        pool.map(vols[:2 * workers])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pool.map(vols)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print("batch of %d x 52^3, Pool(%d): %.1f ms total, %.0f meshes/s" % (n, workers, dt * 1e3, n / dt))
        pool.close()
