#!/usr/bin/env python3
"""One stacked batch of n x 52^3 objects (config 5) from pinned host memory: total time, for a launch list under ncu."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
stack = torch.empty((n, 52, 52, 52), dtype=torch.float32).pin_memory()
host = stack.numpy()
for i in range(n):
    host[i] = synth.random_object(1000 + i)
ctx = mrb.Context(0)
ctx.reconstruct_batch(host[:8])
ctx.reconstruct_batch(host)
ts = []
for _ in range(reps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = ctx.reconstruct_batch(host)
    torch.cuda.synchronize()
    ts.append((time.perf_counter() - t0) * 1e3)
dev = stack.cuda()
ctx.reconstruct_batch(dev)
torch.cuda.synchronize()
t0 = time.perf_counter()
ctx.reconstruct_batch(dev)
torch.cuda.synchronize()
td = (time.perf_counter() - t0) * 1e3
print("batch %d x 52^3: host->host %.2f ms (%.0f meshes/s); volumes resident -> host meshes %.2f ms; faces %d" %
      (n, min(ts), n / min(ts) * 1e3, td, sum(len(m[1]) for m in out)))
