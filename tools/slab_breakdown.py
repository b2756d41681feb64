#!/usr/bin/env python3
"""Where the time of the slab-sharded step goes (torchrun, N ranks): wall-clock per phase with a device synchronise after each.
Usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/slab_breakdown.py [size]"""
import os, sys, time, json, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import slab, synth

size = int(sys.argv[1]) if len(sys.argv) > 1 else 256
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = mrb.Context(local)
v = synth.sphere_t(size, 110.0 if size == 256 else size / 2.0 - 18.0, device="cuda")
torch.cuda.synchronize()
solver = slab.FusedSolver(ctx, slab.max_bricks_for(v.shape) // 4)
T = {}
def tick(name, t0):
    torch.cuda.synchronize()
    T[name] = T.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return time.perf_counter()
for rep in range(4):
    if rep == 1:
        T.clear()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    plan = slab.Plan(ctx, v, 4); t = tick("plan_create", t)
    tm = {}
    x = solver.solve_plan(plan, 500, tm); t = tick("solve_plan_total", t)
    F = plan.assemble(x); t = tick("assemble", t)
    plan.close(); t = tick("plan_close", t)
    m = ctx.mcubes(F, 0.0); t = tick("mcubes", t)
    T["kernel_ms"] = T.get("kernel_ms", 0.0) + tm["kernel_ms"]
if dist.get_rank() == 0:
    print(json.dumps({k: round(val / 3, 3) for k, val in T.items()}))
ph = {}
for rep in range(4):
    if rep == 1:
        ph.clear()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    solver.reconstruct_sharded(v, iters=500, timings={"phases": ph})
if dist.get_rank() == 0:
    print("sharded back end:", json.dumps({k: round(val / 3, 3) for k, val in ph.items()}))
dist.destroy_process_group()
