#!/usr/bin/env python3
"""torchrun check of the NCCL slab-sharded QP: F must equal the single-GPU mrb_optimize bit for bit on every rank.
Usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/slab_check.py [size] [iters]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import slab, synth

size = int(sys.argv[1]) if len(sys.argv) > 1 else 128
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 100
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = mrb.Context(local)
v = synth.sphere(size, size / 2.0 - 18.0)
vt = torch.from_numpy(v).cuda(local)
F_ref, st = ctx.optimize(vt, iters=iters, return_stats=True)
mode = sys.argv[3] if len(sys.argv) > 3 else "fused"
t = {}
if mode == "nccl":
    F = slab.solve_distributed(ctx, vt, iters=iters, timings=t)
else:
    slab.solve_fused(ctx, vt, iters=min(iters, 20))        # warm-up (module load, symmetric-memory setup)
    F = slab.solve_fused(ctx, vt, iters=iters, timings=t)
    t["solve_ms"] = t["kernel_ms"]
ok = bool(torch.equal(F, F_ref))
if mode != "nccl" and ok:
    # the whole path: mesh of the sharded run == mesh of the single-GPU run
    v1, f1 = ctx.reconstruct(vt, iters=iters)
    v2, f2 = slab.reconstruct_fused(ctx, vt, iters=iters)
    host = lambda a: a.cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    ok = bool(np.array_equal(host(f1), host(f2)) and np.array_equal(host(v1), host(v2)))
if not ok and dist.get_rank() == 0:
    bad = (F != F_ref).nonzero()
    print("mismatch voxels", len(bad), "x range", int(bad[:, 0].min()), int(bad[:, 0].max()), "cuts", t.get("cuts"),
          "max abs diff", float((F - F_ref).abs().max()), flush=True)
flag = torch.tensor([1 if ok else 0], device=vt.device)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if dist.get_rank() == 0:
    print(json.dumps({"slab_check": "ok" if int(flag) else "MISMATCH", "world": dist.get_world_size(), "size": size, "iters": iters,
                      "n_band": st["n_band"], "single_gpu_kernel_ms": st["ms_qp_kernel"], "mode": mode, "slab_solve_ms": t["solve_ms"],
                      "us_per_sweep_slab": 1000.0 * t["solve_ms"] / iters}))
dist.destroy_process_group()
sys.exit(0 if int(flag) else 1)
