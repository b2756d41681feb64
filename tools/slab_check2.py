#!/usr/bin/env python3
"""torchrun check of the fused slab-sharded solve on a BASELINE config volume: F (and the mesh) must equal this rank's single-GPU
result bit for bit, several times in a row; prints where a mismatch lies.
Usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/slab_check2.py [cfg=4] [reps=3] [iters=500]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import slab, synth

cfgs = [int(c) for c in (sys.argv[1] if len(sys.argv) > 1 else "4").split(",")]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 500
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
ctx = mrb.Context(local)
solver = slab.FusedSolver(ctx, slab.max_bricks_for((512, 512, 512)) // 4)
res = []
for cfg in cfgs:
    v = synth.config_volume_t(cfg, device="cuda") if cfg in (1, 3, 4) else synth.sphere_t(cfg, cfg / 2.0 - 18.0, device="cuda")
    torch.cuda.synchronize()
    F1 = ctx.optimize(v, iters=iters)
    v1, f1 = ctx.reconstruct(v, iters=iters)
    mode = os.environ.get("SLAB_MODE", "local")
    for r in range(reps):
        t = {"phases": {}}
        fn = {"local": solver.reconstruct_local, "sharded": solver.reconstruct_sharded}[mode]
        (v2, f2), Fs, (xa, xb) = fn(v, iters=iters, timings=t)
        ok = bool(torch.equal(F1[xa:xb], Fs))
        info = {"cfg": cfg, "rep": r, "mode": mode, "ok": ok, "kernel_ms": round(t["kernel_ms"], 3), "planes": [xa, xb],
                "phases": {k: round(x, 2) for k, x in t["phases"].items()}}
        if not ok:
            bad = (F1[xa:xb] != Fs).nonzero()
            info.update(n_bad=int(len(bad)), x_min=xa + int(bad[:, 0].min()), x_max=xa + int(bad[:, 0].max()), cuts=[2 * c for c in t["cuts"]],
                        max_abs=float((F1[xa:xb] - Fs).abs().max()))
        info["mesh_ok"] = bool(v1.shape == v2.shape and f1.shape == f2.shape and torch.equal(v1.view(torch.int32), v2.view(torch.int32)) and torch.equal(f1, f2))
        info["ok"] = ok and info["mesh_ok"]
        res.append(info)
    del v, F1
cfg = cfgs
flag = torch.tensor([1 if all(i["ok"] for i in res) else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
for rr in range(world):
    dist.barrier()
    if rr == rank and (rank == 0 or not all(i["ok"] for i in res)):
        print("rank", rank, json.dumps(res), flush=True)
if rank == 0:
    print(json.dumps({"slab_check2": "ok" if int(flag) else "MISMATCH", "world": world, "cfg": cfg}), flush=True)
dist.destroy_process_group()
