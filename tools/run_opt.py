#!/usr/bin/env python3
"""One mrb_optimize on the 256^3 workload (tuning aid): prints the sweep kernel's time; checks F against a saved reference if asked.
Usage: run_opt.py [reps] [size]   env: the MRB_* knobs"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
size = int(sys.argv[2]) if len(sys.argv) > 2 else 256
v = synth.sphere_t(size, 110.0 if size == 256 else size / 2.0 - 18.0, device="cuda")
torch.cuda.synchronize()
ctx = mrb.Context(0)
ts = []
for _ in range(reps):
    F, st = ctx.optimize(v, return_stats=True)
    ts.append(st["ms_qp_kernel"])
print("kernel ms: min %.3f median %.3f" % (min(ts), float(np.median(ts))), "B", st["n_band"], "delta", st["last_max_delta"],
      "sha", __import__("hashlib").sha256(F.cpu().numpy().tobytes()).hexdigest()[:16])
