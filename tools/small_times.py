#!/usr/bin/env python3
"""Stage times of small single objects (the reference's own use case: grids of about 40..64 voxels per edge)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
ctx = mrb.Context(0)
for name, v in (("52^3 sphere r17 (config 1)", synth.sphere(52, 17.0)), ("42^3 sphere r13", synth.sphere(42, 13.0)),
                ("64^3 sphere r24", synth.sphere(64, 24.0)), ("96^3 sphere r40", synth.sphere(96, 40.0)), ("128^3 sphere r54", synth.sphere(128, 54.0))):
    vd = torch.from_numpy(v).cuda()
    for _ in range(3):
        ctx.reconstruct(vd, return_stats=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        _, _, st = ctx.reconstruct(vd, return_stats=True)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 20 * 1e3
    print("%-28s step %.3f ms  B=%d  %s  us/sweep %.2f" % (name, ms, st["n_band"], {k: round(st[k], 3) for k in st if k.startswith("ms_")},
                                                          st["ms_qp_kernel"] * 1000 / 500))
