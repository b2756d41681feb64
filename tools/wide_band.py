import sys, os, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
ctx = mrb.Context(0)
v = torch.from_numpy(synth.sphere(256, 110.0)).cuda()
for W in (4, 8):
    for _ in range(3): ctx.reconstruct(v, band_size=W, return_stats=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): _, _, st = ctx.reconstruct(v, band_size=W, return_stats=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    frac = 12.0 * st["n_band"] * 500 / (st["ms_qp_kernel"] * 1e6) / 6551.0
    print("band_size", W, "step ms %.3f" % (dt * 1e3), {k: round(st[k], 3) for k in st if k.startswith("ms_")}, "B", st["n_band"], "qp frac %.3f" % frac)
