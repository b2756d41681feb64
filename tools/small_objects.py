import sys, os, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
ctx = mrb.Context(0)
sizes = [int(a) for a in sys.argv[1:]] or [52, 92, 172]
for n, r in [(n, n / 2.0 - 16.0) for n in sizes]:
    v = torch.from_numpy(synth.sphere(n, r)).cuda()
    for _ in range(3): ctx.reconstruct(v, return_stats=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): _, _, st = ctx.reconstruct(v, return_stats=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    print(os.environ.get("MRB_QP_LAYOUT", "bricks"), os.environ.get("MRB_QP_BRICK_THREADS", ""), n, "step ms %.3f" % (dt * 1e3), "qp_kernel %.3f" % st["ms_qp_kernel"], "B", st["n_band"])
