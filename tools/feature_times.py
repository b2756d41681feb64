#!/usr/bin/env python3
"""Step time of the --feature-detection path (BASELINE config 2 shape) on one GPU (tuning aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth
ctx = mrb.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 172
v = synth.config_volume(2) if n == 172 else synth.union(synth.box(n, (n // 3, n // 4, n // 5), None, 20.0, 35.0), synth.sphere(n, n / 4.0))
vt = torch.from_numpy(v).cuda()
for feat in (False, True):
    for _ in range(2): ctx.reconstruct(vt, using_features=feat)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): verts, faces = ctx.reconstruct(vt, using_features=feat)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    print("features" if feat else "default ", v.shape, "step ms %.3f" % (dt * 1e3), "verts", len(verts), "faces", len(faces))
# pieces of the feature path
t = {}
def timed(name, fn):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); t[name] = (time.perf_counter() - t0) * 1e3; return r
s = timed("getSurface", lambda: ctx.getSurface(vt))
nrm = timed("getSurfaceNormals", lambda: ctx.getSurfaceNormals(vt, s))
smap = timed("getIndexMap", lambda: ctx.getIndexMap(v.shape, s))
fm = timed("getFeatureMap", lambda: ctx.getFeatureMap(vt, smap, nrm, 0.75))
F = timed("optimize", lambda: ctx.optimize(vt, fm, True))
m = timed("mcubes_features", lambda: ctx.mcubes_features(F, smap, nrm, 0.0, 0.75, 0.8))
print({k: round(x, 3) for k, x in t.items()})
