import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import slab, synth
ctx = mrb.Context(0)
v = synth.sphere_t(256, 110.0, device="cuda"); torch.cuda.synchronize()
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
def whole():
    p = slab.Plan(ctx, v); p.close()
def sub(a, b):
    def f():
        p = slab.Plan(ctx, v[a:b], slab=True); p.close()
    return f
print("whole grid plan  %.3f ms" % t(whole))
print("sub [0:140]      %.3f ms" % t(sub(0, 140)))
print("sub [116:256]    %.3f ms" % t(sub(116, 256)))
print("sub [60:200]     %.3f ms" % t(sub(60, 200)))
print("whole again      %.3f ms" % t(whole))
