#!/usr/bin/env python3
"""Small runs of every synchronising kernel for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_run.py
Cases: z-strip kernel (grid barrier), on-chip brick kernel (tagged-word exchange), tiled brick kernel (release/acquire flags, both
publishers), the slab-sharded kernel with one rank, flipEdges (ticket order + per-cell flags), the no-allocation entry point.
Each result is compared with the CPU oracle, so a sanitizer-clean run is also a correct run."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mesh_reconstruction_b200 as mrb
from mesh_reconstruction_b200 import synth, slab
from oracle import oracle as O

iters = int(os.environ.get("SAN_ITERS", "12"))
v = synth.union(synth.sphere(48, 14.0), synth.box(48, (16, 6, 6), None, 0.0, 0.0))
F_ref = O.optimize(v, iters=iters)


def run(env, what):
    for k in ("MRB_QP_LAYOUT", "MRB_QP_ONCHIP", "MRB_QP_DEBUG", "MRB_QP_ORDER"):
        os.environ.pop(k, None)
    os.environ.update(env)
    with mrb.Context(0) as c:
        F = c.optimize(v, iters=iters)
    ok = np.array_equal(F, F_ref)
    print("%-34s %s" % (what, "ok" if ok else "MISMATCH"), flush=True)
    return ok


ok = run({"MRB_QP_LAYOUT": "strips"}, "k_qp_strips (grid barrier)")
ok &= run({"MRB_QP_LAYOUT": "bricks", "MRB_QP_ONCHIP": "1"}, "k_qp_bricks_onchip (tagged words)")
ok &= run({"MRB_QP_LAYOUT": "bricks"}, "k_qp_bricks_tiled (flags)")
ok &= run({"MRB_QP_LAYOUT": "bricks", "MRB_QP_DEBUG": "32"}, "k_qp_bricks_tiled (thread-0 flags)")
ok &= run({"MRB_QP_LAYOUT": "bricks", "MRB_QP_ORDER": "xyz"}, "k_qp_bricks (range flags)")
for k in ("MRB_QP_LAYOUT", "MRB_QP_ONCHIP", "MRB_QP_DEBUG", "MRB_QP_ORDER"):
    os.environ.pop(k, None)
with mrb.Context(0) as c:
    F = slab.solve_fused(c, v, iters=iters).cpu().numpy()
    good = np.array_equal(F, F_ref)
    print("%-34s %s" % ("sharded kernel, world 1", "ok" if good else "MISMATCH"), flush=True)
    ok &= good
    vb = synth.box(40, (9, 7, 6), None, 20.0, 35.0)
    s = O.get_surface(vb)
    n = O.surface_normals(vb, s)
    smap = O.index_map(vb.shape, s)
    Ff = O.optimize(vb, O.feature_map(vb, smap, n, 0.75), using_features=True, iters=iters)
    vr, fr = O.mcubes_features(Ff, smap, n, 0.0, 0.75, 0.8)
    vg, fg = c.mcubes_features(Ff, smap, n, 0.0, 0.75, 0.8)
    good = np.array_equal(vr, vg) and np.array_equal(fr, fg)
    print("%-34s %s" % ("k_flip_edges_par", "ok" if good else "MISMATCH"), flush=True)
    ok &= good
    verts, faces = c.reconstruct(v, iters=iters)
    vr, fr = O.mcubes(F_ref, 0.0)
    good = np.array_equal(verts, vr) and np.array_equal(faces, fr)
    print("%-34s %s" % ("mrb_reconstruct (stages 1-4)", "ok" if good else "MISMATCH"), flush=True)
    ok &= good
sys.exit(0 if ok else 1)
