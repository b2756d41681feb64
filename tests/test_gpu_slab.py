"""Slab-sharded stage 3: 1/2/3/4/8 slabs give the single-GPU result bit for bit (ranks emulated on one GPU:
separate iterate buffers per rank, explicit halo copies between them)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_plan_sweeps_equal_optimize(ctx, O, synth):
    from mesh_reconstruction_b200 import slab
    v = synth.torus(56, 14.0, 5.0)
    plan = slab.Plan(ctx, v)
    band, tight = O.create_bands(O.dfield(v), 4.0)
    assert plan.n_band == int(band.sum()) and plan.n_tight == int(tight.sum())
    assert plan.plane_off[0] == 0 and plan.plane_off[-1] == plan.Q and np.all(np.diff(plan.plane_off) >= 0)
    a, b = plan.new_x(), plan.new_x(init=False)
    for _ in range(30):
        plan.sweep(a, b, 0, plan.Q)
        a, b = b, a
    F = plan.assemble(a).cpu().numpy()
    assert np.array_equal(F, O.optimize(v, iters=30))


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 8])
def test_emulated_slabs_bit_identical(ctx, synth, nranks):
    from mesh_reconstruction_b200 import slab
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    F_ref = ctx.optimize(v, iters=40)
    F = slab.solve_emulated(ctx, v, nranks, iters=40).cpu().numpy()
    assert np.array_equal(F, F_ref)


def test_slab_planes_partition():
    from mesh_reconstruction_b200 import slab
    off = np.array([0, 0, 10, 30, 60, 60, 100])
    for w in (1, 2, 3, 8):
        g = slab.slab_planes(off, w)
        assert g[0][0] == 0 and g[-1][1] == 6
        assert all(g[i][1] == g[i + 1][0] for i in range(w - 1))


def test_fused_sharded_kernel_world1(ctx, synth):
    """The slab-sharded persistent kernel with a single rank (no peers) == mrb_optimize."""
    from mesh_reconstruction_b200 import slab
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    for iters in (1, 40, 41):
        F = slab.solve_fused(ctx, v, iters=iters).cpu().numpy()
        assert np.array_equal(F, ctx.optimize(v, iters=iters)), iters


def _slab_mesh(ctx, F, cuts, iso):
    """Marching cubes slab by slab (mrb_mcubes_slab) + merge_slab_meshes, all on one GPU."""
    import torch
    from mesh_reconstruction_b200 import slab
    nx = F.shape[0]
    parts = []
    for x0, x1 in zip(cuts[:-1], cuts[1:]):
        v, f = ctx.mcubes_slab(np.ascontiguousarray(F[x0:min(x1 + 1, nx)]), x0, iso)
        parts.append((torch.from_numpy(v), torch.from_numpy(f), x0, x1))
    v, f = slab.merge_slab_meshes(parts, F.shape)
    return v.numpy(), f.numpy()


@pytest.mark.parametrize("cuts", [[0, 64], [0, 30, 64], [0, 22, 23, 41, 64], [0, 2, 62, 64]])
def test_sharded_marching_cubes_equals_whole_grid(ctx, synth, cuts):
    """Stage 4 slab by slab: same vertices (bits), same first-appearance ids, same faces as mcubes on the whole grid, for
    cuts through the surface, a one-plane slab and slabs that hold no surface at all."""
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    F = ctx.optimize(v, iters=40)
    verts, faces = ctx.mcubes(F, 0.0)
    sv, sf = _slab_mesh(ctx, F, cuts, 0.0)
    assert sv.shape == verts.shape and np.array_equal(sv.view(np.uint32), verts.view(np.uint32))
    assert np.array_equal(sf, faces)


def test_sharded_marching_cubes_exact_zeros_and_binary_volume(ctx, synth):
    """Vertices that fall exactly on grid points of the cut plane (F == isolevel there) and the degenerate-triangle filter:
    a binary volume at isolevel 0.5 (m_cubes.cpp) and an embedding with exact zeros on the cut plane."""
    v = synth.union(synth.sphere(48, 13.0), synth.box(48, (15, 5, 5), None, 0.0, 0.0))
    verts, faces = ctx.mcubes(v, 0.5)
    sv, sf = _slab_mesh(ctx, v, [0, 17, 24, 31, 48], 0.5)
    assert np.array_equal(sv.view(np.uint32), verts.view(np.uint32)) and np.array_equal(sf, faces)
    F = ctx.optimize(v, iters=30)
    F[24][np.abs(F[24]) < 0.3] = 0.0                      # exact zeros on plane 24 (and the cells around it change sign there)
    verts, faces = ctx.mcubes(F, 0.0)
    sv, sf = _slab_mesh(ctx, F, [0, 24, 48], 0.0)
    assert np.array_equal(sv.view(np.uint32), verts.view(np.uint32)) and np.array_equal(sf, faces)


def test_fused_sharded_back_end_world1(ctx, synth):
    """FusedSolver.reconstruct_sharded with a single rank: slab assembly + slab marching cubes == mrb_reconstruct."""
    import torch
    from mesh_reconstruction_b200 import slab
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    vt = torch.from_numpy(v).cuda()
    solver = slab.FusedSolver(ctx, slab.max_bricks_for(v.shape))
    (sv, sf), F_slab, (x0, x1) = solver.reconstruct_sharded(vt, iters=40)
    verts, faces = ctx.reconstruct(v, iters=40)
    assert (x0, x1) == (0, 64) and np.array_equal(F_slab.cpu().numpy(), ctx.optimize(v, iters=40))
    assert np.array_equal(sv.cpu().numpy().view(np.uint32), verts.view(np.uint32)) and np.array_equal(sf.cpu().numpy(), faces)


def test_slab_local_plan_world1(ctx, synth):
    """FusedSolver.reconstruct_local with a single rank (the sub-grid is the whole grid): == mrb_reconstruct."""
    import torch
    from mesh_reconstruction_b200 import slab
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    vt = torch.from_numpy(v).cuda()
    solver = slab.FusedSolver(ctx, slab.max_bricks_for(v.shape))
    (sv, sf), F_slab, (x0, x1) = solver.reconstruct_local(vt, iters=40)
    verts, faces = ctx.reconstruct(v, iters=40)
    assert (x0, x1) == (0, 64) and np.array_equal(F_slab.cpu().numpy(), ctx.optimize(v, iters=40))
    assert np.array_equal(sv.cpu().numpy().view(np.uint32), verts.view(np.uint32)) and np.array_equal(sf.cpu().numpy(), faces)


@pytest.mark.parametrize("lo,hi", [(10, 40), (0, 33), (27, 64)])
def test_slab_local_stage12_equals_whole_grid(ctx, synth, lo, hi):
    """The exactness claim behind slab-local plans: a plan built from x-planes [lo - margin, hi + margin) assembles, for the
    initial iterate, the same F on [lo, hi) as the whole grid's plan (same band, same signed margins, same +-5 outside)."""
    import torch
    from mesh_reconstruction_b200 import slab
    v = synth.union(synth.sphere(64, 20.0), synth.box(64, (24, 8, 8), None, 0.0, 0.0))
    vt = torch.from_numpy(v).cuda()
    full = slab.Plan(ctx, vt)
    F_full = full.assemble(full.new_x()).cpu().numpy()
    full.close()
    margin = 4 + 5
    a, b = max(0, lo - margin) // 2 * 2, min(64, hi + margin + 1) // 2 * 2
    sub = slab.Plan(ctx, vt[a:b], slab=True)
    F_sub = torch.empty((hi - lo, 64, 64), dtype=torch.float32, device="cuda")
    import ctypes as C
    x = sub.new_x()
    torch.cuda.synchronize()
    sub._check(sub.lib.mrb_qp_plan_assemble_slab(sub.h, C.c_void_p(x.data_ptr()), lo - a, hi - a, C.c_void_p(F_sub.data_ptr())))
    sub.close()
    assert np.array_equal(F_sub.cpu().numpy(), F_full[lo:hi])
