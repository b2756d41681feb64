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
