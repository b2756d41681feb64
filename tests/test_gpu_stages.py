"""GPU parity tests, stage by stage, through the C-ABI (libmrb.so) against the CPU oracle.

Integer / index / topology outputs must be bit-exact; the margin and the QP iterate are compared
with tolerance 0 as well (design target: bit-exact), the documented bound being 1e-4 voxel units.
"""
import os

import numpy as np
import pytest

from conftest import random_blobs

pytestmark = pytest.mark.gpu

TOL = 1e-4   # north_star tolerance for float outputs (voxel units); we assert exact equality first


def volumes(synth):
    yield "sphere52", synth.sphere(52, 17.0)
    yield "blobs40", random_blobs((40, 44, 48), seed=1)
    yield "box_rot", synth.box(48, (12, 9, 7), None, 20.0, 35.0)
    yield "torus", synth.torus(56, 14.0, 5.0)


def test_fast_perim(ctx, O, synth):
    for name, v in volumes(synth):
        assert np.array_equal(ctx.fastPerim(v), O.fast_perim(v)), name
    # arbitrary float values (the reference compares with !=)
    rng = np.random.default_rng(0)
    v = rng.integers(0, 3, size=(9, 10, 11)).astype(np.float32) * 0.5
    assert np.array_equal(ctx.fastPerim(v), O.fast_perim(v))


def test_sqdist_exact(ctx, O, synth):
    for name, v in volumes(synth):
        p = O.fast_perim(v)
        assert np.array_equal(ctx.getsqdist(p), O.sqdist(p, brute=True)), name
    # sparse sites, non-cubic grid
    rng = np.random.default_rng(3)
    g = (rng.random((17, 23, 29)) > 0.002).astype(np.float32)
    assert np.array_equal(ctx.getsqdist(g), O.sqdist(g, brute=True))
    # no zero voxel at all -> 100000 everywhere (dfields.cpp:50-58)
    g = np.ones((8, 9, 10), np.float32)
    assert np.array_equal(ctx.getsqdist(g), np.full(g.shape, 100000.0, np.float32))


def test_sqdist_clamp(ctx, O):
    # a single zero voxel in a long grid: squared distances beyond 100000 are clamped (needs > 317 voxels)
    g = np.ones((340, 3, 4), np.float32)
    g[2, 1, 1] = 0.0
    got = ctx.getsqdist(g)
    ref = O.sqdist(g, brute=True)
    assert got.max() == 100000.0
    assert np.array_equal(got, ref)
    g = np.ones((3, 4, 350), np.float32)
    g[1, 2, 340] = 0.0
    assert np.array_equal(ctx.getsqdist(g), O.sqdist(g, brute=True))
    g = np.ones((3, 345, 4), np.float32)
    g[1, 5, 2] = 0.0
    assert np.array_equal(ctx.getsqdist(g), O.sqdist(g, brute=True))


def test_sqrt_and_dfield(ctx, O, synth):
    for name, v in volumes(synth):
        m_ref = O.dfield(v, brute=True)
        m = ctx.dfield_gpu(v)
        assert np.array_equal(m, m_ref), name
        s = O.sqdist(O.fast_perim(v))
        assert np.array_equal(ctx.getsqrt(s), O.sqrt(s)), name


def test_bands(ctx, O, synth):
    for name, v in volumes(synth):
        m = O.dfield(v)
        band_ref, tight_ref = O.create_bands(m, 4.0)
        band, tight = ctx.createBands(m, 4.0)
        assert np.array_equal(band, band_ref) and np.array_equal(tight, tight_ref), name
        assert np.array_equal(ctx.erode_grid(band_ref), O.erode(band_ref)), name
        assert np.array_equal(ctx.dilate_grid(tight_ref), O.dilate(tight_ref)), name
    # border voxels are copied, not morphed (narrowBand.cpp:11-14)
    rng = np.random.default_rng(5)
    g = (rng.random((7, 8, 9)) > 0.4).astype(np.float32)
    assert np.array_equal(ctx.erode_grid(g), O.erode(g))
    assert np.array_equal(ctx.dilate_grid(g), O.dilate(g))


def test_find_indexes_and_index_map(ctx, O, synth):
    for name, v in volumes(synth):
        band, _ = O.create_bands(O.dfield(v), 4.0)
        idx_ref = O.find_indexes(band)
        idx = ctx.findIndexes(band)
        assert np.array_equal(idx, idx_ref), name
        assert np.array_equal(ctx.getIndexMap(band.shape, idx), O.index_map(band.shape, idx_ref)), name
    empty = np.zeros((5, 6, 7), np.float32)
    assert len(ctx.findIndexes(empty)) == 0


def test_prime_qp(ctx, O, synth):
    for name, v in volumes(synth):
        m = O.dfield(v)
        band, tight = O.create_bands(m, 4.0)
        ref = O.prime_qp(v, m, band, tight)
        got = ctx.primeQP(v, m, band, tight)
        assert got["n"] == ref["n"], name
        for k in ("lb", "ub", "x0", "invdg", "jc", "ir", "pr"):
            assert np.array_equal(got[k], ref[k]), (name, k)


def test_run_qp_csc(ctx, O, synth):
    for name, v in volumes(synth):
        m = O.dfield(v)
        band, tight = O.create_bands(m, 4.0)
        q = O.prime_qp(v, m, band, tight)
        for iters in (1, 2, 37):
            assert np.array_equal(ctx.runQP(q, iters=iters), O.run_qp(q, iters=iters)), (name, iters)


def test_run_qp_feature_reset(ctx, O, synth):
    v = synth.sphere(40, 11.0)
    m = O.dfield(v)
    band, tight = O.create_bands(m, 4.0)
    q = O.prime_qp(v, m, band, tight)
    n = q["n"]
    for feat in ([5, 17, 300], [-1, 5, 17], [n - 1], []):
        f = np.array(feat, np.int32)
        ref = O.run_qp(q, iters=5, feature_idx=f, using_features=True)
        got = ctx.runQP(q, f, True, iters=5)
        assert np.array_equal(got, ref), feat


def test_optimize_matrix_free(ctx, O, synth):
    """The matrix-free persistent stencil kernel == 500 CSC sweeps of the reference, bit for bit."""
    for name, v in volumes(synth):
        for iters in (3, 500):
            F_ref, info = O.optimize(v, iters=iters, return_info=True)
            F, st = ctx.optimize(v, iters=iters, return_stats=True)
            assert st["n_band"] == info["n_band"], name
            assert np.abs(F - F_ref).max() <= TOL, name
            assert np.array_equal(F, F_ref), (name, iters)


def test_optimize_sync_and_layout_variants(ctx, O, synth, monkeypatch):
    """Every selectable stage-3 variant (tuning knobs) must give the same bits as the default."""
    v = synth.torus(56, 14.0, 5.0)
    F_ref = O.optimize(v, iters=60)
    deltas = []
    B = {"MRB_QP_LAYOUT": "bricks"}          # grids this small default to the strip kernel
    X = dict(B, MRB_QP_ORDER="xyz")          # the older (X,Y,Z)-ordered brick kernel
    for env in (B, dict(B, MRB_QP_BRICK_THREADS="512"), dict(B, MRB_QP_BRICK_THREADS="256"), dict(B, MRB_QP_BRICK_COMPACT="1"),
                dict(B, MRB_QP_BRICK_NOSMEM="1"), dict(B, MRB_QP_TILE="3"), dict(B, MRB_QP_MIN_THREADS="128"),
                dict(B, MRB_QP_DEBUG="32"),      # sweep flag published by thread 0 (the slab-sharded kernel's publisher)
                # the opt-in on-chip kernel (iterate + halo in shared memory, tagged-word exchange): both fetch modes, two block sizes
                dict(B, MRB_QP_ONCHIP="1"), dict(B, MRB_QP_ONCHIP="1", MRB_QP_OC_MODE="2"),
                dict(B, MRB_QP_ONCHIP="1", MRB_QP_OC_THREADS="352"), dict(B, MRB_QP_ONCHIP="1", MRB_QP_ORDER="morton"),
                dict(B, MRB_QP_ORDER="morton"), dict(B, MRB_QP_ORDER="morton:2,2,8"),
                X, dict(X, MRB_QP_SYNC="flow"), dict(X, MRB_QP_BRICK_THREADS="512"), dict(X, MRB_QP_BRICK_THREADS="128"),
                dict(X, MRB_QP_BRICK_NOSMEM="1"), {"MRB_QP_LAYOUT": "strips"},
                # strips: the default for a grid this small is the thread-block-cluster kernel (band in distributed shared
                # memory); cluster sizes 16 / 8 / 4 and the L2-resident kernel it replaces
                {"MRB_QP_CLUSTER": "16"}, {"MRB_QP_CLUSTER": "8"}, {"MRB_QP_CLUSTER": "4"}, {"MRB_QP_CLUSTER": "0"},
                {"MRB_QP_LAYOUT": "strips", "MRB_QP_VARIANT": "0:1024"}, {"MRB_QP_LAYOUT": "strips", "MRB_QP_VARIANT": "2:512"}):
        for k, val in env.items():
            monkeypatch.setenv(k, val)
        F, st = ctx.optimize(v, iters=60, return_stats=True)
        assert np.array_equal(F, F_ref), env
        # the residual of the last sweep (max |x_60 - x_59|) is the same number whichever kernel reduced it
        deltas.append(st["last_max_delta"])
        assert 0.0 < deltas[-1] < 1.0 and deltas[-1] == deltas[0], env
        for k in env:
            monkeypatch.delenv(k)


def test_optimize_band_sizes(ctx, O, synth):
    v = synth.sphere(64, 20.0)
    for bs in (2, 8):
        F_ref = O.optimize(v, band_size=bs, iters=20)
        F = ctx.optimize(v, band_size=bs, iters=20)
        assert np.array_equal(F, F_ref), bs


def test_optimize_sign_constraints(ctx, O, synth):
    v = synth.sphere(52, 17.0)
    F = ctx.optimize(v)
    m = O.dfield(v)
    band, _ = O.create_bands(m, 4.0)
    inb = band != 0
    inside = v == 1
    assert np.all(F[inb & inside] >= m[inb & inside])     # x >= lb = margin inside
    assert np.all(F[inb & ~inside] <= -m[inb & ~inside])  # x <= ub = -margin outside
    assert np.all(F[~inb & inside] == 5.0) and np.all(F[~inb & ~inside] == -5.0)


def test_border_band_is_an_error(ctx):
    import mesh_reconstruction_b200 as mrb
    v = np.zeros((20, 20, 20), np.float32)
    v[0:8, 5:15, 5:15] = 1.0      # object touches the border -> tight band on the border
    with pytest.raises(mrb.MrbError) as e:
        ctx.optimize(v)
    assert e.value.code == -4


def test_mcubes_topology(ctx, O, synth):
    for name, v in volumes(synth):
        F = O.optimize(v)
        verts_ref, faces_ref = O.mcubes(F, 0.0)
        verts, faces = ctx.mcubes(F, 0.0)
        assert faces.shape == faces_ref.shape and verts.shape == verts_ref.shape, name
        assert np.array_equal(faces, faces_ref), name
        assert np.array_equal(verts, verts_ref), name


def test_mcubes_binary_volume_isolevel_half(ctx, O, synth):
    # m_cubes.cpp:80: plain MC on a 0/1 grid at isolevel 0.5 (exact corner hits, degenerate filter, hash merges)
    v = synth.box(30, (6, 5, 4), None, 0.0, 0.0)
    verts_ref, faces_ref = O.mcubes(v, 0.5)
    verts, faces = ctx.mcubes(v, 0.5)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)
    # isolevel 0 on a grid with exact zeros: vertices land on grid corners, triangles degenerate
    g = np.zeros((6, 6, 6), np.float32) - 1.0
    g[2:4, 2:4, 2:4] = 1.0
    g[3, 3, 3] = 0.0
    g[1, 2, 2] = 0.0
    verts_ref, faces_ref = O.mcubes(g, 0.0)
    verts, faces = ctx.mcubes(g, 0.0)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)


def test_hash_points(ctx, O):
    rng = np.random.default_rng(11)
    pts = (rng.random((5000, 3)) * np.array([255, 255, 255])).astype(np.float32)
    pts[:100] = np.round(pts[:100])
    assert np.array_equal(ctx.hashXYZ(pts, (256, 256, 256)), O.hash_points(pts, (256, 256, 256)))


def test_reconstruct_end_to_end(ctx, O, synth, tmp_path):
    for name, v in volumes(synth):
        verts_ref, faces_ref = O.mcubes(O.optimize(v), 0.0)
        verts, faces, st = ctx.reconstruct(v, return_stats=True)
        assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref), name
        assert st["iters"] == 500
    # PLY bytes identical to the oracle writer (mcubes.cpp:626-660)
    a, b = tmp_path / "a.ply", tmp_path / "b.ply"
    ctx.write_ply(str(a), verts, faces, (0.01, 0.02, 0.03), (-0.5, 0.25, 1.0), 6)
    O.write_ply(str(b), verts_ref, faces_ref, (0.01, 0.02, 0.03), (-0.5, 0.25, 1.0), 6)
    assert a.read_bytes() == b.read_bytes()


def test_device_pointers(ctx, O, synth):
    """Same entry points with device buffers (torch tensors passed by pointer)."""
    import torch
    v = synth.sphere(52, 17.0)
    vt = torch.from_numpy(v).cuda()
    F = ctx.optimize(vt)
    assert F.is_cuda
    F_ref = O.optimize(v)
    assert np.array_equal(F.cpu().numpy(), F_ref)
    verts, faces = ctx.reconstruct(vt)
    verts_ref, faces_ref = O.mcubes(F_ref, 0.0)
    assert np.array_equal(verts.cpu().numpy(), verts_ref) and np.array_equal(faces.cpu().numpy(), faces_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["bricks", "strips"])
@pytest.mark.parametrize("band_size", [1, 2, 3, 4, 5, 6, 7, 8, 9])
def test_fused_stages_every_band_size(ctx, O, synth, band_size, layout, monkeypatch):
    """The fused stage-1+2 tile kernels are instantiated per band_size (bit-sliced <= 4, byte arrays <= 8, global
    passes above): each against the oracle, on a grid whose dims are not multiples of the 8 x 8 x 32 tile."""
    monkeypatch.setenv("MRB_QP_LAYOUT", layout)
    v = synth.random_spheres((70, 61, 75), 4, 9, 14, seed=100 + band_size, pad=band_size + 3)
    F_ref, info = O.optimize(v, band_size=band_size, iters=12, brute_edt=False, return_info=True)
    F, st = ctx.optimize(v, band_size=band_size, iters=12, return_stats=True)
    assert st["n_band"] == info["n_band"]
    assert np.array_equal(F, F_ref)


@pytest.mark.gpu
def test_sweep_synchronisation_is_repeatable(ctx, synth):
    """The persistent sweep kernel synchronises blocks with flags in global memory: a lost wait would show up as a
    run-to-run difference.  Many short runs (different grid sizes -> different block / neighbour-list shapes)."""
    import os
    os.environ["MRB_QP_LAYOUT"] = "bricks"        # the flag-synchronised kernel, also on the small grids
    try:
        _repeat(ctx, synth)
    finally:
        del os.environ["MRB_QP_LAYOUT"]


def _repeat(ctx, synth):
    for n, r, iters in ((96, 30.0, 40), (160, 60.0, 30), (256, 110.0, 12)):
        v = synth.sphere(n, r)
        first = ctx.optimize(v, iters=iters)
        for _ in range(12 if n < 256 else 4):
            again = ctx.optimize(v, iters=iters)
            assert np.array_equal(first, again)


@pytest.mark.gpu
def test_pool_of_contexts(ctx, synth):
    """Pool = several contexts on one GPU, one host thread each: same meshes as one context, input order kept."""
    import mesh_reconstruction_b200 as mrb
    vols = [synth.random_object(3000 + i) for i in range(7)]
    with mrb.Pool(ctx.device if hasattr(ctx, "device") else 0, workers=3) as pool:
        got = pool.map(vols, iters=60)
    for v, (verts, faces) in zip(vols, got):
        v1, f1 = ctx.reconstruct(v, iters=60)
        assert np.array_equal(faces, f1) and np.array_equal(verts, v1)


def test_cluster_kernel_sizes_and_sweep_parity(ctx, O, synth):
    """k_qp_cluster (small bands, one thread-block cluster): odd and even sweep counts (the result lies in the other buffer),
    one sweep, bands of a few strips, ragged grids; and contexts on several host threads with DIFFERENT sizes at once (the
    shared-memory opt-in of a kernel is process-wide)."""
    import mesh_reconstruction_b200 as mrb
    for v, its in ((synth.sphere(52, 17.0), 1), (synth.sphere(52, 17.0), 7), (synth.sphere(42, 13.0), 500), (synth.sphere(30, 4.0), 33),
                   (synth.ellipsoid((30, 26, 33), (6.0, 4.5, 7.5), (14.5, 12.0, 16.0)), 40), (synth.torus(56, 14.0, 5.0), 64)):
        F, st = ctx.optimize(v, iters=its, return_stats=True)
        assert np.array_equal(F, O.optimize(v, iters=its)), (v.shape, its)
    vols = [synth.sphere(52, 17.0), synth.sphere(30, 4.0), synth.sphere(42, 13.0), synth.torus(56, 14.0, 5.0), synth.sphere(64, 24.0),
            synth.sphere(36, 9.0)] * 3
    refs = [ctx.reconstruct(v, iters=40) for v in vols[:6]]
    with mrb.Pool(0, workers=4) as pool:
        got = pool.map(vols, iters=40)
    for i, (verts, faces) in enumerate(got):
        assert np.array_equal(faces, refs[i % 6][1]) and np.array_equal(verts, refs[i % 6][0]), i


@pytest.mark.gpu
def test_degenerate_volumes_through_the_fused_path(ctx, O, synth):
    """Empty, full, one-voxel and tiny grids: the tile / brick machinery must cope and agree with the oracle."""
    for shape in ((9, 9, 9), (33, 17, 40), (64, 64, 64)):
        for fill in (0.0, 1.0):
            v = np.full(shape, fill, np.float32)
            F_ref = O.optimize(v, iters=5)
            F, st = ctx.optimize(v, iters=5, return_stats=True)
            assert st["n_band"] == 0 and np.array_equal(F, F_ref), (shape, fill)
            verts, faces = ctx.reconstruct(v, iters=5)
            assert len(verts) == 0 and len(faces) == 0
    v = np.zeros((40, 40, 40), np.float32)
    v[20, 20, 20] = 1.0                                     # a single voxel, far from every border
    F_ref = O.optimize(v, iters=30)
    assert np.array_equal(ctx.optimize(v, iters=30), F_ref)
    vr, fr = O.mcubes(F_ref, 0.0)
    verts, faces = ctx.reconstruct(v, iters=30)
    assert np.array_equal(faces, fr) and np.array_equal(verts, vr)
    v = np.zeros((40, 41, 37), np.float32)
    v[12:20, 14:29, 11:17] = 1.0                            # non-cubic grid, box
    for layout in ("bricks", "strips"):
        os.environ["MRB_QP_LAYOUT"] = layout
        try:
            assert np.array_equal(ctx.optimize(v, iters=40), O.optimize(v, iters=40)), layout
        finally:
            del os.environ["MRB_QP_LAYOUT"]
