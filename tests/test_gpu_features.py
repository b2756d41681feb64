"""Feature pre-pass (--feature-detection): surface, normals, feature map and the QP feature reset, GPU vs oracle.
The eigen-solve is our own deterministic definition (include/mrb_eig3.h): bit-exact GPU == oracle is required;
against the reference (Eigen) this sub-path is parity-unpinned."""
import numpy as np
import pytest

from conftest import random_blobs

pytestmark = pytest.mark.gpu


def shapes(synth):
    yield "box", synth.box(48, (12, 9, 7), None, 0.0, 0.0)
    yield "box_rot", synth.box(56, (13, 10, 8), None, 20.0, 35.0)
    yield "blobs", random_blobs((44, 48, 52), seed=9)
    yield "cfg2_small", synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0))


def test_surface_normals_feature_map(ctx, O, synth):
    for name, v in shapes(synth):
        s_ref = O.get_surface(v)
        s = ctx.getSurface(v)
        assert np.array_equal(s, s_ref), name
        n_ref = O.surface_normals(v, s_ref)
        n = ctx.getSurfaceNormals(v, s)
        assert np.array_equal(n, n_ref), name
        assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-6)
        smap = O.index_map(v.shape, s_ref)
        assert np.array_equal(ctx.getIndexMap(v.shape, s), smap), name
        for thr in (0.75, 0.9):
            assert np.array_equal(ctx.getFeatureMap(v, smap, n, thr), O.feature_map(v, smap, n_ref, thr)), (name, thr)


def test_optimize_with_feature_reset(ctx, O, synth):
    """optimize(..., USING_FEATURES=true): band voxels on the feature map are reset to 0.2 (quadprog.cpp:92-101)."""
    for name, v in shapes(synth):
        s = O.get_surface(v)
        n = O.surface_normals(v, s)
        fm = O.feature_map(v, O.index_map(v.shape, s), n, 0.75)
        F_ref = O.optimize(v, fm, using_features=True, iters=40)
        F = ctx.optimize(v, fm, True, iters=40)
        assert np.array_equal(F, F_ref), name
        if fm.sum() > 0:
            assert np.all(F[fm == 1] == np.float32(0.2)), name     # all feature voxels lie in the band here


def test_border_volume_is_an_error(ctx):
    import mesh_reconstruction_b200 as mrb
    v = np.zeros((10, 10, 10), np.float32)
    v[0:4, 3:7, 3:7] = 1.0
    with pytest.raises(mrb.MrbError):
        ctx.getSurface(v)


def test_feature_marching_cubes(ctx, O, synth):
    """mcubes with USING_FEATURES: feature points, fans, edge flips -- GPU == oracle, topology and coordinates."""
    for name, v in shapes(synth):
        s = O.get_surface(v)
        n = O.surface_normals(v, s)
        smap = O.index_map(v.shape, s)
        fm = O.feature_map(v, smap, n, 0.75)
        F = O.optimize(v, fm, using_features=True)
        verts_ref, faces_ref = O.mcubes_features(F, smap, n, 0.0, 0.75, 0.8)
        verts, faces = ctx.mcubes_features(F, smap, n, 0.0, 0.75, 0.8)
        assert faces.shape == faces_ref.shape and verts.shape == verts_ref.shape, name
        assert np.array_equal(faces, faces_ref), name
        assert np.array_equal(verts, verts_ref), name
        # and it really differs from plain marching cubes on shapes with sharp edges
        if name.startswith("box"):
            v0, f0 = O.mcubes(F, 0.0)
            assert len(faces) != len(f0)


def test_reconstruct_with_feature_detection(ctx, O, synth):
    """The whole --feature-detection pipeline in one call (mesh_reconstruction.cpp:123-134) == the oracle's composition."""
    v = synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0))
    s = O.get_surface(v)
    n = O.surface_normals(v, s)
    smap = O.index_map(v.shape, s)
    fm = O.feature_map(v, smap, n, 0.75)
    F = O.optimize(v, fm, using_features=True)
    verts_ref, faces_ref = O.mcubes_features(F, smap, n, 0.0, 0.75, 0.8)
    verts, faces = ctx.reconstruct(v, using_features=True)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)


def test_config2_feature_detection_172(ctx, O, synth):
    """BASELINE config 2 as stated: 172^3 rotated box + sphere WITH --feature-detection (thresholds 0.75 / 0.8), the
    whole pipeline of mesh_reconstruction.cpp:123-134 in one call == the oracle's composition, bit for bit."""
    v = synth.config_volume(2)
    assert v.shape == (172, 172, 172)
    s = O.get_surface(v)
    n = O.surface_normals(v, s)
    smap = O.index_map(v.shape, s)
    fm = O.feature_map(v, smap, n, 0.75)
    assert fm.sum() > 1000                                             # the box edges are detected
    F = O.optimize(v, fm, using_features=True, brute_edt=False)        # (fast exact EDT == brute force, test_oracle_dfield)
    verts_ref, faces_ref = O.mcubes_features(F, smap, n, 0.0, 0.75, 0.8)
    assert np.array_equal(ctx.getSurface(v), s) and np.array_equal(ctx.getSurfaceNormals(v, s), n)
    assert np.array_equal(ctx.getFeatureMap(v, smap, n, 0.75), fm)
    assert np.array_equal(ctx.optimize(v, fm, True), F)
    verts, faces = ctx.reconstruct(v, using_features=True, feature_threshold=0.75, corner_threshold=0.8)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)
    v0, f0 = ctx.reconstruct(v)                                        # and the default path differs (features are real here)
    assert len(f0) != len(faces)


def test_repeated_feature_calls_on_one_context(synth, O):
    """Host-buffer feature calls that grow the pinned staging must not invalidate the side buffers (regression:
    pinned_reserve once freed them): small mesh, then larger meshes, then the first again, all on ONE context."""
    import mesh_reconstruction_b200 as mrb
    shapes_ = [synth.box(40, (8, 6, 5), None, 0.0, 0.0), synth.box(56, (13, 10, 8), None, 20.0, 35.0),
               synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0))]
    refs = []
    for v in shapes_:
        s = O.get_surface(v)
        n = O.surface_normals(v, s)
        smap = O.index_map(v.shape, s)
        F = O.optimize(v, O.feature_map(v, smap, n, 0.75), using_features=True)
        refs.append(O.mcubes_features(F, smap, n, 0.0, 0.75, 0.8))
    with mrb.Context(0) as c:
        for i in (0, 1, 2, 2, 1, 0, 2):
            verts, faces = c.reconstruct(shapes_[i], using_features=True)
            assert np.array_equal(faces, refs[i][1]) and np.array_equal(verts, refs[i][0]), i
