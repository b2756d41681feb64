"""Front end: binvox completion + observed points -> occupancy volume.

Three-way parity:
  reference's own code (oracle/_ref/libmrb_ref_front.so, compiled unmodified from /root/reference against PCL stand-ins; where
  that is absent, the fixtures tests/golden/ref_front_golden.* it generated)  ==  oracle/frontend_oracle.cpp  ==  the CUDA
  library through the C-ABI (mrb_binvox_to_points, mrb_nn_stats, mrb_voxelize_leaf, mrb_front_end), bit for bit.
What the stand-ins leave unpinned (inside PCL: kd-tree traversal, VoxelGrid's summation order, octree alignment) is listed in
oracle/ref_stubs/pcl/*.h and DESIGN.md section 4."""
import hashlib
import json
import os

import numpy as np
import pytest

from front_cases import scenes, pack_volume

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def golden():
    meta = json.load(open(os.path.join(HERE, "golden", "ref_front_golden.json")))
    arrays = np.load(os.path.join(HERE, "golden", "ref_front_golden.npz"))
    return meta, arrays


@pytest.fixture(scope="module")
def S():
    return scenes()


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, np.float32).view(np.uint32), np.ascontiguousarray(b, np.float32).view(np.uint32))


# ---- CPU: the scenes are what the fixtures were made from; the oracle reproduces the reference's outputs -----------------------
def test_scenes_match_fixture_inputs(S, golden):
    meta, arrays = golden
    assert set(S) == set(meta)
    for name, (vox, dwh, xf, obs) in S.items():
        assert np.array_equal(arrays[name + "/voxels"], vox), name
        assert bits_equal(arrays[name + "/observed"], obs), name
        assert list(dwh) == meta[name]["dwh"]


def check_against_fixture(impl_points, impl_stats, impl_front, name, scene, m, arrays):
    vox, dwh, xf, obs = scene
    for res in (1, 2, 3):
        p = impl_points(vox, dwh, xf, res)
        assert len(p) == m["points"][str(res)]["n"], (name, res)
        assert sha(p) == m["points"][str(res)]["sha256"], (name, res)
        if res == 2:
            dt, cnt = impl_stats(p)
            assert np.float32(cnt) / np.float32(dt) == np.float32(m["density_res2"]), name          # getDensity
            assert np.float32(2.0 * float(dt) / float(np.float32(cnt))) == np.float32(m["resolution_res2"]), name   # getResolution
    dt, cnt = impl_stats(obs)
    assert np.float32(cnt) / np.float32(dt) == np.float32(m["observed_density"]), name
    fe = impl_front(vox, dwh, xf, obs)
    assert fe["res_factor"] == m["res_factor"] and fe["n_predicted"] == m["n_predicted"] and fe["n_cloud"] == m["n_cloud"], name
    assert np.float32(fe["leaf"]) == np.float32(m["leaf"]), name
    assert list(fe["dims"]) == m["dims"] and fe["pad"] == m["pad"], name
    assert np.array_equal(np.asarray(fe["scale"], np.float32), np.asarray(m["scale"], np.float32)), name
    assert np.array_equal(np.asarray(fe["shift"], np.float32), np.asarray(m["shift"], np.float32)), name
    vol = np.asarray(fe["volume"])
    assert set(np.unique(vol)) <= {0.0, 1.0}
    assert np.array_equal(pack_volume(vol), arrays[name + "/volume_bits"]), name
    assert int(vol.sum()) == m["occupied"], name


def test_oracle_reproduces_the_reference_fixtures(O, S, golden):
    meta, arrays = golden
    for name, scene in S.items():
        check_against_fixture(O.binvox_to_points, O.nn_stats, O.front_end, name, scene, meta[name], arrays)


def test_oracle_against_the_live_reference(O, tmp_path):
    """Fresh scenes (not the fixtures') through the reference's compiled code, where /root/reference or a shipped oracle/_ref exists."""
    from oracle import ref as R
    from mesh_reconstruction_b200 import binvox, synth
    if not R.front_available():
        pytest.skip("oracle/_ref/libmrb_ref_front.so not built (no /root/reference here)")
    for i, (kind, n, zs, step) in enumerate((("blobs", 21, 1.2, 0.4), ("box", 17, 2.0, 0.3), ("sphere", 19, 0.0, 0.9))):
        vox, dwh, xf, obs = synth.completion_scene(n, kind, seed=40 + i, z_shift=zs, observed_step=step)
        path = str(tmp_path / ("s%d.binvox" % i))
        binvox.write(path, vox, dwh, xf)
        rv, rdwh, rxf = R.read_binvox(path)
        bv, bdwh, bxf = binvox.read(path)                               # the product's reader against read_binvox
        assert np.array_equal(rv, vox) and np.array_equal(bv, vox) and tuple(rdwh) == tuple(dwh) == tuple(bdwh)
        assert np.array_equal(np.asarray(rxf, np.float32), np.asarray(bxf, np.float32))
        for res in (1, 2, 4):
            assert bits_equal(R.binvox_to_points(path, res), O.binvox_to_points(vox, dwh, xf, res)), (kind, res)
        p = O.binvox_to_points(vox, dwh, xf, 2)
        dt, cnt = O.nn_stats(p)
        assert R.density(p) == np.float32(cnt) / dt
        assert R.resolution(p) == np.float32(2.0 * float(dt) / float(np.float32(cnt)))
        a, b = R.front_end(path, obs), O.front_end(vox, dwh, xf, obs)
        assert a["res_factor"] == b["res_factor"] and a["n_cloud"] == b["n_cloud"] and a["leaf"] == b["leaf"]
        assert a["dims"] == b["dims"] and np.array_equal(a["scale"], b["scale"]) and np.array_equal(a["shift"], b["shift"])
        assert np.array_equal(a["volume"], b["volume"])
        # createGrid with other gap widths on the merged cloud (fillGrid alone)
        allp = np.concatenate([obs, O.binvox_to_points(vox, dwh, xf, b["res_factor"])])
        for rf in (1, 3, 6):
            ra, oa = R.volume_of_cloud(allp, rf), O.voxelize_leaf(allp, b["leaf"], rf)
            assert ra["leaf"] == b["leaf"] and np.array_equal(ra["volume"], oa["volume"]), (kind, rf)


def test_fill_grid_closes_only_the_first_narrow_gap(O):
    """grid.cpp:213-265 on hand-made columns: one point per leaf, leaf size 1."""
    def column(ks, res):
        pts = np.array([[0.5, 0.5, k + 0.5] for k in ks] + [[3.5, 0.5, 0.5]], np.float32)
        v = O.voxelize_leaf(pts, 1.0, res)["volume"]
        return v[6, 6, 6:-6].astype(int).tolist()
    assert column([0, 1, 4, 5, 8], 2) == [1, 1, 1, 1, 1, 1, 0, 0, 1]          # gap of 2 < 3 closed, the second gap never
    assert column([0, 1, 4, 5, 8], 1) == [1, 1, 0, 0, 1, 1, 0, 0, 1]          # gap of 2 >= 2 stays
    assert column([0, 1, 4, 5, 8], 0) == [1, 1, 0, 0, 1, 1, 0, 0, 1]          # res_factor 0: no fillGrid at all
    assert column([2, 9], 7) == [0, 0, 1, 1, 1, 1, 1, 1, 1, 1]                # gap of 6 < 8


# ---- GPU: the CUDA library against the oracle and the fixtures -------------------------------------------------------
@pytest.mark.gpu
def test_gpu_reproduces_the_reference_fixtures(ctx, S, golden):
    meta, arrays = golden
    for name, scene in S.items():
        check_against_fixture(ctx.binvox_to_points, ctx.nn_stats, ctx.front_end, name, scene, meta[name], arrays)


@pytest.mark.gpu
def test_gpu_front_end_pieces_equal_the_oracle(ctx, O, synth):
    rng = np.random.default_rng(7)
    for i, (kind, n, zs, step) in enumerate((("blobs", 33, 1.0, 0.3), ("box", 26, 2.5, 0.25), ("sphere", 40, 0.0, 0.5))):
        vox, dwh, xf, obs = synth.completion_scene(n, kind, seed=60 + i, z_shift=zs, observed_step=step)
        for res in (1, 2, 3, 5):
            assert bits_equal(ctx.binvox_to_points(vox, dwh, xf, res), O.binvox_to_points(vox, dwh, xf, res)), (kind, res)
        p = O.binvox_to_points(vox, dwh, xf, 3)
        allp = np.concatenate([obs, p])
        for cloud in (obs, p, allp, allp[rng.permutation(len(allp))], allp[:7], allp[:11]):
            a, b = ctx.nn_stats(cloud), O.nn_stats(cloud)
            assert a[1] == b[1] and a[0] == b[0], (kind, len(cloud))
        dt, cnt = O.nn_stats(allp)
        leaf = np.float32(2.0 * float(dt) / float(np.float32(cnt)))
        for lf, rf in ((leaf, 3), (leaf, 0), (np.float32(leaf * 0.6), 1), (np.float32(leaf * 2.5), 8)):
            a, b = ctx.voxelize_leaf(allp, lf, rf), O.voxelize_leaf(allp, lf, rf)
            assert a["dims"] == b["dims"] and a["pad"] == b["pad"] and a["occupied_leaves"] == b["occupied_leaves"], (kind, lf, rf)
            assert np.array_equal(a["scale"], b["scale"]) and np.array_equal(a["shift"], b["shift"]), (kind, lf, rf)
            assert np.array_equal(a["volume"], b["volume"]), (kind, lf, rf)
        a, b = ctx.front_end(vox, dwh, xf, obs), O.front_end(vox, dwh, xf, obs)
        for k in ("res_factor", "n_predicted", "n_cloud", "observed_count", "completed_count", "leaf", "observed_dist_total",
                  "completed_dist_total", "observed_density", "completed_density", "dims"):
            assert a[k] == b[k], (kind, k, a[k], b[k])
        assert np.array_equal(a["volume"], b["volume"]) and np.array_equal(a["scale"], b["scale"]) and np.array_equal(a["shift"], b["shift"])


@pytest.mark.gpu
def test_gpu_nn_stats_outliers_and_clusters(ctx, O):
    """Queries far outside the reference points' box, two distant clusters, duplicate points, a flat cloud."""
    rng = np.random.default_rng(11)
    a = rng.normal(0, 1, (3000, 3)).astype(np.float32)
    b = (rng.normal(0, 0.01, (2000, 3)) + [500.0, -300.0, 40.0]).astype(np.float32)
    far = np.concatenate([a, b])
    far[0] = [1e4, 1e4, -1e4]                       # query 0 is an outlier
    far[10] = [-3e6, 0, 0]                          # and one absurdly far away
    flat = a.copy(); flat[:, 2] = 0.25
    dup = np.repeat(a[:500], 3, axis=0)
    for name, cloud in (("far", far), ("flat", flat), ("dup", dup), ("tiny", a[:3]), ("one", a[:1])):
        g, o = ctx.nn_stats(cloud), O.nn_stats(cloud)
        assert g[1] == o[1] and g[0] == o[0], (name, g, o)


@pytest.mark.gpu
def test_gpu_front_end_to_mesh(ctx, O, synth, tmp_path):
    """The whole CLI pipeline (mesh_reconstruction.cpp:99-134): front end on the GPU, volume kept on the device, reconstruct,
    PLY -- equal to the oracle's composition byte for byte."""
    import torch
    vox, dwh, xf, obs = synth.completion_scene(36, "blobs", seed=77, z_shift=1.0, observed_step=0.3)
    fe = ctx.front_end(vox, dwh, xf, obs, device=True)
    assert isinstance(fe["volume"], torch.Tensor) and fe["volume"].is_cuda
    ref = O.front_end(vox, dwh, xf, obs)
    assert np.array_equal(fe["volume"].cpu().numpy(), ref["volume"])
    verts, faces = ctx.reconstruct(fe["volume"])
    F = O.optimize(ref["volume"])
    verts_ref, faces_ref = O.mcubes(F, 0.0)
    assert np.array_equal(faces.cpu().numpy(), faces_ref) and np.array_equal(verts.cpu().numpy(), verts_ref)
    p1, p2 = str(tmp_path / "a.ply"), str(tmp_path / "b.ply")
    ctx.write_ply(p1, verts, faces, fe["scale"], fe["shift"], fe["pad"])
    O.write_ply(p2, verts_ref, faces_ref, ref["scale"], ref["shift"], ref["pad"])
    assert open(p1, "rb").read() == open(p2, "rb").read()


@pytest.mark.gpu
def test_gpu_front_end_argument_errors(ctx):
    import mesh_reconstruction_b200 as mrb
    v = np.zeros(27, np.uint8)
    with pytest.raises(mrb.MrbError):
        ctx.binvox_to_points(v, (3, 3, 3), (1, 0, 0, 0), 0)              # "Resolution factor must be at least 1"
    assert ctx.binvox_to_points(v, (3, 3, 3), (1, 0, 0, 0), 2).shape == (0, 3)      # an empty completion: no points, no error
    with pytest.raises(mrb.MrbError):
        ctx.voxelize_leaf(np.zeros((4, 3), np.float32), 0.0, 1)


@pytest.mark.gpu
def test_gpu_voxel_grid_on_random_clouds(ctx, O):
    """Negative coordinates, large offsets (few mantissa bits left for the leaf index), planar and clustered clouds, duplicates,
    leaves with hundreds of points (float summation order inside a leaf), res_factor 0..5."""
    rng = np.random.default_rng(21)
    def clouds():
        yield "uniform_neg", rng.uniform(-3.0, -1.0, (20000, 3)), 0.05
        yield "offset_1e3", rng.uniform(0.0, 1.0, (15000, 3)) + [1000.0, -2000.0, 500.0], 0.04
        yield "planar", np.stack([rng.uniform(0, 2, 12000), rng.uniform(0, 1, 12000), np.full(12000, 0.3)], 1), 0.03
        yield "clusters", np.concatenate([rng.normal(c, 0.02, (4000, 3)) for c in ([0, 0, 0], [0.5, 0.1, -0.4], [0.2, 0.9, 0.3])]), 0.015
        yield "dense_leaves", rng.uniform(0.0, 0.3, (30000, 3)), 0.1
        yield "duplicates", np.repeat(rng.uniform(0, 1, (3000, 3)), 4, axis=0), 0.06
        g = np.arange(0, 24, dtype=np.float64) * 0.125                       # points exactly on leaf boundaries
        yield "lattice", np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3), 0.25
    for name, pts, leaf in clouds():
        pts = np.ascontiguousarray(pts, np.float32)
        for rf in (0, 1, 5):
            a, b = ctx.voxelize_leaf(pts, leaf, rf), O.voxelize_leaf(pts, leaf, rf)
            assert a["dims"] == b["dims"] and a["occupied_leaves"] == b["occupied_leaves"], (name, rf)
            assert np.array_equal(a["scale"], b["scale"]) and np.array_equal(a["shift"], b["shift"]), (name, rf, a["scale"], b["scale"])
            assert np.array_equal(a["volume"], b["volume"]), (name, rf)
        ga, oa = ctx.nn_stats(pts), O.nn_stats(pts)
        assert ga[1] == oa[1] and ga[0] == oa[0], (name, ga, oa)
