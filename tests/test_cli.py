"""mesh_reconstruction CLI: argument handling (CPU), the front end and the full run (GPU)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMPAT = os.path.join(ROOT, "mesh_reconstruction_b200", "compat")
CLI = os.path.join(ROOT, "mesh_reconstruction_b200", "cli", "mesh_reconstruction")


@pytest.fixture(scope="module")
def cli():
    subprocess.check_call(["make", "-C", COMPAT, "-s"])
    return CLI


def read_volume(path):
    raw = np.fromfile(path, dtype=np.uint8)
    dims = raw[:12].view(np.int32)
    return raw[12:].view(np.float32).reshape(tuple(dims))


def test_usage_and_exit_codes(cli):
    r = subprocess.run([cli], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage: mesh_reconstruction <binvox file> <pcd file> <output file> <options>" in r.stdout
    r = subprocess.run([cli, "a", "b", "c", "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--feature-detection" in r.stdout and "--corner-threshold" in r.stdout
    r = subprocess.run([cli, "a", "b", "c", "--bogus"], capture_output=True, text=True)
    assert r.returncode == 1


def read_pcd_ascii(path):
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith("DATA")) + 1
    return np.array([l.split()[:3] for l in lines[start:] if l.strip()], dtype=np.float64).astype(np.float32)


@pytest.mark.gpu
def test_front_end_volume(cli, synth, O, tmp_path):
    """The CLI's front end runs on the GPU (mrb_front_end); its occupancy grid equals the CPU oracle's on the same files."""
    from mesh_reconstruction_b200 import binvox
    bv, pcd = synth.cli_inputs(str(tmp_path))
    vol_path = tmp_path / "vol.bin"
    r = subprocess.run([cli, bv, pcd, str(tmp_path / "out.ply"), "--feature-threshold", "0.7", "--dump-volume", str(vol_path)],
                       capture_output=True, text=True)
    assert r.returncode == 1, r.stdout + r.stderr
    for line in ("Not using feature detection", "Not using CUDA", "Feature threshold is 0.7", "Corner threshold is 0.8",
                 "reading binvox version 1", "Resolution Factor:", "Leaf size is"):
        assert line in r.stdout, line
    v = read_volume(vol_path)
    assert set(np.unique(v)) <= {0.0, 1.0}
    vox, dwh, xf = binvox.read(bv)
    ref = O.front_end(vox, dwh, xf, read_pcd_ascii(pcd))
    assert ("Resolution Factor: %d" % ref["res_factor"]) in r.stdout
    assert ("PointCloud before filtering: %d data points" % ref["n_cloud"]) in r.stdout
    assert ("PointCloud after filtering: %d data points" % ref["occupied_leaves"]) in r.stdout
    assert v.shape == ref["volume"].shape and np.array_equal(v, ref["volume"])
    occ = v == 1
    assert occ.sum() > 1000
    # pad = 6 empty voxels on every side (grid.cpp:91-114)
    assert not occ[:6].any() and not occ[-6:].any() and not occ[:, :6].any() and not occ[:, -6:].any()
    assert not occ[:, :, :6].any() and not occ[:, :, -6:].any()
    # the merged cloud is a ball: its occupancy fills roughly pi/6 of the padded-out bounding box
    inner = occ[6:-6, 6:-6, 6:-6]
    assert 0.35 < inner.mean() < 0.7


@pytest.mark.gpu
def test_cli_end_to_end(cli, synth, ctx, tmp_path):
    bv, pcd = synth.cli_inputs(str(tmp_path))
    out = tmp_path / "out.ply"
    r = subprocess.run([cli, bv, pcd, str(out), "--cuda"], capture_output=True, text=True)
    assert r.returncode == 1, r.stdout + r.stderr           # the reference returns 1 on success (mesh_reconstruction.cpp:136)
    assert "Using CUDA" in r.stdout and "quadratic program finished" in r.stdout
    lines = out.read_text().splitlines()
    assert lines[0] == "ply" and lines[1] == "format ascii 1.0"
    nv = int(lines[2].split()[-1]); nf = int(lines[6].split()[-1])
    assert nv > 1000 and nf > 2000
    faces = np.array([l.split()[1:] for l in lines[9 + nv:9 + nv + nf]], dtype=np.int64)
    e = np.sort(np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]), axis=1)
    uniq, cnt = np.unique(e, axis=0, return_counts=True)
    # a voxelised ball: essentially closed (the reference's ambiguous-case table and its zero-snapping can leave a
    # few non-manifold edges on rough input, so this is not required to be exact)
    assert (cnt == 2).mean() > 0.99
    # same mesh as the Python host side on the same occupancy grid
    vol_path = tmp_path / "vol.bin"
    subprocess.run([cli, bv, pcd, str(out) + ".x", "--dump-volume", str(vol_path)], capture_output=True, text=True)
    verts, f2 = ctx.reconstruct(read_volume(vol_path))
    assert len(verts) == nv and np.array_equal(f2, faces)


@pytest.mark.gpu
def test_cli_feature_detection(cli, synth, tmp_path):
    bv, pcd = synth.cli_inputs(str(tmp_path))
    out = tmp_path / "feat.ply"
    r = subprocess.run([cli, bv, pcd, str(out), "--feature-detection", "--cuda", "--feature-threshold", "0.8"],
                       capture_output=True, text=True)
    assert r.returncode == 1, r.stdout + r.stderr
    assert "Using feature detection" in r.stdout and "Feature threshold is 0.8" in r.stdout
    lines = out.read_text().splitlines()
    assert lines[0] == "ply" and int(lines[2].split()[-1]) > 1000 and int(lines[6].split()[-1]) > 2000


def test_m_cubes_usage(cli):
    exe = os.path.join(os.path.dirname(cli), "m_cubes")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage: <pcd filename> <output filename (.ply)> [optional]leaf_size" in r.stderr


@pytest.mark.gpu
def test_m_cubes_meshes_a_point_cloud(cli, ctx, synth, tmp_path):
    """m_cubes (m_cubes.cpp:18-81): voxelise a pcd at the given leaf size, marching cubes at isolevel 0.5.  The mesh must be
    what mrb_mcubes gives on the same occupancy grid (rebuilt here from the same points with numpy)."""
    exe = os.path.join(os.path.dirname(cli), "m_cubes")
    rng = np.random.default_rng(5)
    pts = rng.normal(size=(40000, 3))
    pts = pts / np.linalg.norm(pts, axis=1, keepdims=True) * rng.uniform(0.0, 1.0, (40000, 1)) ** (1 / 3) * 0.5
    pts = (np.round(pts * 128) / 128).astype(np.float32)       # k / 128: exact in fp32 and in the "%.7g" text of the pcd
    pcd = str(tmp_path / "ball.pcd")
    synth.write_pcd_ascii(pcd, pts)
    out = tmp_path / "ball.ply"
    leaf = 0.02
    r = subprocess.run([exe, pcd, str(out), str(leaf)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Running Standard Marching Cubes" in r.stdout and "Leaf size is 0.02" in r.stdout
    lines = out.read_text().splitlines()
    nv, nf = int(lines[2].split()[-1]), int(lines[6].split()[-1])
    # the same occupancy grid: one voxel per occupied leaf of size 2 * leaf, pad 6 (voxelize.cpp / grid.cpp)
    inv = np.float32(1.0) / np.float32(2 * leaf)
    ijk = np.floor(pts * inv).astype(np.int64)
    mn = np.floor(pts.min(axis=0) * inv).astype(np.int64)
    div = np.floor(pts.max(axis=0) * inv).astype(np.int64) - mn + 1
    vol = np.zeros(tuple(div + 12), np.float32)
    q = ijk - mn + 6
    vol[q[:, 0], q[:, 1], q[:, 2]] = 1.0
    verts, faces = ctx.mcubes(vol, 0.5)
    assert (nv, nf) == (len(verts), len(faces)) and nf > 1000
