"""Pins the oracle's stage 4 (CPU only): hand-derived cells, hash known answers, mesh invariants, PLY bytes."""
import numpy as np

from conftest import random_blobs


def euler(verts, faces):
    E = set()
    for f in faces:
        for a, b in ((f[0], f[1]), (f[1], f[2]), (f[2], f[0])):
            E.add((min(a, b), max(a, b)))
    return len(verts) - len(E) + len(faces), E


def test_single_corner_cell(O):
    # one cell, corner 0 below the isolevel: case 1 -> triangle on edges (0, 8, 3) (Bourke table row 1)
    F = np.ones((2, 2, 2), np.float32)
    F[0, 0, 0] = -1.0
    verts, faces = O.mcubes(F, 0.0)
    # edge 0: corners 0-1 along +x; edge 8: corners 0-4 along +z; edge 3: corners 3-0 along y
    assert faces.tolist() == [[0, 1, 2]]
    assert verts.tolist() == [[0.5, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.5, 0.0]]


def test_vertex_interp_endpoint_swap(O):
    # both cells sharing an edge must produce bit-identical vertices (mcubes.cpp:370-378)
    F = np.ones((3, 2, 2), np.float32)
    F[1, :, :] = -0.3
    verts, faces = O.mcubes(F, 0.0)
    xs = sorted(set(np.round(verts[:, 0], 6).tolist()))
    assert len(xs) == 2 and len(verts) == 8          # 4 edge crossings per side, all shared
    assert len(faces) == 4                           # one quad (2 triangles) per cell
    # x = 1 + (-0.3 / -1.3) and 1 + (1 - 1/1.3): evaluated from the smaller endpoint in both cells
    f32 = np.float32
    lo = f32(0.0) + f32(f32(f32(1.0) / f32(f32(-0.3) - f32(1.0))) * f32(f32(0.0) - f32(1.0)))
    hi = f32(1.0) + f32(f32(f32(1.0) / f32(f32(1.0) - f32(-0.3))) * f32(f32(0.0) - f32(-0.3)))
    assert sorted(set(verts[:, 0].tolist())) == sorted([float(lo), float(hi)])


def test_hash_known_answer(O):
    dims = (7, 11, 13)
    pts = np.array([[1.5, 2.25, 3.125], [0.0, 0.0, 0.0], [6.0, 10.0, 12.0]], np.float32)
    keys = O.hash_points(pts, dims)
    exp = []
    for x, y, z in pts.astype(np.float64):
        k = int(x * 1000000000.0 * float(11 * 13)) + int(y * 1000000.0 * 13.0) + int(z * 1000.0)
        exp.append(k)
    assert keys.tolist() == exp


def test_closed_manifold_and_first_appearance(O, synth):
    for v, chi in ((synth.sphere(40, 11.0), 2), (synth.torus(48, 12.0, 4.5), 0)):
        F = O.optimize(v)
        verts, faces = O.mcubes(F, 0.0)
        e, E = euler(verts, faces)
        assert e == chi
        # closed 2-manifold: every edge is shared by exactly two triangles
        cnt = {}
        for f in faces:
            for a, b in ((f[0], f[1]), (f[1], f[2]), (f[2], f[0])):
                k = (min(a, b), max(a, b)); cnt[k] = cnt.get(k, 0) + 1
        assert set(cnt.values()) == {2}
        # ids are handed out in first-appearance order (hasher.cpp:13-23)
        flat = faces.reshape(-1)
        first_seen = {}
        for i, vid in enumerate(flat):
            first_seen.setdefault(int(vid), i)
        order = [vid for vid, _ in sorted(first_seen.items(), key=lambda kv: kv[1])]
        assert order == list(range(len(verts)))


def test_degenerate_triangles_dropped_and_hash_merge(O):
    # exact zeros put vertices on grid corners: zero-length sides are filtered (mcubes.cpp:227-229)
    g = -np.ones((5, 5, 5), np.float32)
    g[2, 2, 2] = 0.0
    verts, faces = O.mcubes(g, 0.0)
    assert len(faces) == 0 and len(verts) == 0
    g[2, 2, 2] = 1.0
    verts, faces = O.mcubes(g, 0.0)
    assert len(verts) == 6 and len(faces) == 8      # octahedron around the single positive voxel
    assert euler(verts, faces)[0] == 2


def test_ply_bytes(O, tmp_path):
    verts = np.array([[6.0, 6.0, 6.0], [7.5, 6.0, 6.25], [6.0, 1234568.0, 6.0]], np.float32)
    faces = np.array([[0, 1, 2]], np.int32)
    p = tmp_path / "t.ply"
    O.write_ply(str(p), verts, faces, (0.5, 0.5, 0.5), (1.0, 2.0, 3.0), 6)
    expect = ("ply\nformat ascii 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
              "element face 1\nproperty list uchar int vertex_indices\nend_header\n"
              "1 2 3\n1.75 2 3.125\n1 617283 3\n3 0 1 2\n")
    assert p.read_text() == expect
