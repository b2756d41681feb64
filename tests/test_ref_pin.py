"""Pins against the REFERENCE'S OWN CODE.

oracle/build_ref.sh compiles the reference's hot-path translation units unmodified from /root/reference (stand-in
headers for the absent Eigen / Boost / PCL) into oracle/_ref/libmrb_ref.so.  Three layers:
  * committed fixtures generated from it (tests/golden/ref_golden.*, made by tests/golden/make_ref_golden.py): the restated
    oracle (CPU) and the CUDA library (GPU) must reproduce them bit for bit -- these run everywhere;
  * the live library, where it exists (this container; the GPU box when oracle/_ref/ travelled with the snapshot):
    oracle == reference and GPU == reference on more shapes and the edge cases;
  * what the build cannot pin -- Eigen's eigen-solver / SVD arithmetic -- is bounded instead: include/mrb_eig3.h
    against LAPACK (numpy) and against the double-precision stand-in build.
Everything is bit-exact (floats compared through their bit patterns, so -0.0 != +0.0) unless a tolerance is written.
"""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import random_blobs

HERE = os.path.dirname(os.path.abspath(__file__))
META = json.load(open(os.path.join(HERE, "golden", "ref_golden.json")))
DATA = np.load(os.path.join(HERE, "golden", "ref_golden.npz"))
NAMES = sorted(k for k in META if not k.startswith("_"))
FRAME = META["_ply_frame"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def same(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return a.shape == b.shape and np.array_equal(bits(a), bits(b))


def volume(name, synth):
    if name + "_volume" in DATA:
        return DATA[name + "_volume"]
    return {
        "config1_sphere52": lambda: synth.sphere(52, 17.0),
        "config1_spheres_seed1": lambda: synth.random_spheres(52, 3, 6, 12, 1),
        "box_rot56": lambda: synth.box(56, (13, 10, 8), None, 20.0, 35.0),
        "cfg2_small64": lambda: synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0)),
    }[name]()


@pytest.fixture(scope="module")
def R():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built (no /root/reference here); the committed fixtures still run")
    return ref


def check_against_fixture(name, m, r):
    """r: dict of arrays produced by some implementation; every key present must hash to the reference's value."""
    for k, a in r.items():
        assert sha(a) == m["sha_" + k], "%s: %s differs from the reference fixture" % (name, k)
        if name + "_" + k in DATA:
            assert same(a, DATA[name + "_" + k]), (name, k)


def ply_bytes(write, verts, faces, tmp_path):
    p = str(tmp_path / "m.ply")
    write(p, verts, faces, scale=tuple(FRAME["scale"]), shift=tuple(FRAME["shift"]), pad=FRAME["pad"])
    return np.frombuffer(open(p, "rb").read(), np.uint8)


# ---- CPU: the restated oracle reproduces the reference's outputs -------------------------------------------------
@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_the_reference_fixtures(O, synth, tmp_path, name):
    m = META[name]
    v = volume(name, synth)
    sq = O.sqdist(O.fast_perim(v))
    margin = O.sqrt(sq)
    band, tight = O.create_bands(margin, 4.0)
    q = O.prime_qp(v, margin, band, tight)
    F = O.optimize(v)
    verts, faces = O.mcubes(F, 0.0)
    r = dict(sq=sq.astype(np.int32), margin=margin, band_idx=np.asarray(O.find_indexes(band), np.int32), F=F, verts=verts,
             faces=np.asarray(faces, np.int32), ply=ply_bytes(O.write_ply, verts, faces, tmp_path),
             **{k: q[k] for k in ("lb", "ub", "x0", "invdg", "jc", "ir", "pr")})
    assert int((tight != 0).sum()) == m["n_tight"]
    if m["features"]:
        s = O.get_surface(v)
        n = O.surface_normals(v, s)
        smap = O.index_map(v.shape, s)
        fm = O.feature_map(v, smap, n, 0.75)
        Ff = O.optimize(v, fm, using_features=True)
        vf, ff = O.mcubes_features(Ff, smap, n, 0.0, 0.75, 0.8)
        r.update(surface=np.asarray(s, np.int32), normals=n, feature_map=fm, F_feat=Ff, verts_feat=vf,
                 faces_feat=np.asarray(ff, np.int32))
    check_against_fixture(name, m, r)


def test_fixtures_come_from_the_reference_build(R, synth):
    """Provenance: the live reference library still produces the committed vectors."""
    assert "reference sources" in META["_provenance"]
    for name in ("cube5", "config1_spheres_seed1"):
        v = volume(name, synth)
        F = R.optimize(v)
        verts, faces = R.mcubes(F, 0.0)
        check_against_fixture(name, META[name], dict(F=F, verts=verts, faces=np.asarray(faces, np.int32)))


def edge_volumes(synth):
    z = np.zeros((9, 8, 7), np.float32)
    yield "empty", z
    yield "full", np.ones((7, 9, 8), np.float32)
    one = z.copy(); one[4, 4, 3] = 1.0
    yield "single_voxel", one
    yield "touching_border", np.pad(np.ones((3, 3, 3), np.float32), ((0, 5), (2, 3), (1, 4)))
    yield "blobs_ragged", random_blobs((31, 27, 35), seed=5, rmax=6.0)
    nb = random_blobs((24, 24, 24), seed=6, rmax=4.5).copy()
    nb[nb > 0] = 0.5                                   # non-binary occupancy: != comparisons, not thresholds
    yield "non_binary", nb
    far = np.zeros((330, 6, 5), np.float32); far[3, 2, 2] = 1.0
    yield "clamp_330", far                             # distances beyond 317 hit the 100000 sentinel (dfields.cpp:50,71,87)


def test_oracle_equals_reference_stage_by_stage(R, O, synth):
    for name, v in edge_volumes(synth):
        p = R.fast_perim(v)
        assert same(p, O.fast_perim(v)), name
        sq = R.sqdist(p)
        assert same(sq, O.sqdist(p)) and same(sq, O.sqdist(p, brute=False)), name
        margin = R.sqrt(sq)
        assert same(margin, O.sqrt(sq)), name
        for bs in (4.0, 2.0, 8.0):
            band, tight = R.create_bands(margin, bs)
            bo, to = O.create_bands(margin, bs)
            assert same(band, bo) and same(tight, to), (name, bs)
        band, tight = R.create_bands(margin, 4.0)
        assert same(R.erode(band), O.erode(band)) and same(R.dilate(tight), O.dilate(tight)), name
        idx = R.find_indexes(band)
        assert same(idx, O.find_indexes(band)), name
        assert same(R.index_map(v.shape, idx), O.index_map(v.shape, idx)), name
    # getsqdist on an arbitrary float grid (zeros are the sites; non-zero values are ignored by pass 1)
    g = np.random.default_rng(3).integers(0, 3, (12, 10, 11)).astype(np.float32)
    assert same(R.sqdist(g), O.sqdist(g))


def test_oracle_equals_reference_qp(R, O, synth):
    for name, v in (("blobs", random_blobs((31, 27, 35), seed=5, rmax=6.0)), ("ellipsoid", DATA["ellipsoid_volume"])):
        margin = R.dfield(v)
        band, tight = R.create_bands(margin, 4.0)
        qr = R.prime_qp(v, margin, band, tight)
        qo = O.prime_qp(v, margin, band, tight)
        for k in ("lb", "ub", "x0", "invdg", "jc", "ir", "pr"):
            assert same(qr[k], qo[k]), (name, k)
        # rows ascending inside every column: the order doStep sums in
        for c in range(0, qr["n"], 97):
            rows = qr["ir"][qr["jc"][c]:qr["jc"][c + 1]]
            assert np.all(np.diff(rows) > 0)
        for iters in (1, 2, 7, 500):
            assert same(R.do_steps(qr, iters), O.run_qp(qo, iters=iters)), (name, iters)
        # sign constraints hold element-wise after the solve
        x = R.do_steps(qr, 500)
        assert np.all(x >= qr["lb"]) and np.all(x <= qr["ub"])


def test_oracle_equals_reference_feature_reset_quirks(R, O, synth):
    """getFeatureIndexes + the merge loop of runQP (primeqp.cpp:231-244, quadprog.cpp:92-101): a feature voxel outside
    the band contributes -1, which blocks every reset."""
    v = synth.box(40, (9, 7, 6), None, 0.0, 0.0)
    s = R.get_surface(v)
    n = R.surface_normals(v, s)
    smap = R.index_map(v.shape, s)
    fm = R.feature_map(v, smap, n, 0.75)
    assert fm.sum() > 0
    margin = R.dfield(v)
    band, _ = R.create_bands(margin, 4.0)
    imap = R.index_map(v.shape, R.find_indexes(band))
    assert same(R.feature_indexes(fm, imap), O.feature_indexes(fm, imap))
    assert same(R.optimize(v, fm, True), O.optimize(v, fm, using_features=True))
    fm2 = fm.copy(); fm2[1, 1, 1] = 1.0               # off-band feature voxel -> index -1 sorts first
    fi = R.feature_indexes(fm2, imap)
    assert fi[0] == -1 and same(fi, O.feature_indexes(fm2, imap))
    F2 = R.optimize(v, fm2, True)
    assert same(F2, O.optimize(v, fm2, using_features=True))
    assert same(F2, R.optimize(v))                    # ... so nothing was reset


def test_oracle_equals_reference_marching_cubes(R, O, synth, tmp_path):
    # F from the QP, and a raw binary volume at isolevel 0.5 (exact zeros in v1 - v2, degenerate triangles: m_cubes.cpp)
    v = random_blobs((31, 27, 35), seed=5, rmax=6.0)
    for name, F, iso in (("qp", R.optimize(v), 0.0), ("binary", v, 0.5)):
        vr, fr = R.mcubes(F, iso)
        vo, fo = O.mcubes(F, iso)
        assert same(vr, vo) and same(fr, fo), name
    F = R.optimize(v)
    p = str(tmp_path / "r.ply")
    R.mcubes_ply(p, F, 0.0, scale=tuple(FRAME["scale"]), shift=tuple(FRAME["shift"]), pad=FRAME["pad"])
    vo, fo = O.mcubes(F, 0.0)
    assert np.array_equal(np.frombuffer(open(p, "rb").read(), np.uint8), ply_bytes(O.write_ply, vo, fo, tmp_path))
    # VertexInterp incl. the endpoint swap and the |v1 - v2| <= 1e-5 branch (mcubes.cpp:369-392)
    for p1, p2, v1, v2 in (((1, 2, 3), (2, 2, 3), -0.3, 0.7), ((2, 2, 3), (1, 2, 3), 0.7, -0.3), ((1, 2, 3), (1, 2, 4), 1e-6, 2e-6),
                           ((5, 5, 5), (5, 6, 5), -4.99999, 1e-3)):
        vr, fr = R.vertex_interp(0.0, p1, p2, v1, v2), None
        lib = O.lib()
        if hasattr(lib, "mrbo_vertex_interp"):
            out = np.empty(3, np.float32)
            lib.mrbo_vertex_interp(C.c_float(0.0), (C.c_float * 3)(*p1), (C.c_float * 3)(*p2), C.c_float(v1), C.c_float(v2),
                                   out.ctypes.data_as(C.POINTER(C.c_float)))
            assert same(vr, out)


def feature_shapes(synth):
    yield "box", synth.box(48, (12, 9, 7), None, 0.0, 0.0)
    yield "blobs", random_blobs((44, 48, 52), seed=9)


def test_oracle_equals_reference_feature_path(R, O, synth):
    """isFeature / getVec / computePoint / fans / flipEdges (mcubes.cpp:9-152,234-311,396-588) and the normal pre-pass
    (dfields.cpp:199-349,447-555) as compiled from the reference (eigen-solver stand-in = mrb_eig3.h) == the oracle."""
    for name, v in feature_shapes(synth):
        s = R.get_surface(v)
        assert same(s, O.get_surface(v)), name
        n = R.surface_normals(v, s)
        assert same(n, O.surface_normals(v, s)), name
        smap = R.index_map(v.shape, s)
        for thr in (0.75, 0.9):
            assert same(R.feature_map(v, smap, n, thr), O.feature_map(v, smap, n, thr)), (name, thr)
        fm = R.feature_map(v, smap, n, 0.75)
        F = R.optimize(v, fm, True)
        assert same(F, O.optimize(v, fm, using_features=True)), name
        vr, fr = R.mcubes(F, 0.0, smap, n, 0.75, 0.8)
        vo, fo = O.mcubes_features(F, smap, n, 0.0, 0.75, 0.8)
        assert same(vr, vo) and same(fr, fo), name


# ---- what the build cannot pin: the eigen-solver / SVD arithmetic inside Eigen -----------------------------------
def _eig3_lib():
    """include/mrb_eig3.h compiled as plain C (the header both the kernels and the oracle use)."""
    import subprocess
    import tempfile
    root = os.path.dirname(HERE)
    d = tempfile.mkdtemp()
    src = os.path.join(d, "e.c")
    open(src, "w").write('#include "mrb_eig3.h"\n'
                         'void t_normal(const float* c, float* n) { mrb_normal_from_covariance(c, n); }\n'
                         'void t_eig(float* a, float* w, float* v) { mrb_eig3_jacobi(a, w, v); }\n'
                         'void t_point(const float* M, const float* vec, int n, const float* sh, int e, float* o) '
                         '{ mrb_feature_point(M, vec, n, sh, e, o); }\n')
    so = os.path.join(d, "e.so")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I", os.path.join(root, "include"),
                           src, "-o", so, "-lm"])
    return C.CDLL(so)


def test_eig3_against_lapack():
    """mrb_eig3.h (10 cyclic Jacobi sweeps in fp32) against numpy.linalg.eigh / lstsq in double."""
    lib = _eig3_lib()
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    rng = np.random.default_rng(11)
    worst_val = worst_vec = 0.0
    for trial in range(400):
        pts = rng.integers(-1, 2, (rng.integers(4, 20), 3)).astype(np.float64)       # covariances of 3x3x3 neighbourhoods
        if trial % 5 == 0:
            pts[:, 2] = 0.0                                                          # planar: a zero eigenvalue
        c = np.cov(pts.T, bias=True)
        a = np.ascontiguousarray(c, np.float32).reshape(9)
        w = np.empty(3, np.float32); v = np.empty(9, np.float32)
        lib.t_eig(fp(a.copy()), fp(w), fp(v))
        wl, vl = np.linalg.eigh(c)
        order = np.argsort(w, kind="stable")
        worst_val = max(worst_val, np.abs(w[order] - wl).max())
        V = v.reshape(3, 3)[:, order].astype(np.float64)
        assert np.abs(V.T @ V - np.eye(3)).max() < 5e-6                              # orthonormal
        assert np.abs(c @ V - V * w[order]).max() < 5e-6                             # eigen-pairs of c
        gaps = np.diff(wl)
        if gaps.min() > 1e-3:                                                        # direction is only defined off ties
            for k in range(3):
                worst_vec = max(worst_vec, 1.0 - abs(float(V[:, k] @ vl[:, k])))
    assert worst_val < 2e-6 and worst_vec < 1e-6, (worst_val, worst_vec)
    # least-squares feature point: corners (full rank) against lstsq, edges (rank 2) against the truncated SVD
    worst = worst50 = 0.0
    for trial in range(300):
        n = int(rng.integers(3, 9))
        base = rng.normal(size=(3, 3)); base /= np.linalg.norm(base, axis=1, keepdims=True)
        is_edge = trial % 2
        if is_edge:
            base[2] = base[0]                                                        # two distinct normal directions only
        M = base[rng.integers(0, 3, n)] + rng.normal(scale=1e-3, size=(n, 3))
        M[:3] = base + rng.normal(scale=1e-3, size=(3, 3))
        M /= np.linalg.norm(M, axis=1, keepdims=True)
        M32 = np.ascontiguousarray(M, np.float32)
        vec = np.ascontiguousarray(rng.uniform(-0.5, 0.5, n), np.float32)
        shift = np.ascontiguousarray(rng.uniform(3, 40, 3), np.float32)
        out = np.empty(3, np.float32)
        lib.t_point(fp(M32), fp(vec), n, fp(shift), int(is_edge), fp(out))
        U, S, Vt = np.linalg.svd(M32.astype(np.float64), full_matrices=False)
        keep = S > (S[-1] / (S[0] - 0.0001) if is_edge else 3 * 1.1920929e-7) * S[0]     # mcubes.cpp:144-149 + Eigen's rank rule
        p = (Vt.T[:, keep] / S[keep]) @ (U.T[keep] @ vec.astype(np.float64)) + shift
        cond = S[keep][0] / S[keep][-1]
        err = np.abs(out - p).max()
        if np.abs(p - shift).max() > 1.0:
            continue                                 # the reference discards feature points further than 1 voxel (mcubes.cpp:249)
        if cond < 10:
            worst = max(worst, err)
        elif cond < 50:
            worst50 = max(worst50, err)
    # fp32 normal equations lose cond^2: within the north star's 1e-4 voxel tolerance for well-separated normals
    # (cond < 10, i.e. opening angles above ~12 degrees), within 1e-3 up to cond 50
    assert worst < 1e-4 and worst50 < 1e-3, (worst, worst50)


def test_double_standin_bounds_the_unpinned_arithmetic(R, synth):
    """Reference code with an independent double-precision eigen-solver / SVD stand-in against the mrb_eig3.h one:
    normals agree to 1e-5 except where the covariance has a (near-)double smallest eigenvalue (the direction is then
    undefined for ANY solver, Eigen included); the feature meshes agree in all but a few cells."""
    from oracle import ref
    if not ref.available(double=True):
        pytest.skip("double stand-in not built")
    v = synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0))
    s = R.get_surface(v)
    n = R.surface_normals(v, s)
    nd = R.surface_normals(v, s, double=True)
    d = np.abs(n - nd).max(axis=1)
    assert (d > 1e-5).mean() < 0.01, (d > 1e-5).mean()
    smap = R.index_map(v.shape, s)
    fm = R.feature_map(v, smap, n, 0.75)
    F = R.optimize(v, fm, True)
    cid, feat, fp_ = R.feature_cells(F, smap, n)
    cid2, feat2, fp2 = R.feature_cells(F, smap, n, double=True)         # same normals, other SVD
    assert same(cid, cid2)
    agree = feat == feat2
    assert agree.mean() > 0.995, agree.mean()
    both = agree & (feat != 0)
    assert both.sum() > 100
    err = np.abs(fp_[both] - fp2[both]).max(axis=1)
    assert np.quantile(err, 0.99) < 1e-4 and np.median(err) < 1e-5, (np.quantile(err, 0.99), np.median(err))


# ---- GPU: the CUDA library reproduces the reference's outputs, through the C-ABI -----------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_gpu_reproduces_the_reference_fixtures(ctx, synth, tmp_path, name):
    m = META[name]
    v = volume(name, synth)
    sq = ctx.getsqdist(ctx.fastPerim(v))
    margin = ctx.dfield_gpu(v)
    band, tight = ctx.createBands(margin, 4.0)
    q = ctx.primeQP(v, margin, band, tight)
    F = ctx.optimize(v)
    verts, faces = ctx.reconstruct(v)
    v2, f2 = ctx.mcubes(F, 0.0)
    assert same(verts, v2) and same(faces, f2)
    r = dict(sq=sq.astype(np.int32), margin=margin, band_idx=np.asarray(ctx.findIndexes(band), np.int32), F=F, verts=verts,
             faces=np.asarray(faces, np.int32), ply=ply_bytes(ctx.write_ply, verts, faces, tmp_path),
             **{k: q[k] for k in ("lb", "ub", "x0", "invdg", "jc", "ir", "pr")})
    assert int((tight != 0).sum()) == m["n_tight"]
    r["x_csc"] = None
    del r["x_csc"]
    if m["features"]:
        s = ctx.getSurface(v)
        n = ctx.getSurfaceNormals(v, s)
        smap = ctx.getIndexMap(v.shape, s)
        fm = ctx.getFeatureMap(v, smap, n, 0.75)
        Ff = ctx.optimize(v, fm, True)
        vf, ff = ctx.mcubes_features(Ff, smap, n, 0.0, 0.75, 0.8)
        vf2, ff2 = ctx.reconstruct(v, using_features=True)
        assert same(vf, vf2) and same(ff, ff2)
        r.update(surface=np.asarray(s, np.int32), normals=n, feature_map=fm, F_feat=Ff, verts_feat=vf,
                 faces_feat=np.asarray(ff, np.int32))
    check_against_fixture(name, m, r)


@pytest.mark.gpu
def test_gpu_equals_live_reference(ctx, R, synth):
    """Where oracle/_ref travelled to the GPU box: the CUDA path against the reference code itself on fresh shapes."""
    for name, v in (("blobs", random_blobs((45, 39, 50), seed=21)), ("torus", synth.torus(48, 12.0, 5.0))):
        F = ctx.optimize(v)
        assert same(F, R.optimize(v)), name
        vr, fr = R.mcubes(F, 0.0)
        verts, faces = ctx.reconstruct(v)
        assert same(verts, vr) and same(faces, fr), name
    v = synth.box(44, (10, 8, 6), None, 15.0, 25.0)
    s = ctx.getSurface(v)
    n = ctx.getSurfaceNormals(v, s)
    assert same(s, R.get_surface(v)) and same(n, R.surface_normals(v, s))
    smap = ctx.getIndexMap(v.shape, s)
    fm = ctx.getFeatureMap(v, smap, n, 0.75)
    assert same(fm, R.feature_map(v, smap, n, 0.75))
    verts, faces = ctx.reconstruct(v, using_features=True)
    vr, fr = R.mcubes(R.optimize(v, fm, True), 0.0, smap, n, 0.75, 0.8)
    assert same(verts, vr) and same(faces, fr)
