"""ctypes binding of oracle/_ref/libmrb_ref.so: the reference's own translation units, compiled unmodified from
/root/reference by oracle/build_ref.sh against stand-in third-party headers.  TEST INFRASTRUCTURE ONLY.

Same call shapes as oracle/oracle.py so that a test can run one input through the reference code, the restated oracle
and the CUDA library.  available() is False where the library was never built (no /root/reference and no shipped
oracle/_ref/): tests then fall back to the fixtures that tests/golden/make_ref_golden.py generated from it.
"""
import ctypes as C
import os
import subprocess
import threading
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def _so(double):
    return os.path.join(_HERE, "_ref", "libmrb_ref_dbl.so" if double else "libmrb_ref.so")


def build():
    subprocess.check_call(["sh", os.path.join(_HERE, "build_ref.sh")])


def available(double=False):
    if not os.path.exists(_so(double)) and os.path.isdir(os.environ.get("MRB_REFERENCE_DIR", "/root/reference")):
        build()
    return os.path.exists(_so(double))


def lib(double=False):
    if double not in _libs:
        if not available(double):
            raise RuntimeError("oracle/_ref is not built (needs /root/reference; run oracle/build_ref.sh)")
        l = C.CDLL(_so(double))
        l.ref_build_info.restype = C.c_char_p
        _libs[double] = l
    return _libs[double]


def _big_stack(fn, *a, **kw):
    """primeqp.cpp:13 keeps one Vector3i per tight voxel in a stack VLA: run on a thread with a large stack."""
    out = {}

    def run():
        try:
            out["r"] = fn(*a, **kw)
        except BaseException as e:  # noqa
            out["e"] = e
    old = threading.stack_size(1 << 30)
    try:
        t = threading.Thread(target=run)
        t.start()
        t.join()
    finally:
        threading.stack_size(old)
    if "e" in out:
        raise out["e"]
    return out["r"]


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int))


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def _dims(shape):
    return (C.c_int * 3)(*shape)


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("reference %s failed (%d)" % (what, rc))


def _grid_op(name, g, *extra, **kw):
    g, gp = _f(g)
    out = np.empty_like(g)
    _check(getattr(lib(kw.get("double", False)), name)(gp, _dims(g.shape), *extra, _fp(out)), name)
    return out


def fast_perim(vol):
    return _grid_op("ref_fast_perim", vol)


def sqdist(g):
    return _grid_op("ref_sqdist", g)


def sqrt(g):
    return _grid_op("ref_sqrt", g)


def dfield(vol):
    return sqrt(sqdist(fast_perim(vol)))


def erode(g):
    return _grid_op("ref_erode", g)


def dilate(g):
    return _grid_op("ref_dilate", g)


def create_bands(margin, band_size=4.0):
    m, mp = _f(margin)
    band = np.empty_like(m)
    tight = np.empty_like(m)
    _check(lib().ref_create_bands(mp, _dims(m.shape), C.c_float(band_size), _fp(band), _fp(tight)), "createBands")
    return band, tight


def _int_list(call):
    n = C.c_int(0)
    _check(call(None, 0, C.byref(n)), "count")
    out = np.empty(max(n.value, 1), np.int32)
    _check(call(_ip(out), n.value, C.byref(n)), "fill")
    return out[:n.value].copy()


def find_indexes(g):
    g, gp = _f(g)
    return _int_list(lambda o, c, n: lib().ref_find_indexes(gp, _dims(g.shape), o, c, n))


def index_map(shape, indexes):
    idx, ip = _i(indexes)
    out = np.empty(shape, np.float32)
    _check(lib().ref_index_map(_dims(shape), ip, len(idx), _fp(out)), "getIndexMap")
    return out


def feature_indexes(feature_map, imap):
    fm, fp = _f(feature_map)
    im, ip = _f(imap)
    return _int_list(lambda o, c, n: lib().ref_feature_indexes(fp, ip, _dims(fm.shape), o, c, n))


def prime_qp(volume, margin, band, tight):
    v, vp = _f(volume)
    m, mp = _f(margin)
    b, bp = _f(band)
    t, tp = _f(tight)
    n = int(np.count_nonzero(b))
    lb = np.empty(n, np.float32); ub = np.empty(n, np.float32)
    x0 = np.empty(n, np.float32); invdg = np.empty(n, np.float32)
    jc = np.empty(n + 1, np.int32)
    cap = 13 * n + 16
    ir = np.empty(cap, np.int32); pr = np.empty(cap, np.float32)
    nnz = C.c_int(0)

    def call():
        _check(lib().ref_prime_qp(vp, mp, bp, tp, _dims(v.shape), n, _fp(lb), _fp(ub), _fp(x0), _fp(invdg), _ip(jc),
                                  _ip(ir), _fp(pr), cap, C.byref(nnz)), "primeQP")
    _big_stack(call)
    return dict(lb=lb, ub=ub, x0=x0, invdg=invdg, jc=jc, ir=ir[:nnz.value].copy(), pr=pr[:nnz.value].copy(), n=n)


def do_steps(q, iters):
    n = q["n"]
    x = np.empty(n, np.float32)
    a = {k: np.ascontiguousarray(q[k]) for k in ("jc", "ir", "pr", "invdg", "lb", "ub", "x0")}
    _check(lib().ref_do_steps(n, _ip(a["jc"]), _ip(a["ir"]), _fp(a["pr"]), _fp(a["invdg"]), _fp(a["lb"]), _fp(a["ub"]),
                              _fp(a["x0"]), iters, _fp(x)), "doStep")
    return x


def optimize(volume, feature_map=None, using_features=False):
    v, vp = _f(volume)
    fmp = None
    if feature_map is not None:
        fm, fmp = _f(feature_map)
    F = np.empty_like(v)
    _big_stack(lambda: _check(lib().ref_optimize(vp, fmp, _dims(v.shape), 1 if using_features else 0, _fp(F)), "optimize"))
    return F


def get_surface(vol):
    v, vp = _f(vol)
    return _int_list(lambda o, c, n: lib().ref_get_surface(vp, _dims(v.shape), o, c, n))


def surface_normals(vol, surface, double=False):
    v, vp = _f(vol)
    s, sp = _i(surface)
    out = np.empty((max(len(s), 1), 3), np.float32)
    _check(lib(double).ref_surface_normals(vp, _dims(v.shape), sp, len(s), _fp(out)), "getSurfaceNormals")
    return out[:len(s)].copy()


def feature_map(vol, surface_map, normals, threshold):
    v, vp = _f(vol)
    m, mp = _f(surface_map)
    nr, nrp = _f(np.asarray(normals).reshape(-1, 3))
    out = np.empty_like(v)
    _check(lib().ref_feature_map(vp, mp, nrp, len(nr), _dims(v.shape), C.c_float(threshold), _fp(out)), "getFeatureMap")
    return out


def vertex_interp(iso, p1, p2, v1, v2):
    a, ap = _f(p1)
    b, bp = _f(p2)
    out = np.empty(3, np.float32)
    _check(lib().ref_vertex_interp(C.c_float(iso), ap, bp, C.c_float(v1), C.c_float(v2), _fp(out)), "VertexInterp")
    return out


def _feat_args(surface_map, normals):
    if surface_map is None:
        return None, None, None, None, 0
    m, mp = _f(surface_map)
    nr, nrp = _f(np.asarray(normals).reshape(-1, 3))
    return m, mp, nr, nrp, len(nr)


def mcubes(F, isolevel=0.0, surface_map=None, normals=None, feat_thresh=0.75, corner_thresh=0.8, double=False):
    """Raw vertices / faces of the reference's mcubes body (features when surface_map / normals are given)."""
    f, fp = _f(F)
    m, mp, nr, nrp, nn = _feat_args(surface_map, normals)
    uf = 0 if surface_map is None else 1
    nv = C.c_int(0); nf = C.c_int(0)
    args = (fp, mp, nrp, nn, _dims(f.shape), C.c_float(isolevel), C.c_float(feat_thresh), C.c_float(corner_thresh), uf)
    _check(lib(double).ref_mcubes_mesh(*args, None, None, C.byref(nv), C.byref(nf)), "mcubes")
    verts = np.empty((max(nv.value, 1), 3), np.float32)
    faces = np.empty((max(nf.value, 1), 3), np.int32)
    _check(lib(double).ref_mcubes_mesh(*args, _fp(verts), _ip(faces), C.byref(nv), C.byref(nf)), "mcubes")
    return verts[:nv.value].copy(), faces[:nf.value].copy()


def mcubes_ply(path, F, isolevel=0.0, surface_map=None, normals=None, feat_thresh=0.75, corner_thresh=0.8,
               scale=(1, 1, 1), shift=(0, 0, 0), pad=6):
    """The reference's mcubes() itself: writes the PLY file."""
    f, fp = _f(F)
    m, mp, nr, nrp, nn = _feat_args(surface_map, normals)
    uf = 0 if surface_map is None else 1
    sc = (C.c_float * 3)(*scale)
    sh = (C.c_float * 3)(*shift)
    _check(lib().ref_mcubes_ply(fp, mp, nrp, nn, _dims(f.shape), C.c_float(isolevel), C.c_float(feat_thresh),
                                C.c_float(corner_thresh), path.encode(), uf, sc, sh, pad), "mcubes")


def feature_cells(F, surface_map, normals, isolevel=0.0, feat_thresh=0.75, corner_thresh=0.8, double=False):
    f, fp = _f(F)
    m, mp, nr, nrp, nn = _feat_args(surface_map, normals)
    n = C.c_int(0)
    args = (fp, mp, nrp, nn, _dims(f.shape), C.c_float(isolevel), C.c_float(feat_thresh), C.c_float(corner_thresh))
    _check(lib(double).ref_feature_cells(*args, None, None, None, 0, C.byref(n)), "cells")
    cid = np.empty(max(n.value, 1), np.int32); feat = np.empty(max(n.value, 1), np.int32)
    fpnt = np.empty((max(n.value, 1), 3), np.float32)
    _check(lib(double).ref_feature_cells(*args, _ip(cid), _ip(feat), _fp(fpnt), n.value, C.byref(n)), "cells")
    return cid[:n.value].copy(), feat[:n.value].copy(), fpnt[:n.value].copy()


def cloud_point(p, scale, shift, pad):
    a, ap = _f(p)
    s, sp = _f(scale)
    h, hp = _f(shift)
    out = np.empty(3, np.float32)
    _check(lib().ref_cloud_point(ap, sp, hp, pad, _fp(out)), "getCloudPoint")
    return out


# ---- front end (row f1): oracle/_ref/libmrb_ref_front.so = binvoxToPcl.cpp, confidence_pcl/, voxelize.cpp, grid.cpp ----------
_front = None


def front_available():
    so = os.path.join(_HERE, "_ref", "libmrb_ref_front.so")
    if not os.path.exists(so) and os.path.isdir(os.environ.get("MRB_REFERENCE_DIR", "/root/reference")):
        build()
    return os.path.exists(so)


def front():
    global _front
    if _front is None:
        if not front_available():
            raise RuntimeError("oracle/_ref/libmrb_ref_front.so is not built (needs /root/reference)")
        _front = C.CDLL(os.path.join(_HERE, "_ref", "libmrb_ref_front.so"))
        _front.reff_density.restype = C.c_float
        _front.reff_resolution.restype = C.c_float
    return _front


def read_binvox(path):
    """read_binvox (binvoxToPcl.cpp:15-103): (voxel bytes, (depth, width, height), (scale, tx, ty, tz))."""
    hdr, fh = (C.c_int * 3)(), (C.c_float * 4)()
    n = front().reff_read_binvox(path.encode(), hdr, fh, None, C.c_longlong(0))
    if n < 0:
        raise RuntimeError("read_binvox failed")
    v = np.empty(n, np.uint8)
    front().reff_read_binvox(path.encode(), hdr, fh, v.ctypes.data_as(C.POINTER(C.c_ubyte)), C.c_longlong(n))
    return v, (hdr[0], hdr[2], hdr[1]), (fh[3], fh[0], fh[1], fh[2])


def binvox_to_points(path, res):
    n = front().reff_binvox_to_points(path.encode(), C.c_int(res), None, C.c_int(0))
    out = np.empty((n, 3), np.float32)
    front().reff_binvox_to_points(path.encode(), C.c_int(res), _fp(out), C.c_int(n))
    return out


def density(xyz):
    a = np.ascontiguousarray(xyz, np.float32)
    return np.float32(front().reff_density(_fp(a), C.c_int(len(a))))


def resolution(xyz):
    a = np.ascontiguousarray(xyz, np.float32)
    return np.float32(front().reff_resolution(_fp(a), C.c_int(len(a))))


def resolution_factor(path, observed):
    a = np.ascontiguousarray(observed, np.float32)
    return front().reff_resolution_factor(path.encode(), _fp(a), C.c_int(len(a)))


def _vol_out(meta, fm, vol):
    return {"volume": vol, "dims": tuple(meta[:3]), "pad": meta[3], "scale": np.array(fm[:3], np.float32), "shift": np.array(fm[3:6], np.float32)}


def volume_of_cloud(xyz, res_factor):
    """voxelizeData + createGrid(res_factor) + getBinaryVolume on a cloud whose strengths are all 1."""
    a = np.ascontiguousarray(xyz, np.float32)
    meta, fm, leaf = (C.c_int * 4)(), (C.c_float * 6)(), C.c_float()
    rc = front().reff_volume_of_cloud(_fp(a), C.c_int(len(a)), C.c_int(res_factor), meta, fm, C.byref(leaf), None, C.c_longlong(0))
    assert rc == 0
    vol = np.empty(tuple(meta[:3]), np.float32)
    rc = front().reff_volume_of_cloud(_fp(a), C.c_int(len(a)), C.c_int(res_factor), meta, fm, C.byref(leaf), _fp(vol), C.c_longlong(vol.size))
    assert rc == 0
    out = _vol_out(meta, fm, vol)
    out["leaf"] = np.float32(leaf.value)
    return out


def front_end(binvox_path, observed):
    """mesh_reconstruction.cpp:99-120."""
    a = np.ascontiguousarray(observed, np.float32)
    meta, fm, info, leaf = (C.c_int * 4)(), (C.c_float * 6)(), (C.c_int * 3)(), C.c_float()
    rc = front().reff_front_end(binvox_path.encode(), _fp(a), C.c_int(len(a)), meta, fm, info, C.byref(leaf), None, C.c_longlong(0))
    assert rc == 0
    vol = np.empty(tuple(meta[:3]), np.float32)
    rc = front().reff_front_end(binvox_path.encode(), _fp(a), C.c_int(len(a)), meta, fm, info, C.byref(leaf), _fp(vol), C.c_longlong(vol.size))
    assert rc == 0
    out = _vol_out(meta, fm, vol)
    out.update(res_factor=info[0], n_predicted=info[1], n_cloud=info[2], leaf=np.float32(leaf.value))
    return out
