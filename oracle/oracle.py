"""ctypes binding of the CPU oracle (oracle/_build/libmrb_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product (mesh_reconstruction_b200) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmrb_oracle.so")
_lib = None


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("mrb_oracle.cpp", "frontend_oracle.cpp")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


class _Mesh(C.Structure):
    _fields_ = [("verts", C.POINTER(C.c_float)), ("faces", C.POINTER(C.c_int)),
                ("nverts", C.c_int), ("nfaces", C.c_int)]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.mrbo_last_error.restype = C.c_char_p
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int))


def _dims(shape):
    return (C.c_int * 3)(*shape)


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, lib().mrbo_last_error().decode()))


def _grid_op(name, g, *extra):
    g, gp = _f(g)
    out = np.empty_like(g)
    _check(getattr(lib(), name)(gp, _dims(g.shape), out.ctypes.data_as(C.POINTER(C.c_float)), *extra), name)
    return out


def fast_perim(vol):
    return _grid_op("mrbo_fast_perim", vol)


def sqdist(g, brute=True):
    return _grid_op("mrbo_sqdist_brute" if brute else "mrbo_sqdist_fast", g)


def sqrt(g):
    return _grid_op("mrbo_sqrt", g)


def dfield(vol, brute=True):
    return _grid_op("mrbo_dfield", vol, C.c_int(1 if brute else 0))


def erode(g):
    return _grid_op("mrbo_erode", g)


def dilate(g):
    return _grid_op("mrbo_dilate", g)


def create_bands(margin, band_size=4.0):
    m, mp = _f(margin)
    band = np.empty_like(m)
    tight = np.empty_like(m)
    _check(lib().mrbo_create_bands(mp, _dims(m.shape), C.c_float(band_size),
                                   band.ctypes.data_as(C.POINTER(C.c_float)),
                                   tight.ctypes.data_as(C.POINTER(C.c_float))), "create_bands")
    return band, tight


def find_indexes(g):
    g, gp = _f(g)
    n = C.c_int(0)
    _check(lib().mrbo_find_indexes(gp, _dims(g.shape), None, 0, C.byref(n)), "find_indexes")
    idx = np.empty(max(n.value, 1), dtype=np.int32)
    _check(lib().mrbo_find_indexes(gp, _dims(g.shape), idx.ctypes.data_as(C.POINTER(C.c_int)), n.value, C.byref(n)),
           "find_indexes")
    return idx[:n.value].copy()


def index_map(shape, indexes):
    idx, ip = _i(indexes)
    out = np.empty(shape, dtype=np.float32)
    _check(lib().mrbo_index_map(_dims(shape), ip, len(idx), out.ctypes.data_as(C.POINTER(C.c_float))), "index_map")
    return out


def feature_indexes(feature_map, imap):
    fm, fp = _f(feature_map)
    im, ip = _f(imap)
    n = C.c_int(0)
    _check(lib().mrbo_feature_indexes(fp, ip, _dims(fm.shape), None, 0, C.byref(n)), "feature_indexes")
    out = np.empty(max(n.value, 1), dtype=np.int32)
    _check(lib().mrbo_feature_indexes(fp, ip, _dims(fm.shape), out.ctypes.data_as(C.POINTER(C.c_int)), n.value,
                                      C.byref(n)), "feature_indexes")
    return out[:n.value].copy()


def prime_qp(volume, margin, band, tight):
    """Returns dict(lb, ub, x0, invdg, jc, ir, pr)."""
    v, vp = _f(volume)
    m, mp = _f(margin)
    b, bp = _f(band)
    t, tp = _f(tight)
    n = int(np.count_nonzero(b))
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
    lb = np.empty(n, np.float32); ub = np.empty(n, np.float32)
    x0 = np.empty(n, np.float32); invdg = np.empty(n, np.float32)
    jc = np.empty(n + 1, np.int32)
    nnz = C.c_int(0)
    _check(lib().mrbo_prime_qp(vp, mp, bp, tp, _dims(v.shape), n, fp(lb), fp(ub), fp(x0), fp(invdg), ip(jc),
                               None, None, 0, C.byref(nnz)), "prime_qp")
    ir = np.empty(max(nnz.value, 1), np.int32)
    pr = np.empty(max(nnz.value, 1), np.float32)
    _check(lib().mrbo_prime_qp(vp, mp, bp, tp, _dims(v.shape), n, None, None, None, None, ip(jc),
                               ip(ir), fp(pr), nnz.value, C.byref(nnz)), "prime_qp")
    return dict(lb=lb, ub=ub, x0=x0, invdg=invdg, jc=jc, ir=ir[:nnz.value], pr=pr[:nnz.value], n=n)


def run_qp(q, iters=500, feature_idx=None, using_features=False):
    n = q["n"]
    fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
    ip = lambda a: np.ascontiguousarray(a, np.int32).ctypes.data_as(C.POINTER(C.c_int))
    feat = np.ascontiguousarray(feature_idx if feature_idx is not None else np.zeros(1), np.int32)
    nfeat = 0 if feature_idx is None else len(feature_idx)
    x = np.empty(n, np.float32)
    _check(lib().mrbo_run_qp(n, ip(q["jc"]), ip(q["ir"]), fp(q["pr"]), fp(q["invdg"]), fp(q["lb"]), fp(q["ub"]),
                             fp(q["x0"]), iters, feat.ctypes.data_as(C.POINTER(C.c_int)), nfeat,
                             1 if using_features else 0, x.ctypes.data_as(C.POINTER(C.c_float))), "run_qp")
    return x


def optimize(volume, feature_map=None, using_features=False, band_size=4, iters=500, brute_edt=True, return_info=False):
    v, vp = _f(volume)
    fm = None
    fmp = None
    if feature_map is not None:
        fm, fmp = _f(feature_map)
    F = np.empty_like(v)
    nb = C.c_int(0)
    ms = (C.c_double * 4)()
    _check(lib().mrbo_optimize(vp, fmp, _dims(v.shape), 1 if using_features else 0, band_size, iters,
                               1 if brute_edt else 0, F.ctypes.data_as(C.POINTER(C.c_float)), C.byref(nb), ms),
           "optimize")
    if return_info:
        return F, dict(n_band=nb.value, ms_dfield=ms[0], ms_band=ms[1], ms_prime=ms[2], ms_qp=ms[3])
    return F


def hash_points(xyz, shape):
    p, pp = _f(np.asarray(xyz).reshape(-1, 3))
    keys = np.empty(len(p), np.uint64)
    _check(lib().mrbo_hash_points(pp, len(p), _dims(shape), keys.ctypes.data_as(C.POINTER(C.c_uint64))), "hash")
    return keys


def mcubes(F, isolevel=0.0):
    f, fp = _f(F)
    m = _Mesh()
    _check(lib().mrbo_mcubes(fp, _dims(f.shape), C.c_float(isolevel), C.byref(m)), "mcubes")
    verts = np.ctypeslib.as_array(m.verts, shape=(max(m.nverts, 1), 3))[:m.nverts].copy()
    faces = np.ctypeslib.as_array(m.faces, shape=(max(m.nfaces, 1), 3))[:m.nfaces].copy()
    lib().mrbo_mesh_free(C.byref(m))
    return verts, faces


def write_ply(path, verts, faces, scale=(1, 1, 1), shift=(0, 0, 0), pad=6):
    v, vp = _f(verts)
    f, fpp = _i(faces)
    sc = (C.c_float * 3)(*scale)
    sh = (C.c_float * 3)(*shift)
    _check(lib().mrbo_write_ply(path.encode(), vp, len(v), fpp, len(f), sc, sh, pad), "write_ply")


def get_surface(vol):
    v, vp = _f(vol)
    n = C.c_int(0)
    _check(lib().mrbo_get_surface(vp, _dims(v.shape), None, 0, C.byref(n)), "get_surface")
    out = np.empty(max(n.value, 1), np.int32)
    _check(lib().mrbo_get_surface(vp, _dims(v.shape), out.ctypes.data_as(C.POINTER(C.c_int)), n.value, C.byref(n)), "get_surface")
    return out[:n.value].copy()


def surface_normals(vol, surface):
    v, vp = _f(vol)
    s, sp = _i(surface)
    out = np.empty((max(len(s), 1), 3), np.float32)
    _check(lib().mrbo_surface_normals(vp, _dims(v.shape), sp, len(s), out.ctypes.data_as(C.POINTER(C.c_float))), "surface_normals")
    return out[:len(s)].copy()


def feature_map(vol, surface_map, normals, threshold):
    v, vp = _f(vol)
    m, mp = _f(surface_map)
    nr, nrp = _f(np.asarray(normals).reshape(-1, 3))
    out = np.empty_like(v)
    _check(lib().mrbo_feature_map(vp, mp, nrp, len(nr), _dims(v.shape), C.c_float(threshold),
                                  out.ctypes.data_as(C.POINTER(C.c_float))), "feature_map")
    return out


def mcubes_features(F, surface_map, normals, isolevel=0.0, feat_thresh=0.75, corner_thresh=0.8):
    f, fp = _f(F)
    m, mp = _f(surface_map)
    nr, nrp = _f(np.asarray(normals).reshape(-1, 3))
    mesh = _Mesh()
    _check(lib().mrbo_mcubes_features(fp, mp, nrp, len(nr), _dims(f.shape), C.c_float(isolevel), C.c_float(feat_thresh),
                                      C.c_float(corner_thresh), C.byref(mesh)), "mcubes_features")
    verts = np.ctypeslib.as_array(mesh.verts, shape=(max(mesh.nverts, 1), 3))[:mesh.nverts].copy()
    faces = np.ctypeslib.as_array(mesh.faces, shape=(max(mesh.nfaces, 1), 3))[:mesh.nfaces].copy()
    lib().mrbo_mesh_free(C.byref(mesh))
    return verts, faces


# ---- front end (row f1): oracle/frontend_oracle.cpp ------------------------------------------------------------------
def binvox_to_points(voxels, dwh, xform, res):
    """binvoxToPCL (binvoxToPcl.cpp:105-169) on decoded voxel bytes; dwh = (depth, width, height), xform = (scale, tx, ty, tz)."""
    v = np.ascontiguousarray(voxels, dtype=np.uint8).ravel()
    d = (C.c_int * 3)(*dwh)
    x = (C.c_float * 4)(*xform)
    vp = v.ctypes.data_as(C.POINTER(C.c_ubyte))
    n = lib().fo_binvox_to_points(vp, d, x, C.c_int(res), None, C.c_int(0))
    out = np.empty((n, 3), np.float32)
    lib().fo_binvox_to_points(vp, d, x, C.c_int(res), out.ctypes.data_as(C.POINTER(C.c_float)), C.c_int(n))
    return out


def nn_stats(xyz):
    """(dist_total, count) of getDensity / getResolution (binvoxToPcl.cpp:172-201, voxelize.cpp:6-32), exact brute-force NN."""
    a, ap = _f(np.asarray(xyz).reshape(-1, 3))
    dt, cnt = C.c_float(0), C.c_int(0)
    lib().fo_nn_stats(ap, C.c_int(len(a)), C.byref(dt), C.byref(cnt))
    return np.float32(dt.value), cnt.value


def _front_out(meta, fmeta, vol):
    return {"volume": vol, "dims": tuple(meta[:3]), "pad": meta[3], "occupied_leaves": meta[4],
            "scale": np.array(fmeta[:3], np.float32), "shift": np.array(fmeta[3:6], np.float32)}


def voxelize_leaf(xyz, leaf, res_factor):
    """voxelize + grid(cloud, vox) + fillGrid + getBinaryVolume (voxelize.cpp:34-51, grid.cpp:84-121,213-265,322-334)."""
    a, ap = _f(np.asarray(xyz).reshape(-1, 3))
    meta, fmeta = (C.c_int * 5)(), (C.c_float * 6)()
    rc = lib().fo_voxelize_leaf(ap, C.c_int(len(a)), C.c_float(leaf), C.c_int(res_factor), meta, fmeta, None, C.c_longlong(0))
    if rc != 0:
        raise RuntimeError("fo_voxelize_leaf failed (%d)" % rc)
    vol = np.empty(tuple(meta[:3]), np.float32)
    rc = lib().fo_voxelize_leaf(ap, C.c_int(len(a)), C.c_float(leaf), C.c_int(res_factor), meta, fmeta,
                                vol.ctypes.data_as(C.POINTER(C.c_float)), C.c_longlong(vol.size))
    assert rc == 0
    return _front_out(meta, fmeta, vol)


def front_end(voxels, dwh, xform, observed, res_factor=0, max_voxels=1 << 28):
    """mesh_reconstruction.cpp:99-120: resolution factor, up-sampling, merged cloud, leaf size, occupancy volume."""
    v = np.ascontiguousarray(voxels, dtype=np.uint8).ravel()
    a, ap = _f(np.asarray(observed).reshape(-1, 3))
    meta, fmeta, info, finfo = (C.c_int * 5)(), (C.c_float * 6)(), (C.c_int * 5)(), (C.c_float * 5)()
    args = (v.ctypes.data_as(C.POINTER(C.c_ubyte)), (C.c_int * 3)(*dwh), (C.c_float * 4)(*xform), ap, C.c_int(len(a)), C.c_int(res_factor))
    rc = lib().fo_front_end(*args, meta, fmeta, info, finfo, None, C.c_longlong(0))
    if rc != 0:
        raise RuntimeError("fo_front_end failed (%d)" % rc)
    vol = np.empty(tuple(meta[:3]), np.float32)
    rc = lib().fo_front_end(*args, meta, fmeta, info, finfo, vol.ctypes.data_as(C.POINTER(C.c_float)), C.c_longlong(vol.size))
    assert rc == 0
    out = _front_out(meta, fmeta, vol)
    out.update(res_factor=info[0], n_predicted=info[1], n_cloud=info[2], observed_count=info[3], completed_count=info[4],
               observed_dist_total=np.float32(finfo[0]), completed_dist_total=np.float32(finfo[1]),
               observed_density=np.float32(finfo[2]), completed_density=np.float32(finfo[3]), leaf=np.float32(finfo[4]))
    return out
