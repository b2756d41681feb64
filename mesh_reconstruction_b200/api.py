"""Host-side mirror of the reference's module API on top of the C-ABI (libmrb.so).

Function names follow the reference (dfields.h, narrowBand.h, grid.h, primeqp.h, quadprog.h,
mcubes.h, hasher.h) so that parity tests read like calls into the reference.  Grids are numpy
float32 arrays of shape (nx, ny, nz) in C order (== voxels[x][y][z]) or torch CUDA tensors of the
same shape/dtype (passed by device pointer, results returned as torch tensors on the same device).

All compute happens in the CUDA library; nothing here computes on the CPU.
"""
import ctypes as C
import numpy as np

from . import _lib as L


class MrbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("mrb error %d: %s" % (code, msg))
        self.code = code


def _is_torch(a):
    return type(a).__module__.startswith("torch")


class Context:
    """One GPU context (stream + scratch arena).  Not thread-safe; one per thread / GPU."""

    def __init__(self, device=0):
        self.lib = L.load()
        h = C.c_void_p()
        rc = self.lib.mrb_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise MrbError(rc, self.lib.mrb_last_error().decode())
        self.h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.lib.mrb_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def stream(self):
        return self.lib.mrb_ctx_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.mrb_ctx_launch_count(self.h))

    # ---- helpers -------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise MrbError(rc, self.lib.mrb_last_error().decode())

    @staticmethod
    def _dims(shape):
        if len(shape) != 3:
            raise ValueError("grid must be 3-D, got shape %r" % (tuple(shape),))
        return (C.c_int * 3)(*[int(s) for s in shape])

    @staticmethod
    def _in(a, dtype=np.float32):
        """-> (keepalive, pointer, shape, like_torch)"""
        if _is_torch(a):
            import torch
            want = {np.float32: torch.float32, np.int32: torch.int32}[dtype]
            if a.dtype != want or not a.is_contiguous():
                a = a.to(want).contiguous()
            if a.is_cuda:
                # The library works on its own (non-blocking) stream: work torch has queued on its current stream -- the
                # kernels that are still producing this tensor -- must be finished before the library reads it.
                st = torch.cuda.current_stream(a.device)
                if not st.query():
                    st.synchronize()
            return a, C.c_void_p(a.data_ptr()), tuple(a.shape), a
        arr = np.ascontiguousarray(a, dtype=dtype)
        return arr, C.c_void_p(arr.ctypes.data), arr.shape, None

    @staticmethod
    def _out(shape, like, dtype=np.float32):
        if like is not None:
            import torch
            t = torch.empty(shape, dtype={np.float32: torch.float32, np.int32: torch.int32}[dtype], device=like.device)
            return t, C.c_void_p(t.data_ptr())
        arr = np.empty(shape, dtype=dtype)
        return arr, C.c_void_p(arr.ctypes.data)

    def _grid_op(self, fn, g, *extra):
        keep, p, shape, like = self._in(g)
        out, op = self._out(shape, like)
        self._check(fn(self.h, p, self._dims(shape), *extra, op))
        return out

    # ---- stage 1 (dfields.h:13-22,69) -------------------------------------------------------
    def fastPerim(self, volume):
        return self._grid_op(self.lib.mrb_fast_perim, volume)

    def getsqdist(self, g):
        return self._grid_op(self.lib.mrb_sqdist, g)

    def getsqrt(self, g):
        return self._grid_op(self.lib.mrb_sqrt, g)

    def dfield_gpu(self, volume):
        return self._grid_op(self.lib.mrb_dfield, volume)

    # ---- feature pre-pass (dfields.h:49-67) -------------------------------------------------------
    def getSurface(self, volume):
        keep, p, shape, like = self._in(volume)
        n = C.c_int(0)
        self._check(self.lib.mrb_get_surface(self.h, p, self._dims(shape), None, 0, C.byref(n)))
        out, op = self._out((max(n.value, 1),), like, np.int32)
        if n.value:
            self._check(self.lib.mrb_get_surface(self.h, p, self._dims(shape), op, n.value, C.byref(n)))
        return out[:n.value]

    def getSurfaceNormals(self, volume, surface):
        kv, pv, shape, like = self._in(volume)
        ks, ps, sshape, _ = self._in(surface, np.int32)
        n = int(sshape[0])
        out, op = self._out((max(n, 1), 3), like)
        self._check(self.lib.mrb_surface_normals(self.h, pv, self._dims(shape), ps, n, op))
        return out[:n]

    def getFeatureMap(self, volume, surfaceMap, normals, threshold):
        kv, pv, shape, like = self._in(volume)
        km, pm, _, _ = self._in(surfaceMap)
        kn, pn, nshape, _ = self._in(normals)
        out, op = self._out(shape, like)
        self._check(self.lib.mrb_feature_map(self.h, pv, pm, pn, int(nshape[0]), self._dims(shape), float(threshold), op))
        return out

    # ---- stage 2 (narrowBand.h:18-24, grid.h:95-104) -------------------------------------------
    def erode_grid(self, g):
        return self._grid_op(self.lib.mrb_erode, g)

    def dilate_grid(self, g):
        return self._grid_op(self.lib.mrb_dilate, g)

    def createBands(self, margin, band_size=4.0):
        """-> (band, tight_band)"""
        keep, p, shape, like = self._in(margin)
        band, bp = self._out(shape, like)
        tight, tp = self._out(shape, like)
        self._check(self.lib.mrb_create_bands(self.h, p, self._dims(shape), float(band_size), bp, tp))
        return band, tight

    def findIndexes(self, g):
        keep, p, shape, like = self._in(g)
        n = C.c_int(0)
        self._check(self.lib.mrb_find_indexes(self.h, p, self._dims(shape), None, 0, C.byref(n)))
        out, op = self._out((max(n.value, 1),), like, np.int32)
        if n.value:
            self._check(self.lib.mrb_find_indexes(self.h, p, self._dims(shape), op, n.value, C.byref(n)))
        return out[:n.value]

    def getIndexMap(self, band_or_shape, indexes):
        shape = tuple(band_or_shape) if isinstance(band_or_shape, (tuple, list)) else tuple(band_or_shape.shape)
        keep, p, ishape, like = self._in(indexes, np.int32)
        out, op = self._out(shape, like)
        self._check(self.lib.mrb_index_map(self.h, self._dims(shape), p, int(ishape[0]), op))
        return out

    # ---- stage 3 (primeqp.h:38, quadprog.h:22-26) ------------------------------------------------
    def primeQP(self, volume, margin, band, tight):
        """-> dict(n, lb, ub, x0 (args->x), invdg, jc, ir, pr)   [R = H - diag(H) as CSC]"""
        kv, pv, shape, like = self._in(volume)
        km, pm, _, _ = self._in(margin)
        kb, pb, _, _ = self._in(band)
        kt, pt, _, _ = self._in(tight)
        n = C.c_int(0)
        self._check(self.lib.mrb_find_indexes(self.h, pb, self._dims(shape), None, 0, C.byref(n)))
        n = n.value
        lb, plb = self._out((max(n, 1),), like)
        ub, pub = self._out((max(n, 1),), like)
        x0, px0 = self._out((max(n, 1),), like)
        inv, pinv = self._out((max(n, 1),), like)
        jc, pjc = self._out((n + 1,), like, np.int32)
        nnz = C.c_int(0)
        self._check(self.lib.mrb_prime_qp(self.h, pv, pm, pb, pt, self._dims(shape), n, plb, pub, px0, pinv, pjc,
                                          None, None, 0, C.byref(nnz)))
        ir, pir = self._out((max(nnz.value, 1),), like, np.int32)
        pr, ppr = self._out((max(nnz.value, 1),), like)
        self._check(self.lib.mrb_prime_qp(self.h, pv, pm, pb, pt, self._dims(shape), n, None, None, None, None, pjc,
                                          pir, ppr, nnz.value, C.byref(nnz)))
        return dict(n=n, lb=lb[:n], ub=ub[:n], x0=x0[:n], invdg=inv[:n], jc=jc, ir=ir[:nnz.value], pr=pr[:nnz.value],
                    iter=500)

    def runQP(self, args, featureIndexes=None, USING_FEATURES=False, iters=None):
        n = int(args["n"])
        iters = int(args.get("iter", 500) if iters is None else iters)
        ks = [self._in(args[k], np.int32 if k in ("jc", "ir") else np.float32)
              for k in ("jc", "ir", "pr", "invdg", "lb", "ub", "x0")]
        like = ks[0][3]
        x, px = self._out((max(n, 1),), like)
        if featureIndexes is None or len(featureIndexes) == 0:
            kf, pf, nf = None, None, 0
        else:
            kf, pf, fs, _ = self._in(featureIndexes, np.int32)
            nf = int(fs[0])
        self._check(self.lib.mrb_run_qp(self.h, n, ks[0][1], ks[1][1], ks[2][1], ks[3][1], ks[4][1], ks[5][1], ks[6][1],
                                        iters, pf, nf, 1 if USING_FEATURES else 0, px))
        return x[:n]

    def optimize(self, volume, featureMap=None, USING_FEATURES=False, USING_CUDA=True, band_size=4, iters=500,
                 return_stats=False):
        """quadprog.h:26.  USING_CUDA is accepted for signature parity; the GPU path is the only path."""
        kv, pv, shape, like = self._in(volume)
        pf = None
        if featureMap is not None:
            kf, pf, _, _ = self._in(featureMap)
        F, pF = self._out(shape, like)
        st = L.QPStats()
        self._check(self.lib.mrb_optimize(self.h, pv, pf, self._dims(shape), 1 if USING_FEATURES else 0,
                                          int(band_size), int(iters), pF, C.byref(st)))
        if return_stats:
            return F, _stats(st)
        return F

    # ---- stage 4 (mcubes.h:354, hasher.h:84-106) ---------------------------------------------------
    def _take_mesh(self, m, like):
        nv, nf = m.nverts, m.nfaces
        if m.on_device == 1:
            import torch
            verts = torch.empty((nv, 3), dtype=torch.float32, device=like.device)
            faces = torch.empty((nf, 3), dtype=torch.int32, device=like.device)
            if nv:
                self._check(self.lib.mrb_copy(self.h, verts.data_ptr(), C.cast(m.verts, C.c_void_p), nv * 12))
            if nf:
                self._check(self.lib.mrb_copy(self.h, faces.data_ptr(), C.cast(m.faces, C.c_void_p), nf * 12))
            return verts, faces
        # on_device == 0: malloc'd arrays (freed below); 2: the context's pinned staging (valid until the next call): copy out
        verts = np.ctypeslib.as_array(m.verts, shape=(max(nv, 1), 3))[:nv].copy()
        faces = np.ctypeslib.as_array(m.faces, shape=(max(nf, 1), 3))[:nf].copy()
        self.lib.mrb_mesh_free(C.byref(m))
        return verts, faces

    def mcubes(self, F, isolevel=0.0, filename=None, scale=(1.0, 1.0, 1.0), shift=(0.0, 0.0, 0.0), pad=6):
        """mcubes.h:354 default path; returns (verts [V,3] grid coords, faces [F,3]); writes the PLY if filename."""
        kF, pF, shape, like = self._in(F)
        m = L.Mesh()
        self._check(self.lib.mrb_mcubes(self.h, pF, self._dims(shape), float(isolevel), 1 if like is not None else 2,
                                        C.byref(m)))
        verts, faces = self._take_mesh(m, like)
        if filename is not None:
            self.write_ply(filename, verts, faces, scale, shift, pad)
        return verts, faces

    def mcubes_slab(self, F_slab, x_origin, isolevel=0.0):
        """Marching cubes on planes [x_origin, x_origin + F_slab.shape[0]) of a larger grid: vertices in the whole grid's
        coordinates (bit for bit those of an unsharded run), faces index this slab's own vertex list (mrb_mcubes_slab)."""
        kF, pF, shape, like = self._in(F_slab)
        m = L.Mesh()
        self._check(self.lib.mrb_mcubes_slab(self.h, pF, self._dims(shape), int(x_origin), float(isolevel),
                                             1 if like is not None else 2, C.byref(m)))
        return self._take_mesh(m, like)

    # ---- front end (row f1): binvox completion + observed points -> occupancy volume ------------------------------------
    def _take_volume(self, v, like):
        dims = tuple(v.dims)
        n = dims[0] * dims[1] * dims[2]
        if v.on_device == 1:
            import torch
            vol = torch.empty(dims, dtype=torch.float32, device=like.device)
            self._check(self.lib.mrb_copy(self.h, vol.data_ptr(), C.cast(v.voxels, C.c_void_p), n * 4))
        else:
            vol = np.ctypeslib.as_array(v.voxels, shape=(max(n, 1),))[:n].copy().reshape(dims)
            self.lib.mrb_volume_free(C.byref(v))
        return {"volume": vol, "dims": dims, "pad": v.pad, "occupied_leaves": v.occupied_leaves,
                "scale": np.array(list(v.scale), np.float32), "shift": np.array(list(v.shift), np.float32)}

    def binvox_to_points(self, voxels, dwh, xform, res_factor=1):
        """binvoxToPCL (binvoxToPcl.cpp:105-169): the up-sampled completion as points [n, 3], in the reference's order.
        voxels: decoded binvox bytes (read_binvox), dwh = (depth, width, height), xform = (scale, tx, ty, tz)."""
        v = np.ascontiguousarray(voxels, np.uint8).ravel()
        assert v.size == dwh[0] * dwh[1] * dwh[2]
        p = L.Points()
        self._check(self.lib.mrb_binvox_to_points(self.h, C.c_void_p(v.ctypes.data), (C.c_int * 3)(*dwh), (C.c_float * 4)(*xform),
                                                  int(res_factor), 2, C.byref(p)))
        return np.ctypeslib.as_array(p.xyz, shape=(max(p.n, 1), 3))[:p.n].copy()

    def nn_stats(self, xyz):
        """(dist_total, count) of getDensity / getResolution (binvoxToPcl.cpp:172-201, voxelize.cpp:6-32)."""
        k, p, shape, like = self._in(xyz if _is_torch(xyz) else np.asarray(xyz, np.float32).reshape(-1, 3))
        dt, cnt = C.c_float(0), C.c_int(0)
        self._check(self.lib.mrb_nn_stats(self.h, p, int(shape[0]), C.byref(dt), C.byref(cnt)))
        return np.float32(dt.value), cnt.value

    def voxelize_leaf(self, xyz, leaf, res_factor):
        """voxelize + grid(cloud, vox) + fillGrid + getBinaryVolume for a given leaf size (voxelize.cpp:34-51, grid.cpp:84-121,
        213-265, 322-334).  Returns a dict: volume, dims, pad, scale, shift, occupied_leaves."""
        k, p, shape, like = self._in(xyz if _is_torch(xyz) else np.asarray(xyz, np.float32).reshape(-1, 3))
        v = L.Volume()
        self._check(self.lib.mrb_voxelize_leaf(self.h, p, int(shape[0]), float(leaf), int(res_factor), 1 if like is not None else 2,
                                               C.byref(v)))
        return self._take_volume(v, like)

    def front_end(self, voxels, dwh, xform, observed, res_factor=0, device=False):
        """mesh_reconstruction.cpp:99-120 in one call (mrb_front_end): resolution factor (res_factor < 1: from the densities),
        up-sampled completion, merged cloud, leaf size, occupancy volume.  device=True returns the volume as a CUDA tensor."""
        vb = np.ascontiguousarray(voxels, np.uint8).ravel()
        assert vb.size == dwh[0] * dwh[1] * dwh[2]
        k, p, shape, like = self._in(observed if _is_torch(observed) else np.asarray(observed, np.float32).reshape(-1, 3))
        if device and like is None:
            import torch
            like = torch.empty(0, device="cuda:%d" % self.device)
        v, fi = L.Volume(), L.FrontInfo()
        self._check(self.lib.mrb_front_end(self.h, C.c_void_p(vb.ctypes.data), (C.c_int * 3)(*dwh), (C.c_float * 4)(*xform), p,
                                           int(shape[0]), int(res_factor), 1 if (device and like is not None) else 2, C.byref(v),
                                           C.byref(fi)))
        out = self._take_volume(v, like if device else None)
        for name, _t in L.FrontInfo._fields_:
            val = getattr(fi, name)
            out[name] = np.float32(val) if _t is C.c_float else val
        return out

    def mcubes_features(self, F, surfaceMap, normals, isolevel=0.0, feat_thresh=0.75, corner_thresh=0.8):
        """mcubes(..., USING_FEATURES=true) (mcubes.h:354): feature points, triangle fans and edge flips."""
        kF, pF, shape, like = self._in(F)
        km, pm, _, _ = self._in(surfaceMap)
        kn, pn, nshape, _ = self._in(normals)
        m = L.Mesh()
        self._check(self.lib.mrb_mcubes_features(self.h, pF, pm, pn, int(nshape[0]), self._dims(shape), float(isolevel),
                                                 float(feat_thresh), float(corner_thresh), 1 if like is not None else 2,
                                                 C.byref(m)))
        return self._take_mesh(m, like)

    def hashXYZ(self, xyz, shape):
        kx, px, s, like = self._in(np.asarray(xyz, np.float32).reshape(-1, 3))
        keys = np.empty(s[0], np.uint64)
        self._check(self.lib.mrb_hash_points(self.h, px, int(s[0]), self._dims(shape), C.c_void_p(keys.ctypes.data)))
        return keys

    def write_ply(self, filename, verts, faces, scale=(1.0, 1.0, 1.0), shift=(0.0, 0.0, 0.0), pad=6):
        if _is_torch(verts):
            verts = verts.cpu().numpy()
            faces = faces.cpu().numpy()
        v = np.ascontiguousarray(verts, np.float32)
        f = np.ascontiguousarray(faces, np.int32)
        sc = (C.c_float * 3)(*scale)
        sh = (C.c_float * 3)(*shift)
        self._check(self.lib.mrb_write_ply(str(filename).encode(), C.c_void_p(v.ctypes.data), len(v),
                                           C.c_void_p(f.ctypes.data), len(f), sc, sh, int(pad)))

    # ---- fused ---------------------------------------------------------------------------------
    def reconstruct(self, volume, band_size=4, iters=500, isolevel=0.0, return_stats=False, using_features=False,
                    feature_threshold=0.75, corner_threshold=0.8):
        """optimize + mcubes, device-resident between the stages (mesh_reconstruction.cpp:131-134)."""
        kv, pv, shape, like = self._in(volume)
        opt = L.Options()
        self.lib.mrb_options_default(C.byref(opt))
        opt.band_size, opt.iters, opt.isolevel = int(band_size), int(iters), float(isolevel)
        opt.using_features = 1 if using_features else 0
        opt.feature_threshold, opt.corner_threshold = float(feature_threshold), float(corner_threshold)
        m = L.Mesh()
        st = L.QPStats()
        self._check(self.lib.mrb_reconstruct(self.h, pv, self._dims(shape), C.byref(opt),
                                             1 if like is not None else 2, C.byref(m), C.byref(st)))
        verts, faces = self._take_mesh(m, like)
        if return_stats:
            return verts, faces, _stats(st)
        return verts, faces


def _batch(self, volumes, band_size=4, iters=500, isolevel=0.0):
    """Object-parallel batch: volumes [nobj, nx, ny, nz] -> list of (verts, faces) per object."""
    kv, pv, shape, like = self._in(volumes)
    if len(shape) != 4:
        raise ValueError("volumes must be [nobj, nx, ny, nz]")
    nobj = int(shape[0])
    opt = L.Options()
    self.lib.mrb_options_default(C.byref(opt))
    opt.band_size, opt.iters, opt.isolevel = int(band_size), int(iters), float(isolevel)
    m = L.Mesh()
    voff = (C.c_int * (nobj + 1))()
    foff = (C.c_int * (nobj + 1))()
    self._check(self.lib.mrb_reconstruct_batch(self.h, pv, self._dims(shape[1:]), nobj, C.byref(opt), C.byref(m), voff, foff))
    # No copy: the per-object arrays are views of the library's malloc'd result, which is freed when the last view dies
    # (copying 90 MB of meshes of a 1024-object batch cost 25 ms of a 115 ms call).
    owner = _MeshOwner(self.lib, m)
    verts = _view(m.verts, C.c_float, m.nverts * 3, np.float32, owner).reshape(-1, 3)
    faces = _view(m.faces, C.c_int, m.nfaces * 3, np.int32, owner).reshape(-1, 3)
    return [(verts[voff[o]:voff[o + 1]], faces[foff[o]:foff[o + 1]]) for o in range(nobj)]


class _MeshOwner:
    """Keeps an mrb_mesh with malloc'd arrays alive for the numpy views made of it."""

    def __init__(self, lib, mesh):
        self.lib, self.mesh = lib, mesh

    def __del__(self):
        try:
            self.lib.mrb_mesh_free(C.byref(self.mesh))
        except Exception:
            pass


def _view(ptr, ctype, count, dtype, owner):
    if count == 0:
        return np.empty(0, dtype)
    buf = (ctype * count).from_address(C.addressof(ptr.contents))
    buf._owner = owner                      # the numpy array keeps `buf` alive, `buf` keeps the owner
    return np.frombuffer(buf, dtype=dtype)


Context.reconstruct_batch = _batch


def _stats(st):
    return {k: getattr(st, k) for k, _ in L.QPStats._fields_}




class Pool:
    """Several contexts on ONE GPU, one host thread each: a context is single-threaded and synchronous (like the
    reference's functions), but contexts are independent, so while one thread's object is in its sweep kernel the
    other threads' host->device copies, small kernels and result read-backs overlap with it.  map() keeps the input
    order.  This is host-side plumbing only: every object still goes through mrb_reconstruct with host buffers."""

    def __init__(self, device=0, workers=2):
        self.ctxs = [Context(device) for _ in range(int(workers))]

    def close(self):
        for c in self.ctxs:
            c.close()
        self.ctxs = []

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def map(self, volumes, **kw):
        """reconstruct() of every volume (list of arrays; the same array may appear several times)."""
        import threading
        out = [None] * len(volumes)
        err = []

        def work(w):
            try:
                for i in range(w, len(volumes), len(self.ctxs)):
                    out[i] = self.ctxs[w].reconstruct(volumes[i], **kw)
            except BaseException as e:      # noqa: BLE001 - re-raised in the caller's thread
                err.append(e)

        ts = [threading.Thread(target=work, args=(w,)) for w in range(len(self.ctxs))]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if err:
            raise err[0]
        return out
