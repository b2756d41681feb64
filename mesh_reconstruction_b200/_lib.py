"""ctypes loader for libmrb.so (the C-ABI declared in include/mrb.h).

The library is built in-tree by `make -C mesh_reconstruction_b200/csrc` (or __graft_entry__.build()).
There is no Python or CPU fallback: if the shared library is missing, or no GPU is usable,
importing / creating a context raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("MRB_LIBRARY") or os.path.join(_HERE, "libmrb.so")   # MRB_LIBRARY: A/B runs of two builds

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)


class QPStats(C.Structure):
    _fields_ = [("n_band", C.c_int), ("n_tight", C.c_int), ("iters", C.c_int), ("last_max_delta", C.c_float),
                ("ms_dfield", C.c_float), ("ms_band", C.c_float), ("ms_qp", C.c_float), ("ms_mc", C.c_float),
                ("ms_qp_kernel", C.c_float)]


class Mesh(C.Structure):
    _fields_ = [("verts", c_float_p), ("faces", c_int_p), ("nverts", C.c_int), ("nfaces", C.c_int),
                ("on_device", C.c_int)]


class Points(C.Structure):
    _fields_ = [("xyz", c_float_p), ("n", C.c_int), ("on_device", C.c_int)]


class Volume(C.Structure):
    _fields_ = [("voxels", c_float_p), ("dims", C.c_int * 3), ("scale", C.c_float * 3), ("shift", C.c_float * 3),
                ("pad", C.c_int), ("occupied_leaves", C.c_int), ("on_device", C.c_int)]


class FrontInfo(C.Structure):
    _fields_ = [("observed_dist_total", C.c_float), ("observed_count", C.c_int), ("completed_dist_total", C.c_float),
                ("completed_count", C.c_int), ("observed_density", C.c_float), ("completed_density", C.c_float),
                ("res_factor", C.c_int), ("n_predicted", C.c_int), ("n_cloud", C.c_int), ("leaf", C.c_float)]


class Options(C.Structure):
    _fields_ = [("using_features", C.c_int), ("band_size", C.c_int), ("iters", C.c_int), ("isolevel", C.c_float),
                ("feature_threshold", C.c_float), ("corner_threshold", C.c_float)]


# every symbol include/mrb.h declares: name -> (restype, argtypes)
_VP = C.c_void_p
SYMBOLS = {
    "mrb_ctx_create": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "mrb_ctx_destroy": (None, [_VP]),
    "mrb_ctx_stream": (_VP, [_VP]),
    "mrb_last_error": (C.c_char_p, []),
    "mrb_version": (C.c_char_p, []),
    "mrb_ctx_launch_count": (C.c_uint64, [_VP]),
    "mrb_copy": (C.c_int, [_VP, _VP, _VP, C.c_size_t]),
    "mrb_fast_perim": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_sqdist": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_sqrt": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_dfield": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_get_surface": (C.c_int, [_VP, _VP, c_int_p, _VP, C.c_int, c_int_p]),
    "mrb_surface_normals": (C.c_int, [_VP, _VP, c_int_p, _VP, C.c_int, _VP]),
    "mrb_feature_map": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int, c_int_p, C.c_float, _VP]),
    "mrb_erode": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_dilate": (C.c_int, [_VP, _VP, c_int_p, _VP]),
    "mrb_create_bands": (C.c_int, [_VP, _VP, c_int_p, C.c_float, _VP, _VP]),
    "mrb_find_indexes": (C.c_int, [_VP, _VP, c_int_p, _VP, C.c_int, c_int_p]),
    "mrb_index_map": (C.c_int, [_VP, c_int_p, _VP, C.c_int, _VP]),
    "mrb_find_indexes_noalloc": (C.c_int, [_VP, c_int_p, _VP, C.c_int, c_int_p, _VP]),
    "mrb_prime_qp": (C.c_int, [_VP, _VP, _VP, _VP, _VP, c_int_p, C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, _VP,
                               C.c_int, c_int_p]),
    "mrb_run_qp": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, _VP, C.c_int, _VP, C.c_int, C.c_int, _VP]),
    "mrb_optimize": (C.c_int, [_VP, _VP, _VP, c_int_p, C.c_int, C.c_int, C.c_int, _VP, C.POINTER(QPStats)]),
    "mrb_mcubes": (C.c_int, [_VP, _VP, c_int_p, C.c_float, C.c_int, C.POINTER(Mesh)]),
    "mrb_mesh_free": (None, [C.POINTER(Mesh)]),
    "mrb_mcubes_features": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int, c_int_p, C.c_float, C.c_float, C.c_float, C.c_int,
                                      C.POINTER(Mesh)]),
    "mrb_hash_points": (C.c_int, [_VP, _VP, C.c_int, c_int_p, _VP]),
    "mrb_write_ply": (C.c_int, [C.c_char_p, _VP, C.c_int, _VP, C.c_int, c_float_p, c_float_p, C.c_int]),
    "mrb_options_default": (None, [C.POINTER(Options)]),
    "mrb_reconstruct": (C.c_int, [_VP, _VP, c_int_p, C.POINTER(Options), C.c_int, C.POINTER(Mesh),
                                  C.POINTER(QPStats)]),
    "mrb_qp_plan_create": (C.c_int, [_VP, _VP, c_int_p, C.c_int, C.POINTER(_VP)]),
    "mrb_qp_plan_destroy": (None, [_VP]),
    "mrb_qp_plan_info": (C.c_int, [_VP, c_int_p, c_int_p, c_int_p, c_int_p]),
    "mrb_qp_plan_plane_offsets": (C.c_int, [_VP, c_int_p]),
    "mrb_qp_plan_init_x": (C.c_int, [_VP, _VP]),
    "mrb_qp_plan_sweep": (C.c_int, [_VP, _VP, _VP, C.c_int, C.c_int]),
    "mrb_qp_plan_assemble": (C.c_int, [_VP, _VP, _VP]),
    "mrb_qp_plan_solve_sharded": (C.c_int, [_VP, C.c_int, C.c_int, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP), c_int_p,
                                            C.c_int]),
    "mrb_qp_plan_last_kernel_ms": (C.c_float, [_VP]),
    "mrb_qp_plan_create_slab": (C.c_int, [_VP, _VP, c_int_p, C.c_int, C.POINTER(_VP)]),
    "mrb_qp_plan_solve_sharded_sub": (C.c_int, [_VP, C.c_int, C.c_int, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP), C.c_int, C.c_int,
                                                c_int_p, C.c_int]),
    "mrb_qp_plan_assemble_slab": (C.c_int, [_VP, _VP, C.c_int, C.c_int, _VP]),
    "mrb_mcubes_slab": (C.c_int, [_VP, _VP, c_int_p, C.c_int, C.c_float, C.c_int, C.POINTER(Mesh)]),
    "mrb_binvox_to_points": (C.c_int, [_VP, _VP, c_int_p, c_float_p, C.c_int, C.c_int, C.POINTER(Points)]),
    "mrb_points_free": (None, [C.POINTER(Points)]),
    "mrb_nn_stats": (C.c_int, [_VP, _VP, C.c_int, c_float_p, c_int_p]),
    "mrb_voxelize_leaf": (C.c_int, [_VP, _VP, C.c_int, C.c_float, C.c_int, C.c_int, C.POINTER(Volume)]),
    "mrb_volume_free": (None, [C.POINTER(Volume)]),
    "mrb_front_end": (C.c_int, [_VP, _VP, c_int_p, c_float_p, _VP, C.c_int, C.c_int, C.c_int, C.POINTER(Volume),
                                C.POINTER(FrontInfo)]),
    "mrb_reconstruct_batch": (C.c_int, [_VP, _VP, c_int_p, C.c_int, C.POINTER(Options), C.POINTER(Mesh), c_int_p,
                                        c_int_p]),
}

_lib = None


def load():
    """Loads libmrb.so and binds every declared symbol.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError("libmrb.so is not built (%s). Run `make -C mesh_reconstruction_b200/csrc` or "
                          "__graft_entry__.build(); there is no CPU fallback." % SO_PATH)
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
