"""mesh_reconstruction_b200 -- B200-native volumetric surface-extraction hot path
(exact EDT -> narrow band -> smoothness QP -> marching cubes) behind the module API of
CRLab/Mesh_Reconstruction.  The compute lives in libmrb.so (hand-written sm_100a CUDA, C-ABI in
include/mrb.h); this package is the Python host-side mirror used by the tests and bench.py."""
from .api import Context, MrbError, Pool  # noqa: F401
from . import synth  # noqa: F401

__all__ = ["Context", "MrbError", "Pool", "synth"]
