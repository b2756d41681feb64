"""Shared scenes of the front-end tests (row f1): binvox completion + partial observed cloud, small enough for the CPU checkers."""
import numpy as np

from mesh_reconstruction_b200 import synth


def scenes():
    """name -> (voxels, dwh, xform, observed)"""
    out = {}
    out["sphere24"] = synth.completion_scene(24, "sphere", seed=1)
    out["box20_gap"] = synth.completion_scene(20, "box", seed=2, z_shift=1.5)                 # fillGrid closes real gaps here
    out["blobs28"] = synth.completion_scene(28, "blobs", seed=3, z_shift=0.8, xform=(1.7, 3.0, -2.5, 0.25))
    out["sparse_view"] = synth.completion_scene(22, "sphere", seed=4, observed_step=1.3, view_fraction=0.3)   # completion denser: res 1
    # ragged dims and voxel bytes other than 0 / 1 (the weight is 2 * byte - 1, binvoxToPcl.cpp:148)
    vox, dwh, xf, obs = synth.completion_scene(18, "box", seed=5)
    g = vox.reshape(18, 18, 18)[:15, :17, :].copy()                 # [x][z][y] -> depth 15, height 17, width 18
    g[g > 0] = 3
    out["ragged_bytes"] = (g.ravel().copy(), (15, 18, 17), xf, obs)
    return out


def pack_volume(vol):
    return np.packbits(np.asarray(vol) > 0)
