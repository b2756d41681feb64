"""Committed fixtures (tests/golden/, made by tests/golden/make_golden.py from the oracle): the oracle must still
produce them (CPU), and the GPU path must produce them without the oracle in the loop (GPU)."""
import hashlib
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
META = json.load(open(os.path.join(HERE, "golden", "golden.json")))
DATA = np.load(os.path.join(HERE, "golden", "golden.npz"))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def volume(name, synth):
    if name + "_volume" in DATA:
        return DATA[name + "_volume"]
    assert name == "config1_sphere52"
    return synth.sphere(52, 17.0)


@pytest.mark.parametrize("name", sorted(META))
def test_oracle_reproduces_the_fixtures(O, synth, name):
    m = META[name]
    v = volume(name, synth)
    margin = O.dfield(v)
    band, tight = O.create_bands(margin, float(m["opt"]["band_size"]))
    idx = np.asarray(O.find_indexes(band), np.int32)
    F = O.optimize(v, band_size=m["opt"]["band_size"], iters=m["opt"]["iters"])
    verts, faces = O.mcubes(F, 0.0)
    assert (len(idx), int((tight != 0).sum()), len(verts), len(faces)) == (m["n_band"], m["n_tight"], m["n_verts"], m["n_faces"])
    assert sha(margin) == m["sha_margin"] and sha(idx) == m["sha_band_idx"] and sha(F) == m["sha_F"]
    assert sha(verts) == m["sha_verts"] and sha(np.asarray(faces, np.int32)) == m["sha_faces"]
    if name + "_F" in DATA:
        assert np.array_equal(O.sqdist(O.fast_perim(v)).astype(np.int32), DATA[name + "_sq"])
        assert np.array_equal(idx, DATA[name + "_band_idx"]) and np.array_equal(F, DATA[name + "_F"])
        assert np.array_equal(verts, DATA[name + "_verts"]) and np.array_equal(faces, DATA[name + "_faces"])


def test_cube_fixture_known_answers():
    """The 5^3 cube of distance_fields/main.cpp:5-23: squared distances to the nearest outside voxel, by hand."""
    sq = DATA["cube5_sq"]
    v = DATA["cube5_volume"]
    # fastPerim marks both sides of the surface (the 3x3x3 window is not uniform there): the cube's outer layer (x = 6)
    # and the first empty layer around it (x = 5) are the sites; everything else counts up from them
    assert sq[6, 8, 8] == 0 and sq[5, 8, 8] == 0
    assert sq[7, 8, 8] == 1 and sq[8, 8, 8] == 4                              # inside: centre is two voxels from the shell
    assert sq[4, 8, 8] == 1 and sq[3, 8, 8] == 4 and sq[2, 8, 8] == 9         # outside along an axis
    assert sq[4, 4, 8] == 2 and sq[4, 4, 4] == 3                              # edge and corner directions: site (5,5,8) / (5,5,5)
    assert v.sum() == 125


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(META))
def test_gpu_reproduces_the_fixtures(ctx, synth, name):
    m = META[name]
    v = volume(name, synth)
    margin = ctx.dfield_gpu(v)
    band, tight = ctx.createBands(margin, float(m["opt"]["band_size"]))
    idx = np.asarray(ctx.findIndexes(band), np.int32)
    F = ctx.optimize(v, band_size=m["opt"]["band_size"], iters=m["opt"]["iters"])
    verts, faces = ctx.reconstruct(v, band_size=m["opt"]["band_size"], iters=m["opt"]["iters"])
    assert (len(idx), int((tight != 0).sum()), len(verts), len(faces)) == (m["n_band"], m["n_tight"], m["n_verts"], m["n_faces"])
    assert sha(margin) == m["sha_margin"] and sha(idx) == m["sha_band_idx"] and sha(F) == m["sha_F"]
    assert sha(verts) == m["sha_verts"] and sha(np.asarray(faces, np.int32)) == m["sha_faces"]
