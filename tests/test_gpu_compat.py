"""The C++ host side (compat shim with the reference's own function names) drives the hot path the way
mesh_reconstruction.cpp:119-134 does; its PLY must equal the oracle's byte for byte."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMPAT = os.path.join(ROOT, "mesh_reconstruction_b200", "compat")


def test_compat_pipeline_matches_oracle(O, synth, tmp_path):
    subprocess.check_call(["make", "-C", COMPAT, "-s"])
    out = tmp_path / "compat.ply"
    res = subprocess.run([os.path.join(COMPAT, "test_compat"), "40", "11", str(out)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    # same progress lines as the reference prints (primeqp.cpp:221, quadprog.cpp:103)
    assert "quadratic program ready" in res.stdout and "quadratic program finished" in res.stdout
    v = synth.sphere(40, 11.0)
    F = O.optimize(v)
    verts, faces = O.mcubes(F, 0.0)
    ref = tmp_path / "ref.ply"
    O.write_ply(str(ref), verts, faces, (0.01, 0.01, 0.01), (-0.2, 0.1, 0.3), 6)
    assert out.read_bytes() == ref.read_bytes()
    band, _ = O.create_bands(O.dfield(v), 4.0)
    assert ("band %d " % int(band.sum())) in res.stdout
