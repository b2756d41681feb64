import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA GPU (B200); run with -m gpu")


@pytest.fixture(scope="session")
def O():
    """The CPU oracle (test infrastructure)."""
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def ctx():
    """GPU context of the product library.  Fails loudly when the library or the GPU is missing."""
    import mesh_reconstruction_b200 as mrb
    c = mrb.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def synth():
    from mesh_reconstruction_b200 import synth
    return synth


def random_blobs(n, seed, count=3, rmin=4.0, rmax=9.0, pad=6):
    from mesh_reconstruction_b200 import synth
    return synth.random_spheres(n, count, rmin, rmax, seed, pad)
