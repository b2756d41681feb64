"""The fused multi-GPU path with world > 1, driven through torchrun from pytest (skipped on a box with one GPU):
k_qp_bricks_tiled<...,Sharded=true> with peer stores into the neighbours' halo slots, system-scope flags, slab-local plans,
per-slab marching cubes and the mesh merge -- every rank's slab of F and the merged mesh against its own single-GPU result."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ngpus():
    import torch
    return torch.cuda.device_count()


def run_check(world, args, mode, port):
    env = dict(os.environ, SLAB_MODE=mode)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "slab_check2.py")] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    last = [l for l in r.stdout.splitlines() if l.startswith('{"slab_check2"')]
    assert last, r.stdout[-2000:]
    out = json.loads(last[-1])
    assert out["slab_check2"] == "ok" and out["world"] == world, r.stdout[-3000:]


@pytest.mark.parametrize("mode", ["local", "sharded"])
def test_fused_slab_path_world2(mode):
    if ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    # config 3 (256^3) with 500 sweeps, then 120^3 and 52^3 spheres with an odd number of sweeps (the result lies in the other buffer)
    run_check(2, ["3", 2, 500], mode, 29541)
    run_check(2, ["120,52", 2, 37], mode, 29542)


def test_fused_slab_path_world4():
    if ngpus() < 4:
        pytest.skip("needs 4 GPUs")
    run_check(4, ["3", 2, 500], "local", 29543)
