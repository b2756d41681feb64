#!/usr/bin/env python3
"""Regenerates tests/golden/ref_front_golden.npz + .json from the REFERENCE'S OWN front-end code.

oracle/build_ref.sh compiles binvoxToPCL/binvoxToPcl.cpp, confidence_pcl/*.cpp, voxelize/voxelize.cpp and grid/grid.cpp
unmodified from /root/reference against the PCL stand-ins of oracle/ref_stubs/pcl/ (exact-NN kd-tree, restated VoxelGrid)
into oracle/_ref/libmrb_ref_front.so; this script runs mesh_reconstruction.cpp:99-120 through it on the scenes of
tests/front_cases.py (binvox files written to a temporary directory) and stores what comes back.  /root/reference does not
exist on the GPU box, so the vectors are committed and the tests read only them.

    python tests/golden/make_ref_front_golden.py
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from mesh_reconstruction_b200 import binvox  # noqa: E402
from oracle import ref as R  # noqa: E402
from front_cases import scenes, pack_volume  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    meta, arrays = {}, {}
    with tempfile.TemporaryDirectory() as d:
        for name, (vox, dwh, xf, obs) in scenes().items():
            path = os.path.join(d, name + ".binvox")
            binvox.write(path, vox, dwh, xf)
            rv, rdwh, rxf = R.read_binvox(path)
            assert np.array_equal(rv, vox) and tuple(rdwh) == tuple(dwh)
            m = {"dwh": list(dwh), "xform": [float(np.float32(t)) for t in rxf], "points": {}}
            for res in (1, 2, 3):
                p = R.binvox_to_points(path, res)
                m["points"][str(res)] = {"n": len(p), "sha256": sha(p)}
                if res == 2:
                    m["density_res2"] = float(R.density(p))
                    m["resolution_res2"] = float(R.resolution(p))
            m["observed_density"] = float(R.density(obs))
            m["res_factor"] = int(R.resolution_factor(path, obs))
            fe = R.front_end(path, obs)
            m.update(n_predicted=fe["n_predicted"], n_cloud=fe["n_cloud"], leaf=float(fe["leaf"]), dims=list(fe["dims"]), pad=fe["pad"],
                     scale=[float(t) for t in fe["scale"]], shift=[float(t) for t in fe["shift"]], occupied=int(fe["volume"].sum()))
            arrays[name + "/voxels"] = vox
            arrays[name + "/observed"] = obs
            arrays[name + "/volume_bits"] = pack_volume(fe["volume"])
            meta[name] = m
            print(name, m["res_factor"], m["n_cloud"], m["dims"], m["occupied"])
    np.savez_compressed(os.path.join(HERE, "ref_front_golden.npz"), **arrays)
    with open(os.path.join(HERE, "ref_front_golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
