#!/usr/bin/env python3
"""Regenerates tests/golden/golden.npz + golden.json from the CPU oracle (oracle/mrb_oracle.cpp).

The reference itself cannot be built here (Eigen / PCL / Boost are absent) and ships no fixtures for this path, so
these vectors pin the ORACLE (and through it the GPU path) against silent changes; they are not reference outputs.
Small cases are stored in full, the BASELINE config-1 case as counts + SHA-256 of the raw bytes.

    python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from mesh_reconstruction_b200 import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def cases():
    # (a) the 5^3 solid cube of distance_fields/main.cpp:5-23 with the reference's pad of 6: every stage in full
    v = np.zeros((17, 17, 17), np.float32)
    v[6:11, 6:11, 6:11] = 1.0
    yield "cube5", v, dict(band_size=4, iters=25), True
    # (b) a small off-centre ellipsoid in a non-cubic grid, default options, in full
    yield "ellipsoid", synth.ellipsoid((30, 26, 33), (6.0, 4.5, 7.5), (14.5, 12.0, 16.0)), dict(band_size=4, iters=60), True
    # (c) BASELINE config 1: 52^3 sphere R=17, 500 sweeps: counts + hashes
    yield "config1_sphere52", synth.sphere(52, 17.0), dict(band_size=4, iters=500), False


def main():
    arrays, meta = {}, {}
    for name, v, opt, full in cases():
        perim = O.fast_perim(v)
        sq = O.sqdist(perim)
        margin = O.sqrt(sq)
        band, tight = O.create_bands(margin, float(opt["band_size"]))
        idx = O.find_indexes(band)
        F = O.optimize(v, band_size=opt["band_size"], iters=opt["iters"])
        verts, faces = O.mcubes(F, 0.0)
        m = dict(shape=list(v.shape), opt=opt, n_band=int(len(idx)), n_tight=int((tight != 0).sum()), n_verts=int(len(verts)),
                 n_faces=int(len(faces)), sha_margin=sha(margin), sha_band_idx=sha(np.asarray(idx, np.int32)), sha_F=sha(F),
                 sha_verts=sha(verts), sha_faces=sha(np.asarray(faces, np.int32)))
        if full:
            arrays[name + "_volume"] = v
            arrays[name + "_sq"] = sq.astype(np.int32)
            arrays[name + "_band_idx"] = np.asarray(idx, np.int32)
            arrays[name + "_F"] = F
            arrays[name + "_verts"] = verts
            arrays[name + "_faces"] = np.asarray(faces, np.int32)
        meta[name] = m
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **arrays)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps({k: (m["n_band"], m["n_verts"], m["n_faces"]) for k, m in meta.items()}))


if __name__ == "__main__":
    main()
