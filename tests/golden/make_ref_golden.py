#!/usr/bin/env python3
"""Regenerates tests/golden/ref_golden.npz + ref_golden.json from the REFERENCE'S OWN CODE.

oracle/build_ref.sh compiles the reference's hot-path translation units unmodified from /root/reference (against
stand-in Eigen / Boost / PCL headers, oracle/ref_stubs/) into oracle/_ref/libmrb_ref.so; this script runs them on seeded
synthetic volumes and stores what they return.  Unlike golden.npz (made from our restated oracle) these vectors are
reference outputs: band 4 and 500 sweeps are hard-coded there (quadprog.cpp:112, primeqp.cpp:219).  /root/reference does
not exist on the GPU box, so the vectors are committed and the tests read only them.

    python tests/golden/make_ref_golden.py
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

from mesh_reconstruction_b200 import synth  # noqa: E402
from oracle import ref as R  # noqa: E402

PLY_FRAME = dict(scale=(0.0125, 0.025, 0.0375), shift=(-1.0, 2.0, 0.5), pad=6)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def cases():
    """name, volume, store in full?, feature path too?"""
    v = np.zeros((17, 17, 17), np.float32)
    v[6:11, 6:11, 6:11] = 1.0                                          # the 5^3 cube of distance_fields/main.cpp:5-23, pad 6
    yield "cube5", v, True, True
    yield "ellipsoid", synth.ellipsoid((30, 26, 33), (6.0, 4.5, 7.5), (14.5, 12.0, 16.0)), True, False
    yield "config1_sphere52", synth.sphere(52, 17.0), False, False      # BASELINE config 1
    yield "config1_spheres_seed1", synth.random_spheres(52, 3, 6, 12, 1), False, True
    yield "box_rot56", synth.box(56, (13, 10, 8), None, 20.0, 35.0), False, True
    yield "cfg2_small64", synth.union(synth.box(64, (18, 14, 10), None, 20.0, 35.0), synth.sphere(64, 14.0)), False, True


def run_case(v, features):
    out = {}
    perim = R.fast_perim(v)
    sq = R.sqdist(perim)
    margin = R.sqrt(sq)
    band, tight = R.create_bands(margin, 4.0)
    idx = R.find_indexes(band)
    q = R.prime_qp(v, margin, band, tight)
    F = R.optimize(v)
    verts, faces = R.mcubes(F, 0.0)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "m.ply")
        R.mcubes_ply(p, F, 0.0, **PLY_FRAME)
        ply = open(p, "rb").read()
    out.update(sq=sq.astype(np.int32), margin=margin, band_idx=np.asarray(idx, np.int32), n_tight=int((tight != 0).sum()),
               lb=q["lb"], ub=q["ub"], x0=q["x0"], invdg=q["invdg"], jc=q["jc"], ir=q["ir"], pr=q["pr"],
               F=F, verts=verts, faces=np.asarray(faces, np.int32), ply=np.frombuffer(ply, np.uint8))
    if features:
        s = R.get_surface(v)
        n = R.surface_normals(v, s)
        smap = R.index_map(v.shape, s)
        fm = R.feature_map(v, smap, n, 0.75)
        Ff = R.optimize(v, fm, True)
        vf, ff = R.mcubes(Ff, 0.0, smap, n, 0.75, 0.8)
        out.update(surface=np.asarray(s, np.int32), normals=n, feature_map=fm, F_feat=Ff, verts_feat=vf,
                   faces_feat=np.asarray(ff, np.int32))
    return out


def main():
    if not R.available():
        sys.exit("oracle/_ref is not built: needs /root/reference (run oracle/build_ref.sh)")
    arrays, meta = {}, {"_provenance": R.lib().ref_build_info().decode(), "_ply_frame": PLY_FRAME}
    for name, v, full, features in cases():
        r = run_case(v, features)
        m = dict(shape=list(v.shape), features=features, n_band=int(len(r["band_idx"])), n_tight=r["n_tight"],
                 n_verts=int(len(r["verts"])), n_faces=int(len(r["faces"])), nnz=int(len(r["ir"])))
        if features:
            m.update(n_surface=int(len(r["surface"])), n_feature_voxels=int(r["feature_map"].sum()),
                     n_verts_feat=int(len(r["verts_feat"])), n_faces_feat=int(len(r["faces_feat"])))
        for k, a in r.items():
            if isinstance(a, np.ndarray):
                m["sha_" + k] = sha(a)
        if full:
            arrays[name + "_volume"] = v
            for k in ("sq", "band_idx", "F", "verts", "faces", "invdg", "jc", "ir", "pr"):
                arrays[name + "_" + k] = r[k]
        meta[name] = m
    np.savez_compressed(os.path.join(HERE, "ref_golden.npz"), **arrays)
    with open(os.path.join(HERE, "ref_golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps({k: (m["n_band"], m["n_verts"], m["n_faces"]) for k, m in meta.items() if not k.startswith("_")}))


if __name__ == "__main__":
    main()
