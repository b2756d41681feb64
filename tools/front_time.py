#!/usr/bin/env python3
"""Times the GPU front end (mrb_front_end and its pieces) on synthetic completion scenes; checks the small one against the oracle.
usage: python tools/front_time.py [n ...]      (binvox edge lengths, default 40 96)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mesh_reconstruction_b200 as mrb  # noqa: E402
from mesh_reconstruction_b200 import synth  # noqa: E402


def timed(fn, reps=5):
    fn()
    t = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t) / reps * 1e3, out


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [40, 96]
    with mrb.Context(0) as ctx:
        for n in sizes:
            vox, dwh, xf, obs = synth.completion_scene(n, "blobs", seed=5, z_shift=1.0, observed_step=0.3)
            ms, fe = timed(lambda: ctx.front_end(vox, dwh, xf, obs, device=True))
            res = fe["res_factor"]
            ms_p, pts = timed(lambda: ctx.binvox_to_points(vox, dwh, xf, res))
            allp = np.concatenate([obs, pts])
            ms_n, st = timed(lambda: ctx.nn_stats(allp))
            ms_v, _ = timed(lambda: ctx.voxelize_leaf(allp, fe["leaf"], res))
            print("binvox %d^3: observed %d points, res factor %d, cloud %d points, volume %s (%d leaves)" %
                  (n, len(obs), res, fe["n_cloud"], fe["dims"], fe["occupied_leaves"]))
            print("  mrb_front_end %.2f ms (host binvox bytes + observed points in, volume stays on the device)" % ms)
            print("  pieces with host in/out: binvox_to_points(res %d) %.2f ms, nn_stats(%d points) %.2f ms, voxelize_leaf %.2f ms" %
                  (res, ms_p, len(allp), ms_n, ms_v))
            if n <= 48:
                from oracle import oracle as O
                t = time.perf_counter()
                ref = O.front_end(vox, dwh, xf, obs)
                cpu = (time.perf_counter() - t) * 1e3
                ok = np.array_equal(fe["volume"].cpu().numpy(), ref["volume"]) and fe["leaf"] == ref["leaf"]
                print("  oracle (CPU restatement, brute-force NN, %d threads) %.0f ms; volume and leaf size bit-identical: %s" %
                      (os.cpu_count(), cpu, ok))


if __name__ == "__main__":
    main()
