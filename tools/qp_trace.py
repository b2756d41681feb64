#!/usr/bin/env python3
"""Phase lengths of the on-chip sweep kernel from a MRB_QP_TRACE=1 dump (see k_qp_bricks_onchip): mean over blocks and sweeps, us.
Stamps: 0 sweep start (thread 0), 1 thread 0 past its first brick + send, 2 thread 0 past its last brick, 3 last warp past the
first pass, 4 last warp past the halo fetch, 5 thread 0 past the block barrier that ends the sweep."""
import sys
import numpy as np
raw = open(sys.argv[1] if len(sys.argv) > 1 else "qp_trace.bin", "rb").read()
grid, ns, ne, _ = np.frombuffer(raw[:16], np.int32)
t = np.frombuffer(raw[16:], np.uint64).reshape(grid, ns, ne).astype(np.int64)
sl = (t[:, 1:, 0] - t[:, :-1, 0]) / 1000.0
print("blocks %d, sweeps %d" % (grid, ns))
def row(name, x):
    print("%-46s mean %6.2f  p10 %6.2f  p90 %6.2f" % (name, x.mean(), np.percentile(x, 10), np.percentile(x, 90)))
row("sweep length (start -> next start)", sl)
d = lambda a, b: (t[:, :, b] - t[:, :, a]) / 1000.0
row("thread 0: first brick + send", d(0, 1))
row("thread 0: all its bricks", d(0, 2))
row("last warp: first pass", d(0, 3))
row("last warp: first pass -> halo fetched", d(3, 4))
row("thread 0: last brick -> past the end barrier", d(2, 5))
print("spread of sweep starts across blocks: mean %.2f us" % ((t[:, :, 0].max(0) - t[:, :, 0].min(0)).mean() / 1000.0))
