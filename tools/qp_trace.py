#!/usr/bin/env python3
"""Phase lengths of the sweep kernels from a MRB_QP_TRACE=1 dump: mean over blocks and sweeps, us.
Tiled kernel (header flag 1) stamps: 0 sweep start (thread 0), 1 neighbours' flags seen (warp 0), 2 thread 0 past its interior
bricks, 3 last warp past its interior bricks, 4 last warp past the wait for the flags, 5 last warp past its rim bricks,
6 past the block barrier, 7 past the release store.
On-chip kernel (flag 0) stamps: 0 sweep start, 1 thread 0 past its first brick + send, 2 thread 0 past its last brick, 3 last warp
past the first pass, 4 last warp past the halo fetch, 5 thread 0 past the block barrier that ends the sweep."""
import sys
import numpy as np
raw = open(sys.argv[1] if len(sys.argv) > 1 else "qp_trace.bin", "rb").read()
grid, ns, ne, kind = np.frombuffer(raw[:16], np.int32)
nst = int(grid) * int(ns) * int(ne) * 8
t = np.frombuffer(raw[16:16 + nst], np.uint64).reshape(grid, ns, ne).astype(np.int64)
extra = np.frombuffer(raw[16 + nst:], np.int32)
sl = (t[:, 1:, 0] - t[:, :-1, 0]) / 1000.0
print("blocks %d, sweeps %d, kernel: %s" % (grid, ns, "tiled" if kind == 1 else "on-chip"))
def row(name, x):
    print("%-46s mean %6.2f  p10 %6.2f  p90 %6.2f" % (name, x.mean(), np.percentile(x, 10), np.percentile(x, 90)))
row("sweep length (start -> next start)", sl)
d = lambda a, b: (t[:, :, b] - t[:, :, a]) / 1000.0
if kind == 1:
    row("start -> flags seen by warp 0", d(0, 1))
    row("start -> thread 0 past interior", d(0, 2))
    row("start -> last warp past interior", d(0, 3))
    row("last warp: wait for the flags", d(3, 4))
    row("last warp: rim bricks", d(4, 5))
    row("last warp: block barrier", d(5, 6))
    row("release store (fence + st)", d(6, 7))
    rel = t[:, :-1, 7]; seen = t[:, 1:, 1]
    print("own release(k-1) -> own flags seen(k): mean %.2f us" % ((seen - rel).mean() / 1000.0))
else:
    row("thread 0: first brick + send", d(0, 1))
    row("thread 0: all its bricks", d(0, 2))
    row("last warp: first pass", d(0, 3))
    row("last warp: first pass -> halo fetched", d(3, 4))
    row("thread 0: last brick -> past the end barrier", d(2, 5))
print("spread of sweep starts across blocks: mean %.2f us" % ((t[:, :, 0].max(0) - t[:, :, 0].min(0)).mean() / 1000.0))

if kind == 1 and len(extra) >= 2 * grid:
    rim, size = extra[:grid], extra[grid:2 * grid]
    wait = d(3, 4).mean(1); rimt = d(4, 5).mean(1); intt = d(0, 3).mean(1); seen = d(0, 1).mean(1)
    print("per block: rim bricks min %d median %d p90 %d max %d; bricks %d..%d" % (rim.min(), np.median(rim), np.percentile(rim, 90), rim.max(), size.min(), size.max()))
    order = np.argsort(wait)
    print("blocks that wait least for their neighbours (they set the pace):")
    print("  block  rim  interior  int-phase  flags-seen  wait  rim-phase")
    for b in order[:12]:
        print("  %5d %4d %8d %10.2f %11.2f %5.2f %10.2f" % (b, rim[b], size[b] - rim[b], intt[b], seen[b], wait[b], rimt[b]))
    print("blocks that wait most:")
    for b in order[-6:]:
        print("  %5d %4d %8d %10.2f %11.2f %5.2f %10.2f" % (b, rim[b], size[b] - rim[b], intt[b], seen[b], wait[b], rimt[b]))
    for lo, hi in ((0, 200), (200, 300), (300, 400), (400, 480), (480, 2000)):
        m = (rim >= lo) & (rim < hi)
        if m.any():
            print("rim in [%d,%d): %3d blocks, int-phase %.2f, wait %.2f, rim-phase %.2f, start->rim done %.2f" %
                  (lo, hi, m.sum(), intt[m].mean(), wait[m].mean(), rimt[m].mean(), (intt[m] + wait[m] + rimt[m]).mean()))
