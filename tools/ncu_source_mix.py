#!/usr/bin/env python3
"""Instruction mix of a kernel weighted by executed count, from `ncu -i rep --page source --csv` (SASS view).
Usage: ncu_source_mix.py rep.ncu-rep [min_share]   -> opcode histogram, stall totals, shared-memory conflict summary."""
import csv, subprocess, sys, collections, re
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr_i]
ix = {n: i for i, n in enumerate(h)}
tot = 0
mix = collections.Counter()
stalls = collections.Counter()
sh_w = sh_ideal = 0
hot = []
for r in rows[hdr_i + 1:]:
    if len(r) < len(h):
        continue
    n = int(r[ix["Instructions Executed"]] or 0)
    src = r[ix["Source"]].strip()
    m = re.match(r"(@!?U?P\w+\s+)?([A-Z0-9_]+)", src)
    op = m.group(2) if m else "?"
    mix[op] += n
    tot += n
    for k in h:
        if k.startswith("stall_") and "Not Issued" not in k:
            stalls[k] += int(r[ix[k]] or 0)
    sh_w += int(r[ix["L1 Wavefronts Shared"]] or 0)
    sh_ideal += int(r[ix["L1 Wavefronts Shared Ideal"]] or 0)
    hot.append((int(r[ix["# Samples"]] or 0), n, src))
print("instructions executed (warp level): %d" % tot)
for op, n in mix.most_common(30):
    print("  %-10s %12d  %5.1f%%" % (op, n, 100.0 * n / tot))
print("stall samples:", {k: v for k, v in stalls.most_common(10)})
print("shared wavefronts %d, ideal %d (x%.2f)" % (sh_w, sh_ideal, sh_w / max(sh_ideal, 1)))
print("hottest by samples:")
for s, n, src in sorted(hot, reverse=True)[:25]:
    print("  %6d  %10d  %s" % (s, n, src))
