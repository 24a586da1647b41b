#!/usr/bin/env python
"""Dump a SASS region of an .ncu-rep source page with samples and top stall reason."""
import csv, subprocess, sys
rep, kern, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = []; cur = None
for ln in out.splitlines():
    if ln.startswith('"Kernel Name"'):
        cur = []; blocks.append(cur)
    elif cur is not None:
        cur.append(ln)
rows = list(csv.reader(blocks[kern])); hdr = rows[0]; rows = rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for i in range(lo, min(hi, len(rows))):
    r = rows[i]
    st = {h.replace("stall_", ""): int(r[ix[h]] or 0) for h in stall_cols if int(r[ix[h]] or 0) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:2])
    print(f"{i:5d} {int(r[ix['# Samples']] or 0):5d} {int(r[ix['Instructions Executed']] or 0):9d} {r[ix['Source']][:80]:80s} {st}")
