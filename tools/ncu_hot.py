#!/usr/bin/env python
"""Top stall instructions of an .ncu-rep source page (SASS level), aggregated into coarse regions by opcode context."""
import csv, subprocess, sys
rep = sys.argv[1]; kern = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
# split per kernel
blocks = []; cur = None
for ln in lines:
    if ln.startswith('"Kernel Name"'):
        cur = []; blocks.append(cur)
    elif cur is not None:
        cur.append(ln)
rows = list(csv.reader(blocks[kern]))
hdr = rows[0]; rows = rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix["# Samples"]] or 0) for r in rows)
print("total samples", tot, "instructions", len(rows))
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[ix[h]] or 0) for r in rows) for h in stall_cols}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(rows)), key=lambda i: -int(rows[i][ix["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = rows[i]
    st = {h.replace("stall_", ""): int(r[ix[h]] or 0) for h in stall_cols if int(r[ix[h]] or 0) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{i:5d} {int(r[ix['# Samples']]):6d} {r[ix['Source']][:70]:70s} {st}")
