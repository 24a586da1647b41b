#!/usr/bin/env python
"""Samples per CUDA source line from an .ncu-rep (cuda,sass view), top N lines."""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname = "?"; hdr = None; agg = {}
for row in csv.reader(out.splitlines()):
    if not row:
        continue
    if row[0] == "File Path":
        fname = row[1].split("/")[-1]; continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Line No":
        hdr = row; ix = {h: i for i, h in enumerate(hdr)}; si = hdr.index("# Samples"); continue
    if hdr is None or row[0] == "":
        continue
    try:
        n = int(row[si] or 0)
    except ValueError:
        continue
    if n == 0:
        continue
    st = {}
    for h in hdr:
        if h.startswith("stall_") and "Not Issued" not in h:
            v = int(row[ix[h]] or 0)
            if v:
                st[h[6:]] = v
    key = (fname, int(row[0]))
    a = agg.setdefault(key, [0, row[1].strip()[:70], {}])
    a[0] += n
    for k, v in st.items():
        a[2][k] = a[2].get(k, 0) + v
tot = sum(a[0] for a in agg.values())
print("total samples", tot)
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = dict(sorted(a[2].items(), key=lambda kv: -kv[1])[:3])
    print(f"{a[0]:6d} {100*a[0]/tot:5.1f}% {f}:{ln:<4d} {a[1]:70s} {st}")
