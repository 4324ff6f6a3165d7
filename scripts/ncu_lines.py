#!/usr/bin/env python
"""Per-source-line stall samples of an ncu report (CUDA view): usage ncu_lines.py <report.ncu-rep> [top]
Prints, for every source file of the kernel, the lines with the most warp-stall samples and their stall reasons."""
import csv, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
rows = list(csv.reader(out.splitlines()))
files, cur, hdr = {}, None, None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1]; files[cur] = []; continue
    if len(r) == 2: continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if cur and hdr and len(r) == len(hdr) and r[2] == "-":     # CUDA source rows (SASS rows carry an address)
        files[cur].append(r)
ci = {h: i for i, h in enumerate(hdr)}
si = hdr.index("# Samples")
stall = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si]) for f in files.values() for r in f)
print("total samples", tot)
allrows = [(int(r[si]), f.split("/")[-1], r) for f, rs in files.items() for r in rs if int(r[si])]
allrows.sort(key=lambda t: -t[0])
keep = sorted(allrows[:top], key=lambda t: (t[1], int(t[2][0])))
for n, f, r in keep:
    st = sorted(((int(r[i]), nme) for i, nme in stall if int(r[i])), reverse=True)[:4]
    ie = r[ci["Instructions Executed"]]
    print(f"{f}:{r[0]:>4} {n:7d} {100*n/tot:5.1f}%  inst {ie:>12}  {r[1].strip()[:90]:90s} {[(nme, v) for v, nme in st]}")
