#!/usr/bin/env python
"""Top stall locations of an ncu source page: usage ncu_hot.py <report.ncu-rep> [top]"""
import csv, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr)]
tot = sum(int(r[ci["# Samples"]]) for r in body)
print("total samples", tot, "instructions", len(body))
agg = collections.Counter()
for r in body:
    for s in stall_cols:
        agg[s] += int(r[ci[s]])
print({k: round(v / tot, 3) for k, v in agg.most_common()})
# by opcode
op = collections.Counter(); opn = collections.Counter()
for r in body:
    m = r[ci["Source"]].split()
    o = m[1] if m[0].startswith("@") else m[0]
    op[o.split(".")[0]] += int(r[ci["# Samples"]]); opn[o.split(".")[0]] += int(r[ci["Instructions Executed"]])
print("samples by opcode:", [(k, round(v / tot, 3), opn[k]) for k, v in op.most_common(16)])
idx = sorted(range(len(body)), key=lambda i: -int(body[i][ci["# Samples"]]))[:top]
for i in sorted(idx):
    r = body[i]
    st = {s[6:]: int(r[ci[s]]) for s in stall_cols if int(r[ci[s]])}
    prev = body[i - 1][ci["Source"]].strip() if i else ""
    print(f"{i:5d} {int(r[ci['# Samples']]):6d} {r[ci['Source']].strip():60s} {dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])}   prev: {prev[:40]}")
