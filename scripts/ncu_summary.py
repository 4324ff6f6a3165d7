#!/usr/bin/env python
"""Summarise an ncu report (raw page) and/or a launch list CSV into text for profiles/.
usage: ncu_summary.py raw <report.ncu-rep> | launches <launches.csv>"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum ", "smsp__inst_executed.sum ",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ",
        "smsp__average_warps_issue_stalled", "sm__cycles_elapsed.max ", "smsp__cycles_active.avg "]


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        print(f"== kernel: {name}")
        for h, u, v in zip(hdr, units, r):
            if any((h + " ").startswith(k) for k in KEYS):
                print(f"{h:90s} {v} {u}")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        agg[r[ki][:70]][0] += 1
        agg[r[ki][:70]][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"total device time {tot:.3f} ms over {sum(v[0] for v in agg.values())} launches (ncu: cold cache, serialised)")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:12.3f} ms {100 * t / tot:6.2f}%  n={n:4d}  avg {t / n:10.3f} ms  {k}")


def traffic(path, batch, source):
    """profiles/traffic.json for bench.py's roofline.traffic: DRAM bytes of the captured blind-rotate launch"""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, r = rows[0], rows[1], rows[2]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = hdr.index(name)
        tot += float(r[i].replace(",", "")) * scale[units[i]]
    print(json.dumps({"kernel": r[hdr.index("Kernel Name")], "batch": int(batch), "dram_bytes_per_launch": tot,
                      "source": source}))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        {"raw": raw, "launches": launches}[sys.argv[1]](sys.argv[2])
