#!/usr/bin/env python
"""Summarise ncu outputs brought back from the GPU box into small text files under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv [--last N]   # per-kernel share of a step
  python tools/ncu_summary.py full gpurun_out/prof.ncu-rep                   # key counters per captured launch
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg"]


def short(name):
    name = re.sub(r"\(.*", "", name).replace("void ", "").replace("sdb::", "")
    return name[:70]


def launches(path, last=None):
    with open(path) as f:
        rows = list(csv.DictReader(l for l in f if l.startswith('"')))
    if last:
        rows = rows[-last:]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        k = (short(r["Kernel Name"]), r["Grid Size"], r["Block Size"])
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"]) / 1e3
    tot = sum(v[1] for v in agg.values())
    print(f"# {len(rows)} launches, {tot:.1f} us total (ncu per-launch times: cold-cache, serialised -> compare SHARES)")
    print(f"{'us':>10} {'share':>6} {'n':>4}  kernel grid block")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]:10.1f} {100 * v[1] / tot:5.1f}% {v[0]:4d}  {k[0]} {k[1]} {k[2]}")
    by = collections.defaultdict(float)
    for k, v in agg.items():
        by[k[0]] += v[1]
    print("\n# by kernel")
    for k, v in sorted(by.items(), key=lambda kv: -kv[1]):
        print(f"{v:10.1f} {100 * v / tot:5.1f}%  {k}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("kernel:", short(d.get("Kernel Name", "?")), "grid", d.get("Grid Size"), "block", d.get("Block Size"))
        for k in KEYS:
            if k in d:
                print(f"  {k:70s} {d[k]:>16s} {units[hdr.index(k)]}")
        print()


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else None
        launches(sys.argv[2], last)
    else:
        full(sys.argv[2])
