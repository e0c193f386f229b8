#!/usr/bin/env python
"""Top stall-sampled SASS instructions of one captured kernel: python tools/ncu_hot.py X.ncu-rep [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); h, r = rows[0], rows[-1]
for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
          "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
          "lts__throughput.avg.pct_of_peak_sustained_elapsed"]:
    if k in h: print(f"{k:70s} {r[h.index(k)]}")
for i, n in enumerate(h):
    if "issue_stalled" in n and n.endswith("per_issue_active.ratio") and float(r[i] or 0) > 0.05:
        print(f"  {n.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''):24s} {float(r[i]):.3f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); h = rows[1]; data = rows[2:]
iS, iI, isrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
stall = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(x[iS]) for x in data)
print("total samples", tot, "SASS instructions", len(data))
for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:topn]):
    x = data[i]
    st = sorted(((int(x[j]), h[j][6:]) for j in stall), reverse=True)[:2]
    print(f"{i:6d} {int(x[iS]):6d} {x[iI]:>9s}  {x[isrc].strip()[:64]:64s} {st}")
