#!/usr/bin/env python
"""Summary of an `ncu -i X.ncu-rep --page raw --csv` dump of the SpMV kernel: ncu_summary.py raw.csv"""
import csv
import subprocess
import sys

raw = sys.argv[1]
print(subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "ncu_keys.py"), raw], capture_output=True, text=True).stdout
      .split("----\n")[1], end="")
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
col = {k: i for i, k in enumerate(hdr)}


def val(r, k):
    v, u = float(r[col[k]].replace(",", "")), units[col[k]]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3}.get(u, 1)


print("---- DRAM bytes per launch (%d launches = the slabs of one step)" % (len(rows) - 2))
tot = 0
for r in rows[2:]:
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    tot += rd + wr
    print("  read %.3f Gbyte  write %.3f Mbyte  time %.3f us" % (rd / 1e9, wr / 1e6, val(r, "gpu__time_duration.sum")))
print("  sum over the step: %d bytes (algorithmic 6777995268 on C3)" % tot)
print("---- utilisation metrics > 8 % (first captured launch)")
for k in sorted(col):
    if k.endswith(".avg.pct_of_peak_sustained_elapsed"):
        try:
            v = float(rows[2][col[k]].replace(",", ""))
        except ValueError:
            continue
        if v > 8:
            print("    %6.2f %s" % (v, k))
print("---- warp states, first captured launch (smsp__average_warps_issue_stalled_*_per_issue_active / warp_latency)")
for k in sorted(col):
    if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio"):
        try:
            v = float(rows[2][col[k]].replace(",", ""))
        except ValueError:
            continue
        if v > 0.3:
            print("    %6.2f %s" % (v, k))
