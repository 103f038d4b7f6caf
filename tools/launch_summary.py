#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launch_summary.py launches.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ik, im, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) <= iv or r[im] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"^void |<unnamed>::|\(.*$", "", r[ik])
    tot[name] += float(r[iv].replace(",", ""))
    cnt[name] += 1
allns = sum(tot.values())
print("%-64s %5s %14s %7s" % ("kernel", "n", "total_ns", "share"))
for k, v in tot.most_common():
    print("%-64s %5d %14d %6.1f%%" % (k, cnt[k], v, 100 * v / allns))
step = {k: v for k, v in tot.items() if k.startswith(("spmv_", "vec_accumulate", "x_prefetch"))}
s = sum(step.values())
flag = sum(v for k, v in step.items() if k.startswith("spmv_flag_kernel"))
nflag = sum(cnt[k] for k in step if k.startswith("spmv_flag_kernel"))
if s:
    print("within the SpMV steps: spmv_flag_kernel %.1f%%, everything else (fix-up, accumulate) %.1f%%; mean flag launch %.1f us"
          % (100 * flag / s, 100 * (s - flag) / s, flag / max(1, nflag) / 1e3))
