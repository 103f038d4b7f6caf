#!/usr/bin/env python
"""Top stall sites of an `ncu -i X.ncu-rep --page source --csv` dump (SASS view): ncu_hot.py source.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
execd = sum(int(r[col["Instructions Executed"]] or 0) for r in body)
print("samples %d, warp instructions executed %d, static instructions %d" % (tot, execd, len(body)))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[col[h]] or 0) for r in body) for h in stalls}
print("stall mix: " + "  ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v * 100 > tot))
top = sorted(body, key=lambda r: -int(r[col["# Samples"]] or 0))[:n]
for r in sorted(top, key=lambda r: int(r[col["Address"]], 16) if r[col["Address"]].startswith("0x") else 0):
    s = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print("%6.2f%%  %-8s %-70s %s" % (100.0 * s / max(tot, 1), r[col["Address"]][-5:], r[col["Source"]][:70], " ".join("%s:%d" % (b, a) for a, b in why if a)))
