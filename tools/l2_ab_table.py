#!/usr/bin/env python
"""profiles/r02_l2_ab.txt from the three per-launch ncu CSVs of tools/r02_ncu.sh and the step times of the same commands
run without ncu:  l2_ab_table.py DIR "window step line" "nopersist step line" "1 slab step line" > profiles/r02_l2_ab.txt"""
import collections
import csv
import os
import sys

d = sys.argv[1]
live = sys.argv[2:5]


def load(f):
    rows = [r for r in csv.DictReader(l for l in open(f) if not l.startswith("=="))]
    per = collections.OrderedDict()
    for r in rows:
        per.setdefault(r["ID"], {})[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    return per


print("""Round 2 -- the L2 "vector cache" A/B that north_star asks for, on C3 (uniform 16M x 16M, 32 nnz/row, fp64, y += A*x)
One B200, kernel 3 (spmv_flag_kernel, rows closed by RED.ADD.F64), one step = the slab launches listed; ncu --clock-control none.
Commands: tools/r02_ncu.sh (python tools/prof_case.py c3 kernel=3 persist=1 | persist=0 | slabs=1 persist=1 under
  ncu -k regex:spmv_flag_kernel --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,... ).
Raw CSVs: profiles/r02_l2ab_{3slabs_window,3slabs_nopersist,1slab}.csv.  Step times WITHOUT ncu (same commands, CUDA events):""")
for name, l in zip(("3 slabs + window", "3 slabs, no window", "1 slab + window"), live):
    print("  %-20s %s" % (name, l))
print("""
x gathers are ld.global.nc with an L2::evict_last policy when the window is on (they show up in the evict_last lookup counters) and
plain ld.global.nc without it (evict_normal counters).  hit% below = hits / (hits + misses) of THAT class = the x gathers.

configuration              launch   time us   lts hit rate %   x-gather sectors hit / miss   (hit %)   dram read GB  write MB  l1tex2xbar req %  dram busy %""")
for name, f, cls in (("3 slabs + window", "r02_l2ab_3slabs_window.csv", "evict_last"), ("3 slabs, no window", "r02_l2ab_3slabs_nopersist.csv", "evict_normal"),
                     ("1 slab + window", "r02_l2ab_1slab.csv", "evict_last")):
    rd = wr = t = 0
    for i, m in load(os.path.join(d, f)).items():
        h = m["lts__t_sectors_srcunit_tex_op_read_%s_lookup_hit.sum" % cls]
        mi = m["lts__t_sectors_srcunit_tex_op_read_%s_lookup_miss.sum" % cls]
        print("%-26s slab %s  %8.1f   %6.2f           %11d / %10d   (%5.2f)   %7.3f      %6.1f    %5.2f            %5.2f" % (
            name, i, m["gpu__time_duration.sum"] / 1e3, m["lts__t_sector_hit_rate.pct"], h, mi, 100 * h / (h + mi), m["dram__bytes_read.sum"] / 1e9,
            m["dram__bytes_write.sum"] / 1e6, m["l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed"],
            m.get("dram__cycles_active.avg.pct_of_peak_sustained_elapsed", float("nan"))))
        rd += m["dram__bytes_read.sum"]
        wr += m["dram__bytes_write.sum"]
        t += m["gpu__time_duration.sum"]
    print("%-26s step    %8.1f                    dram read+write %.3f GB (algorithmic 6.778 GB -> x%.2f)\n" % (name, t / 1e3, (rd + wr) / 1e9, (rd + wr) / 6.777995268e9))
