#!/usr/bin/env python
"""One small device-resident run of a named case, for ncu (numpy + ctypes only):

  python tools/prof_case.py CASE [key=value ...]     e.g.  prof_case.py band16 kernel=6 stream_db=0

CASE: c3 | band16 | banded | banded3 | rmat16 | u4m ; options are b200spmv_set_option pairs (slabs=.. before setup).
Runs 2 warm-up steps and 2 timed steps of y += A*x and prints the step time."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200spmv", os.path.join(ROOT, "spmv-vector-cache_b200", "b200spmv.py"))
b2 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b2)

CASES = {"c3": (0, 1 << 24, 32, 0), "band16": (3, 1 << 24, 16, 1), "banded": (2, 1 << 24, 5, 1), "banded3": (2, 1 << 24, 5, 3),
         "rmat16": (1, 1 << 24, 32, 0), "u4m": (0, 1 << 22, 32, 1)}
kind, rows, npr, slabs = CASES[sys.argv[1]]
opts = dict(a.split("=") for a in sys.argv[2:])
flags = b2.ACCUMULATE | int(opts.pop("flags", 0))
e = b2.Engine(0)
e.set_option("slabs", int(opts.pop("slabs", slabs)))
for k, v in opts.items():
    e.set_option(k, int(v))
e.setup_synthetic(kind, rows, rows, npr, 0x5EED5EED, b2.F64)
e.fill_x_synthetic(0x5EED5EEE)
e.set_option("graph", 0)
e.exec_device(2, flags)
ms = e.exec_device(2, flags)
print("%s %s: slabs %d, step %.4f ms, %.1f GNNZ/s" % (sys.argv[1], opts, e.stat("slabs"), float(ms.min()), e.stat("nz") / float(ms.min()) / 1e6))
e.close()
