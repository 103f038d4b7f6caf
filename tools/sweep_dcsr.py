#!/usr/bin/env python
"""Where does the thread-per-row kernel (spmv_dcsr_kernel) beat the flagged kernel?  Short-row matrices on one B200,
each timed with option dcsr_x16 = 0 (never), 1024 (always, rows up to dcsr_maxrow) and -1 (the choice setup measured).  numpy + ctypes only.
  python tools/sweep_dcsr.py > gpurun_out/sweep_dcsr.txt"""
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200spmv", os.path.join(ROOT, "spmv-vector-cache_b200", "b200spmv.py"))
b2 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b2)

CASES = [("uniform 4M x 4M", 0, 1 << 22, npr, 1) for npr in (1, 2, 3, 4, 5, 6, 8, 12)]
CASES += [("uniform 16M x 16M", 0, 1 << 24, npr, 3) for npr in (4, 8, 12, 16, 24)]
CASES += [("banded(11)+5 16M", 2, 1 << 24, 5, s) for s in (2, 3, 4, 6)]
CASES += [("banded(11)+5 128M", 2, 1 << 27, 5, s) for s in (1, 2, 3, 4, 8, 18)]
CASES += [("rmat 16M", 1, 1 << 24, 8, s) for s in (1, 3)]
for name, kind, rows, npr, slabs in CASES:
    e = b2.Engine(0)
    e.set_option("slabs", slabs)
    e.setup_synthetic(kind, rows, rows, npr, 0x5EED5EED, b2.F64)
    e.fill_x_synthetic(0x5EED5EEE)
    nnz, ns = e.stat("nz"), e.stat("slabs")
    per = " ".join("%.2f/%d" % (e.stat("slab_nnz_%d" % i) / max(1, e.stat("slab_rows_%d" % i)), e.stat("slab_maxrow_%d" % i)) for i in range(min(ns, 4)))
    res = []
    for dx in (0, 1024, -1):
        e.set_option("dcsr_x16", dx)
        e.fill_y_zero()
        e.exec_device(3, b2.ACCUMULATE)
        ms = e.exec_device(10, b2.ACCUMULATE)
        res.append(float(np.median(ms)))
    print("%-20s npr %2d slabs %2d  entries per non-empty row / longest row, per slab: %-44s flag %.4f ms %6.1f GNNZ/s | dcsr %.4f ms %6.1f GNNZ/s | dcsr/flag %.2f | as set up (%s; tuning %.1f ms) %.4f ms" % (
        name, npr, ns, per, res[0], nnz / res[0] / 1e6, res[1], nnz / res[1] / 1e6, res[1] / res[0],
        "".join("d" if e.stat("slab_dcsr_%d" % i) else "f" for i in range(ns)), e.stat("tunetime_ns") / 1e6, res[2]), flush=True)
    e.close()
