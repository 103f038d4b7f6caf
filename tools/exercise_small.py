#!/usr/bin/env python
"""A few small invocations of every device path (numpy + ctypes only, no torch; seconds on a B200): strict against
throughput kernels (u64: bit for bit), host-buffer against device-resident execution, sharded and slab-aligned
layouts, the download pipeline.  A quick self-check after a kernel change, before the full `pytest -m gpu`."""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200spmv", os.path.join(ROOT, "spmv-vector-cache_b200", "b200spmv.py"))
b2 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b2)
SEED = 0x5EED5EED
bad = 0


def check(name, ok):
    global bad
    print("%-70s %s" % (name, "ok" if ok else "MISMATCH"), flush=True)
    bad += not ok


# (kind, rows, npr, slabs, options): the 1 MB vectors of the first three switch the download pipeline on
cases = [(0, 1 << 17, 8, 2, {}), (1, 1 << 17, 8, 3, {}), (2, (1 << 17) + 777, 5, 2, {"dcsr_x16": 0}),
         (0, 1 << 17, 4, 1, {"pin_host": 1}),
         (0, 5000, 3, 4, {"prefetch_x": 1}), (1, 4096, 16, 2, {"dcsr_maxrow": 1 << 30, "dcsr_x16": 1 << 20}),
         (2, 4099, 5, 3, {"world": 3, "rank": 1, "own_slabs": 1}), (0, 3000, 32, 1, {"kernel": 3}), (0, 3000, 32, 2, {"kernel": 6, "stream_db": 0, "stream_pf": 2})]
for kind, rows, npr, slabs, opts in cases:
    for dtype in (b2.U64, b2.F64):
        e = b2.Engine(0)
        e.set_option("slabs", slabs)
        for k, v in opts.items():
            e.set_option(k, v)
        e.setup_synthetic(kind, rows, rows, npr, SEED, dtype)
        e.fill_x_synthetic(SEED + 1)
        x = e.get_x()
        npdt = b2.np_dtype(dtype)
        want = np.zeros(rows, dtype=npdt)
        e.exec(x, want, b2.STRICT_ORDER)
        y = np.full(rows, 3, dtype=npdt)
        e.exec(x, y, 0)
        rb, re = e.stat("row_begin"), e.stat("row_end")
        tag = "kind %d rows %d npr %d slabs %d %s %s" % (kind, rows, npr, e.stat("slabs"), "u64" if dtype == b2.U64 else "f64", opts)
        if dtype == b2.U64:
            check(tag + " exec", y[rb:re].tobytes() == want[rb:re].tobytes())
        else:
            check(tag + " exec", np.allclose(y[rb:re], want[rb:re], rtol=0, atol=1e-9))
        y2 = y.copy()
        e.exec(x, y2, b2.ACCUMULATE)
        if dtype == b2.U64:
            check(tag + " accumulate", y2[rb:re].tobytes() == (y[rb:re] + want[rb:re]).tobytes())
        e.exec_device(2, 0)
        d = e.get_y()
        if dtype == b2.U64:
            check(tag + " exec_device", d[rb:re].tobytes() == want[rb:re].tobytes())
        if "world" not in opts:
            e.power_iterate(2, False, 0)
        e.close()
print("SMALL CASES %s" % ("OK" if bad == 0 else "FAILED"))
sys.exit(1 if bad else 0)
