#!/usr/bin/env python
"""Kernel sweep on one B200 (numpy + ctypes only): for a few workloads, set the matrix up once and time the SpMV
step (b200spmv_exec_device, CUDA events, y += A*x) under different launch-time options of kernels 3 and 6.
Timing only -- parity is pytest's job -- but each timed configuration's y is compared with the first
configuration's y (same matrix, same x, fp64 re-association only), so a variant that skips work shows up.

  python tools/sweep_k6.py [--quick] [--only NAME] > gpurun_out/sweep.txt
  B200SPMV_LIB=.../build/libb200spmv_abl3.so python tools/sweep_k6.py --ablate   (measurement builds, no check)
"""
import argparse
import importlib.util
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200spmv", os.path.join(ROOT, "spmv-vector-cache_b200", "b200spmv.py"))
b2 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b2)
SEED = 0x5EED5EED
PEAK = 6548.2
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass

ap = argparse.ArgumentParser()
ap.add_argument("--quick", action="store_true")
ap.add_argument("--ablate", action="store_true", help="measurement build: kernel 6 variants only, no result check")
ap.add_argument("--only", default="")
ap.add_argument("--variants", default="", help="comma-separated variant names to run (default: all)")
ap.add_argument("--steps", type=int, default=20)
args = ap.parse_args()

# (name, kind, rows, npr, slabs list)
WORK = [
    ("uniform C3", 0, 1 << 24, 32, [3]),
    ("band16", 3, 1 << 24, 16, [1, 3]),
    ("banded11+5", 2, 1 << 24, 5, [1, 2, 3]),
    ("uniform 16M x 8", 0, 1 << 24, 8, [3]),
    ("rmat16M", 1, 1 << 24, 32, [0]),
    ("uniform 4M (x in L2)", 0, 1 << 22, 32, [1]),
]
if args.quick:
    WORK = WORK[:3]
if args.only:
    WORK = [w for w in WORK if args.only in w[0]]

VARIANTS = [
    ("k3", {"kernel": 3}),
    ("k3 nopersist", {"kernel": 3, "_flags": 4}),
    ("k3 persist", {"kernel": 3, "persist": 1}),
    ("k3 pf0", {"kernel": 3, "prefetch_tiles": 0}),
    ("k3 pf592", {"kernel": 3, "prefetch_tiles": 592}),
    ("k6 db occ2", {"kernel": 6, "stream_db": 1, "stream_pf": 0, "stream_occ": 0, "stream_vchunks": 1}),
    ("k6 nodb occ3 pf1", {"kernel": 6, "stream_db": 0, "stream_pf": 1, "stream_occ": 0, "stream_vchunks": 1}),
    ("k6 nodb occ3 pf1 v8", {"kernel": 6, "stream_db": 0, "stream_pf": 1, "stream_occ": 0, "stream_vchunks": 8}),
    ("k6 nodb pf1 nopersist", {"kernel": 6, "stream_db": 0, "stream_pf": 1, "stream_occ": 0, "stream_vchunks": 1, "_flags": 4}),
    ("k3 never dcsr", {"kernel": 3, "dcsr_x16": 0}),
    ("k3 never dcsr nopersist", {"kernel": 3, "dcsr_x16": 0, "_flags": 4}),
    ("k6 never dcsr", {"kernel": 6, "stream_db": 0, "stream_pf": 1, "stream_occ": 0, "stream_vchunks": 1, "dcsr_x16": 0}),
    ("k6 never dcsr nopersist", {"kernel": 6, "stream_db": 0, "stream_pf": 1, "stream_occ": 0, "stream_vchunks": 1, "dcsr_x16": 0, "_flags": 4}),
]
if args.variants:
    VARIANTS = [v for v in VARIANTS if v[0] in args.variants.split(",")]
if args.ablate:
    VARIANTS = [v for v in VARIANTS if v[0] in ("k6 db occ2", "k6 nodb occ3", "k6 nodb occ3 pf2")]

print("lib %s  peak %.1f GB/s  steps %d (y += A*x, fp64)" % (b2.LIB_PATH, PEAK, args.steps), flush=True)
for name, kind, rows, npr, slab_list in WORK:
    for slabs in slab_list:
        e = b2.Engine(0)
        e.set_option("slabs", slabs)
        t0 = time.perf_counter()
        e.setup_synthetic(kind, rows, rows, npr, SEED, b2.F64)
        setup_s = time.perf_counter() - t0
        e.fill_x_synthetic(SEED + 1)
        nnz = e.stat("nz")
        balg = 12 * nnz + 4 * (rows + 1) + 16 * rows
        ns = e.stat("slabs")
        print("== %s: rows %d nnz %d slabs %d (setup %.2f s, window measured %s) B_alg %.3f GB; per slab nnz/nonempty/maxrow: %s" % (
            name, rows, nnz, ns, setup_s, "on" if e.stat("persist") else "off", balg / 1e9,
            " ".join("%d/%d/%d" % (e.stat("slab_nnz_%d" % i), e.stat("slab_rows_%d" % i), e.stat("slab_maxrow_%d" % i)) for i in range(ns))),
            flush=True)
        ref = None
        for vname, opts in VARIANTS:
            e.set_option("dcsr_x16", -1)
            e.set_option("persist", -1)
            e.set_option("prefetch_tiles", 148)
            fl = b2.ACCUMULATE | opts.get("_flags", 0)
            for k, v in opts.items():
                if not k.startswith("_"):
                    e.set_option(k, v)
            e.fill_y_zero()
            e.exec_device(3, fl)
            e.fill_y_zero()
            ms = e.exec_device(args.steps, fl)
            y = None if args.ablate else e.get_y()
            e.set_option("profile", 1)
            e.exec_device(5, fl)
            kns, kn = e.stat("merge_kernel_ns"), e.stat("merge_kernel_launches")
            per_slab = [e.stat("slab_ns_%d" % i) * 1e-6 / 5 for i in range(ns)]
            e.set_option("profile", 0)
            kms = kns * 1e-6 / kn * ns if kn else 0.0
            ok = ""
            if y is not None:
                if ref is None:
                    ref = y
                else:
                    scale = np.maximum(np.abs(ref), 1e-300)
                    ok = "  maxrel %.1e" % float(np.max(np.abs(y - ref) / scale))
            best, med = float(ms.min()), float(np.median(ms))
            print("  %-22s step best %.4f med %.4f ms  %6.1f GNNZ/s | kernel %.4f ms %6.1f GNNZ/s frac %.3f | slabs %s%s" % (
                vname, best, med, nnz / med / 1e6, kms, nnz / kms / 1e6 if kms else 0, balg / (kms * 1e-3) / 1e9 / PEAK if kms else 0,
                " ".join("%.3f" % t for t in per_slab), ok), flush=True)
        e.close()
