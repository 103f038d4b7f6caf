#!/usr/bin/env python
"""Regenerate tests/golden/ from the reference (runs only in the build container).

1. Copies the reference's matrix-directory fixtures (data files, not source) from
   /root/reference/matrices/<name>/ to tests/golden/matrices/<name>/ unchanged, incl.
   golden.bin where the reference ships one (matrices/matrixutils.py:108-113).
2. Runs the UNMODIFIED reference SoftwareSpMV (oracle/_ref/libref_{f64,u64}.so, built by
   oracle/Makefile from /root/reference/software/*.cpp) on every directory in its native
   dtype and in the derived other dtype (SURVEY.md 8d "Repo-matrix runs"), with x = 1 and
   with a seeded random x, and stores the result vectors as tests/golden/ref_outputs.npz
   plus their SURVEY 8(c) word hashes in tests/golden/ref_hashes.json.

The GPU box has no /root/reference: tests read only what this script committed.
"""
import ctypes as C
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_MATRICES = "/root/reference/matrices"
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402

FP64_DIRS = ["circuit204", "i64", "i1k", "i64k", "row64k"]
U64_DIRS = ["circuit204-uint64", "dia64-uint64", "i64-uint64", "i1024-uint64", "rowvec64-uint64"]


def main():
    out_m = os.path.join(HERE, "matrices")
    os.makedirs(out_m, exist_ok=True)
    for name in FP64_DIRS + U64_DIRS:
        dst = os.path.join(out_m, name)
        os.makedirs(dst, exist_ok=True)
        for f in sorted(os.listdir(os.path.join(REF_MATRICES, name))):
            if f.endswith(".bin"):
                shutil.copyfile(os.path.join(REF_MATRICES, name, f), os.path.join(dst, f))
                os.chmod(os.path.join(dst, f), 0o644)
    shutil.copyfile(os.path.join(REF_MATRICES, "mtx", "circuit204.mtx"),
                    os.path.join(HERE, "circuit204.mtx"))
    os.chmod(os.path.join(HERE, "circuit204.mtx"), 0o644)

    ref = helpers.RefLib()
    outputs, hashes = {}, {}
    for name in FP64_DIRS + U64_DIRS:
        m = helpers.load_matrix_dir(os.path.join(out_m, name))
        native = "u64" if name.endswith("uint64") else "f64"
        for dtype in ("f64", "u64"):
            vals = helpers.derive_values(m, native, dtype)
            for xkind in ("ones", "rand"):
                x = helpers.make_x(m.cols, dtype, xkind, seed=1234)
                y = np.zeros(m.rows, dtype=np.float64 if dtype == "f64" else np.uint64)
                ref.spmv(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y)
                key = f"{name}|{dtype}|{xkind}"
                outputs[key] = y
                hashes[key] = "%016x" % helpers.hash_words(y)
    np.savez_compressed(os.path.join(HERE, "ref_outputs.npz"), **outputs)
    with open(os.path.join(HERE, "ref_hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1, sort_keys=True)
    print("wrote", len(outputs), "reference outputs")


if __name__ == "__main__":
    main()
