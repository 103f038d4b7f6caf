"""Randomised parity (seeded): CSC matrices the fixtures do not cover -- unsorted and duplicate row indices
inside a column (the reference loop, software/SoftwareSpMV.cpp:59-64, does not care about either), empty
rows and columns, rectangular shapes, one dense row / one dense column, forced column slabs -- through
every throughput kernel and the strict kernel, against the oracle.  u64 and strict fp64: bit-exact;
fp64 throughput kernels: per-row bound 2 n_i 2^-53 sum|a_ij x_j|."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu
EPS = 2.0 ** -53


def random_csc(rng, rows, cols, nnz, hub_row=False, hub_col=False):
    c = rng.integers(0, cols, nnz)
    r = rng.integers(0, rows, nnz)
    if hub_row:
        r[: nnz // 3] = rng.integers(0, rows)
    if hub_col:
        c[nnz // 3: 2 * nnz // 3] = rng.integers(0, cols)
    # leave some columns and rows empty
    c = np.where(c % 7 == 3, (c + 1) % cols, c)
    r = np.where(r % 5 == 2, (r + 1) % rows, r)
    order = np.argsort(c, kind="stable")  # column-major, row indices in arbitrary order inside a column
    c, r = c[order], r[order]
    colptr = np.zeros(cols + 1, dtype=np.uint32)
    np.add.at(colptr, c + 1, 1)
    colptr = np.cumsum(colptr).astype(np.uint32)
    return colptr, r.astype(np.uint32)


CASES = [(1, 1, 1), (1, 50, 40), (50, 1, 40), (7, 3000, 500), (3000, 7, 500), (257, 255, 4000), (2048, 2048, 2049),
         (1000, 1000, 30000), (5000, 3000, 100), (4097, 4099, 20000)]


@pytest.mark.parametrize("case", range(len(CASES)))
@pytest.mark.parametrize("dtype", ["u64", "f64"])
def test_random_matrices_every_kernel(oracle, case, dtype):
    import b200spmv as b2
    rows, cols, nnz = CASES[case]
    rng = np.random.default_rng(1000 + case)
    colptr, rowind = random_csc(rng, rows, cols, nnz, hub_row=case % 2 == 0, hub_col=case % 3 == 0)
    vals = H.rand_values(nnz, dtype, 50 + case)
    x = H.make_x(cols, dtype, "rand", 60 + case)
    npdt = np.uint64 if dtype == "u64" else np.float64
    y0 = H.make_x(rows, dtype, "rand", 70 + case)
    want = y0.copy()
    oracle.spmv_csc(dtype, rows, cols, colptr, rowind, vals, x, want)
    abs_sum = np.zeros(rows)
    row_nnz = np.zeros(rows, dtype=np.uint32)
    if dtype == "f64":
        oracle.lib.oracle_row_abs_sums_f64(H.C.c_uint32(rows), H.C.c_uint32(cols), H.ptr(colptr), H.ptr(rowind),
                                           H.ptr(vals), H.ptr(x), H.ptr(abs_sum), H.ptr(row_nnz))
    for kernel, dcsr, slabs in ((3, 0, 1), (3, 0, 3), (6, 0, 1), (6, 0, 2), (6, 0, 5), (6, 1 << 20, 2), (6, 96, 4)):
        e = b2.Engine(0)
        e.set_option("kernel", kernel)
        e.set_option("dcsr_x16", dcsr)
        e.set_option("slabs", min(slabs, cols))
        e.setup_csc(rows, cols, colptr, rowind, vals, b2.U64 if dtype == "u64" else b2.F64)
        y = y0.copy()
        e.exec(x, y, b2.ACCUMULATE | b2.STRICT_ORDER)
        # strict order is the reference's order whatever the slab count: slabs ascend by column and every pass
        # continues from the y the previous one left, i.e. the same sequence ((y + p1) + p2) + ...
        assert y.tobytes() == want.tobytes(), (kernel, slabs)
        y = y0.copy()
        e.exec(x, y, b2.ACCUMULATE)
        if dtype == "u64":
            assert y.tobytes() == want.tobytes(), (kernel, dcsr, slabs)
        else:
            bound = 2.0 * (row_nnz + 1.0) * EPS * (abs_sum + np.abs(y0))
            assert (np.abs(y - want) <= bound).all(), (kernel, dcsr, slabs, float((np.abs(y - want) - bound).max()))
        # y = A*x (no accumulate): rows without entries must come back as zero
        y = y0.copy()
        e.exec(x, y, 0)
        z = np.zeros(rows, dtype=npdt)
        oracle.spmv_csc(dtype, rows, cols, colptr, rowind, vals, x, z)
        if dtype == "u64":
            assert y.tobytes() == z.tobytes(), (kernel, dcsr, slabs)
        else:
            assert (np.abs(y - z) <= 2.0 * row_nnz * EPS * abs_sum).all() and (y[row_nnz == 0] == 0).all()
        e.close()
