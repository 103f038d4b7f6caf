"""GPU parity tests (-m gpu): the CUDA path, reached through the C ABI, against the oracle,
the committed reference outputs and the reference's golden.bin files.

Bars (BASELINE.json north_star):
  * u64: bit-exact in every mode.
  * fp64 strict mode (B200SPMV_STRICT_ORDER): memcmp-equal to SoftwareSpMV.
  * fp64 merge-path mode: |y - y_ref| <= 2 * n_i * 2^-53 * sum_j |a_ij x_j| per row; rows with
    n_i <= 1 bit-exact.
"""
import os

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

EPS = 2.0 ** -53


@pytest.fixture(scope="module")
def b2():
    import b200spmv
    return b200spmv


# (kernel, options, dcsr_x16).  dcsr_x16 0: never hand a slab to the thread-per-row kernel, so the kernel under test
# sees every matrix; -1: as shipped (measured per slab at setup; slabs too small to time use < 6 entries per non-empty
# row); 96: that rule for every slab; 1 << 20: every slab goes to spmv_dcsr_kernel
ENGINES = {
    "flag": (3, {}, 0),                                               # kernel 3 (CTA tiles), the default
    "stream": (6, {}, 0),                                             # kernel 6 as shipped (single buffer + L2 prefetch)
    "stream-db": (6, {"stream_db": 1, "stream_pf": 0}, 0),            # ... register double buffer
    "stream-v3": (6, {"stream_vchunks": 3, "stream_occ": 1, "stream_pf": 2}, 0),  # ... several chunks per warp
    "default": (3, {}, -1),
    "dcsr-short": (3, {}, 96),
    "dcsr-always": (3, {}, 1 << 20),
}


@pytest.fixture(params=list(ENGINES), ids=list(ENGINES))
def eng(b2, request):
    """every throughput kernel configuration: 3 = flagged segmented sum, one tile per CTA (default),
    6 = the same by persistent warps"""
    kernel, opts, dcsr = ENGINES[request.param]
    e = b2.Engine(0)
    e.set_option("kernel", kernel)
    for k, v in opts.items():
        e.set_option(k, v)
    e.set_option("dcsr_x16", dcsr)
    yield e
    e.close()


def DT(b2, dtype):
    return b2.F64 if dtype == "f64" else b2.U64


def check_f64_tolerance(oracle, m_rows, m_cols, colptr, rowind, vals, x, y_gpu, y_ref):
    abs_sum = np.zeros(m_rows)
    row_nnz = np.zeros(m_rows, dtype=np.uint32)
    oracle.lib.oracle_row_abs_sums_f64(H.C.c_uint32(m_rows), H.C.c_uint32(m_cols), H.ptr(colptr), H.ptr(rowind),
                                       H.ptr(vals), H.ptr(x), H.ptr(abs_sum), H.ptr(row_nnz))
    bound = 2.0 * row_nnz * EPS * abs_sum
    err = np.abs(y_gpu - y_ref)
    assert (err <= bound).all(), "max excess %g" % float((err - bound).max())
    short = row_nnz <= 1
    assert y_gpu[short].tobytes() == y_ref[short].tobytes()


@pytest.mark.parametrize("name", H.ALL_DIRS)
@pytest.mark.parametrize("dtype", ["f64", "u64"])
@pytest.mark.parametrize("xkind", ["ones", "rand"])
def test_repo_matrices_against_reference_outputs(b2, eng, oracle, ref_outputs, name, dtype, xkind):
    """Every matrix directory, native and derived dtype, through b200spmv_exec with host
    buffers (y pre-zeroed, ACCUMULATE = the reference's y += A*x)."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.derive_values(m, H.native_dtype(name), dtype)
    x = H.make_x(m.cols, dtype, xkind)
    want = ref_outputs[f"{name}|{dtype}|{xkind}"]
    eng.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, DT(b2, dtype))
    # strict order: memcmp-equal in both dtypes (HardwareSpMV::compareGolden contract)
    y = np.zeros(m.rows, dtype=vals.dtype)
    eng.exec(x, y, b2.ACCUMULATE | b2.STRICT_ORDER)
    assert y.tobytes() == want.tobytes()
    assert oracle.lib.oracle_compare_golden(H.ptr(want), H.ptr(y), m.rows) == 0
    # merge path
    for flags in (b2.ACCUMULATE, 0, b2.ACCUMULATE | b2.NO_PERSIST):
        y = np.zeros(m.rows, dtype=vals.dtype)
        eng.exec(x, y, flags)
        if dtype == "u64":
            assert y.tobytes() == want.tobytes()
        else:
            check_f64_tolerance(oracle, m.rows, m.cols, m.colptr, m.rowind, vals, x, y, want)


@pytest.mark.parametrize("name", H.FP64_DIRS)
def test_golden_bin_bitwise(b2, eng, name):
    """golden.bin (matrixutils.py:108-113) reproduced bit-for-bit in strict mode."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    golden = np.fromfile(os.path.join(H.MATRICES, name, "golden.bin"), dtype="<f8")
    eng.setup_csc(m.rows, m.cols, m.colptr, m.rowind, m.data_raw.view(np.float64), b2.F64)
    y = np.zeros(m.rows)
    eng.exec(np.ones(m.cols), y, b2.ACCUMULATE | b2.STRICT_ORDER)
    assert y.tobytes() == golden.tobytes()


@pytest.mark.parametrize("name", H.ALL_DIRS)
@pytest.mark.parametrize("dtype", ["f64", "u64"])
def test_nonzero_y_random_values(b2, eng, oracle, name, dtype):
    """y += A*x with a non-zero y, random values and random x: wraps mod 2^64 in mul and add."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.rand_values(m.nnz, dtype, seed=777)
    x = H.make_x(m.cols, dtype, "rand", seed=31337)
    y0 = H.rand_values(m.rows, dtype, seed=5)
    want = oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y0.copy())
    eng.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, DT(b2, dtype))
    y = eng.exec(x, y0.copy(), b2.ACCUMULATE | b2.STRICT_ORDER)
    assert y.tobytes() == want.tobytes()
    y = eng.exec(x, y0.copy(), b2.ACCUMULATE)
    if dtype == "u64":
        assert y.tobytes() == want.tobytes()
    else:
        # the starting y adds one term to each row's sum
        abs_sum = np.zeros(m.rows)
        row_nnz = np.zeros(m.rows, dtype=np.uint32)
        oracle.lib.oracle_row_abs_sums_f64(H.C.c_uint32(m.rows), H.C.c_uint32(m.cols), H.ptr(m.colptr),
                                           H.ptr(m.rowind), H.ptr(vals), H.ptr(x), H.ptr(abs_sum), H.ptr(row_nnz))
        bound = 2.0 * (row_nnz + 1) * EPS * (abs_sum + np.abs(y0))
        assert (np.abs(y - want) <= bound).all()


@pytest.mark.parametrize("name", H.ALL_DIRS)
def test_device_transpose_is_stable(b2, eng, oracle, name):
    """The CSR built on the device equals the stable counting sort of csr2csc.c:11-39."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    dtype = H.native_dtype(name)
    vals = H.derive_values(m, dtype, dtype)
    eng.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, DT(b2, dtype))
    rp, ci, v = eng.get_csr()
    rp_o, ci_o, v_o = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    assert (rp == rp_o).all() and (ci == ci_o).all() and v.tobytes() == v_o.tobytes()


@pytest.mark.parametrize("slabs", [2, 3, 7])
@pytest.mark.parametrize("dtype", ["f64", "u64"])
def test_column_slabs(b2, oracle, slabs, dtype):
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    vals = H.rand_values(m.nnz, dtype, seed=11)
    x = H.make_x(m.cols, dtype, "rand", seed=12)
    y0 = H.rand_values(m.rows, dtype, seed=13)
    want = oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y0.copy())
    e = b2.Engine(0)
    e.set_option("slabs", slabs)
    e.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, DT(b2, dtype))
    assert e.stat("slabs") == slabs
    # strict order stays bit-exact across slabs (slabs are visited in column order)
    assert e.exec(x, y0.copy(), b2.ACCUMULATE | b2.STRICT_ORDER).tobytes() == want.tobytes()
    y = e.exec(x, y0.copy(), b2.ACCUMULATE)
    if dtype == "u64":
        assert y.tobytes() == want.tobytes()
    else:
        # per-row bound of SURVEY 8(c); the starting y adds one term to each row's sum
        abs_sum = np.zeros(m.rows)
        row_nnz = np.zeros(m.rows, dtype=np.uint32)
        oracle.lib.oracle_row_abs_sums_f64(H.C.c_uint32(m.rows), H.C.c_uint32(m.cols), H.ptr(m.colptr),
                                           H.ptr(m.rowind), H.ptr(vals), H.ptr(x), H.ptr(abs_sum), H.ptr(row_nnz))
        assert (np.abs(y - want) <= 2.0 * (row_nnz + 1) * EPS * (abs_sum + np.abs(y0))).all()
    rp, ci, v = e.get_csr()
    rp_o, ci_o, v_o = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    assert (rp == rp_o).all() and (ci == ci_o).all() and v.tobytes() == v_o.tobytes()
    e.close()


def test_marked_matrix_needs_index_mask(b2, eng, oracle):
    """main.cpp:228-229 marks row starts before the HW exec when CMS=1: a drop-in receives
    indices with bit 31 set and must mask them (0x3FFFFFFF, SparseMatrix.cpp:64)."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    inds = m.rowind.copy()
    oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(inds), 0, 31)
    golden = np.fromfile(os.path.join(H.MATRICES, "circuit204", "golden.bin"), dtype="<f8")
    eng.setup_csc(m.rows, m.cols, m.colptr, inds, m.data_raw.view(np.float64), b2.F64, index_mask=0x3FFFFFFF)
    y = eng.exec(np.ones(m.cols), np.zeros(m.rows), b2.ACCUMULATE | b2.STRICT_ORDER)
    assert y.tobytes() == golden.tobytes()


def test_unchecked_indices_are_rejected(b2, eng, oracle):
    """Caller-supplied index arrays are validated on the device before anything indexes with them:
    marked row indices (bits 31 / 30, SparseMatrix.cpp:64) passed with the default mask, a row index >= rows, a
    column pointer that is not non-decreasing, and marked / out-of-range indices into the row-mark entry points
    all give B200SPMV_EINVAL (-1) instead of out-of-bounds device atomics; the context stays usable."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    vals = m.data_raw.view(np.float64)
    marked = m.rowind.copy()
    oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(marked), H.C.c_int(0), H.C.c_int(31))
    with pytest.raises(b2.B200Error, match="index_mask"):
        eng.setup_csc(m.rows, m.cols, m.colptr, marked, vals, b2.F64)
    big = m.rowind.copy()
    big[m.nnz // 2] = m.rows
    with pytest.raises(b2.B200Error, match="-1"):
        eng.setup_csc(m.rows, m.cols, m.colptr, big, vals, b2.F64)
    cp = m.colptr.copy()
    cp[5], cp[6] = cp[6] + 1, cp[5]
    with pytest.raises(b2.B200Error, match="non-decreasing"):
        eng.setup_csc(m.rows, m.cols, cp, m.rowind, vals, b2.F64)
    with pytest.raises(b2.B200Error, match="-1"):
        eng.mark_row_starts(m.rows, big.copy(), 0, 31)
    with pytest.raises(b2.B200Error, match="-1"):
        eng.max_alive(m.rows - 1, m.rowind.copy() | np.uint32(m.rows - 1))
    # and the same context still works
    eng.setup_csc(m.rows, m.cols, m.colptr, marked, vals, b2.F64, index_mask=0x3FFFFFFF)
    golden = np.fromfile(os.path.join(H.MATRICES, "circuit204", "golden.bin"), dtype="<f8")
    y = eng.exec(np.ones(m.cols), np.zeros(m.rows), b2.ACCUMULATE | b2.STRICT_ORDER)
    assert y.tobytes() == golden.tobytes()


def test_error_behaviour(b2, eng):
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "i64"))
    x, y = np.ones(64), np.zeros(64)
    with pytest.raises(b2.B200Error, match="-4"):  # ESTATE: exec before setup
        eng.exec(x, y)
    with pytest.raises(b2.B200Error, match="-1"):  # zero dimension, SparseMatrix.cpp:33-38
        eng.setup_csc(0, 64, m.colptr, m.rowind, m.data_raw, b2.F64)
    eng.set_option("require_aligned", 1)
    raw = np.zeros(64 + 9, dtype=np.float64)
    off = 1 if raw.ctypes.data % 64 == 0 else 0
    mis = raw[off + 0:off + 64] if (raw.ctypes.data + 8 * off) % 64 else raw[off + 1:off + 65]
    eng.set_option("require_aligned", 0)
    eng.setup_csc(m.rows, m.cols, m.colptr, m.rowind, m.data_raw.view(np.float64), b2.F64)
    eng.set_option("require_aligned", 1)
    with pytest.raises(b2.B200Error, match="-6"):  # 64-byte rule, HardwareSpMV.cpp:6-16
        eng.exec(mis, mis.copy())
    assert eng.stat("rows") == 64 and eng.stat("nz") == 64 and eng.stat("gpus") == 1
