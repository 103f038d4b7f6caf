"""CPU suite: pins oracle/spmv_oracle.c against the reference's golden vectors, its
known-answer tests and (where oracle/_ref is built) the reference's own compiled code."""
import json
import os

import numpy as np
import pytest

import helpers as H


@pytest.mark.parametrize("name", H.FP64_DIRS)
def test_oracle_reproduces_golden_bin(oracle, name):
    """matrices/<name>/golden.bin = A*ones (matrixutils.py:108-113); the reference's own
    acceptance test is a bytewise compare (HardwareSpMV.cpp:37-39, chisel/vecdiff.sh)."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    golden = np.fromfile(os.path.join(H.MATRICES, name, "golden.bin"), dtype="<f8")
    y = np.zeros(m.rows)
    oracle.spmv_csc("f64", m.rows, m.cols, m.colptr, m.rowind, m.data_raw.view(np.float64),
                    np.ones(m.cols), y)
    assert y.tobytes() == golden.tobytes()
    assert oracle.lib.oracle_compare_golden(H.ptr(golden), H.ptr(y), m.rows) == 0


def test_survey_hashes_of_uint64_set(oracle):
    """hashes of the u64 results for x = 1 from the reference build."""
    want = {"circuit204-uint64": 0x2631b0afec9686f7, "dia64-uint64": 0x89cb7fa853e9c800,
            "rowvec64-uint64": 0x3fe56df2f3141ec0, "i64-uint64": 0x3f61a84c831c4940,
            "i1024-uint64": 0x6934c93b389a9400}
    for name, h in want.items():
        m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
        y = np.zeros(m.rows, dtype=np.uint64)
        oracle.spmv_csc("u64", m.rows, m.cols, m.colptr, m.rowind, m.data_raw, np.ones(m.cols, np.uint64), y)
        assert H.hash_words(y) == h
        assert int(oracle.lib.oracle_hash_words(H.ptr(y), H.C.c_uint64(m.rows))) == h


@pytest.mark.parametrize("name", H.ALL_DIRS)
@pytest.mark.parametrize("dtype", ["f64", "u64"])
@pytest.mark.parametrize("xkind", ["ones", "rand"])
def test_oracle_matches_committed_reference_outputs(oracle, ref_outputs, name, dtype, xkind):
    """tests/golden/ref_outputs.npz holds outputs of the UNMODIFIED reference SoftwareSpMV
    (made by tests/golden/make_golden.py).  Bit-exact in both dtypes."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.derive_values(m, H.native_dtype(name), dtype)
    x = H.make_x(m.cols, dtype, xkind)
    y = np.zeros(m.rows, dtype=vals.dtype)
    oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y)
    assert y.tobytes() == ref_outputs[f"{name}|{dtype}|{xkind}"].tobytes()
    with open(os.path.join(H.GOLDEN, "ref_hashes.json")) as f:
        assert "%016x" % H.hash_words(y) == json.load(f)[f"{name}|{dtype}|{xkind}"]


def test_known_answers_from_rtl_frontend_tests(oracle):
    """chisel/tests/TestSpMVFrontend.scala:121-144 identity 64x64 with x=1..64 => sum 2080;
    :148-183 1x64 row vector => y0 = 2080 (integer semantics on the raw words)."""
    n = 64
    colptr = np.arange(n + 1, dtype=np.uint32)
    x = np.arange(1, n + 1, dtype=np.uint64)
    y = np.zeros(n, dtype=np.uint64)
    oracle.spmv_csc("u64", n, n, colptr, np.arange(n, dtype=np.uint32), np.ones(n, np.uint64), x, y)
    assert int(y.sum()) == 2080 and (y == x).all()
    y = np.zeros(n, dtype=np.uint64)
    oracle.spmv_csc("u64", n, n, colptr, np.zeros(n, np.uint32), np.ones(n, np.uint64), x, y)
    assert int(y[0]) == 2080 and not y[1:].any()


def test_u64_wraps_mod_2_64(oracle):
    colptr = np.array([0, 2], dtype=np.uint32)
    rowind = np.array([0, 0], dtype=np.uint32)
    vals = np.array([2**63, 2**63 + 5], dtype=np.uint64)
    x = np.array([3], dtype=np.uint64)
    y = np.array([7], dtype=np.uint64)
    oracle.spmv_csc("u64", 1, 1, colptr, rowind, vals, x, y)
    assert int(y[0]) == (7 + 3 * 2**63 + 3 * (2**63 + 5)) % 2**64


@pytest.mark.parametrize("name", H.ALL_DIRS)
def test_transpose_roundtrip_and_csr_equivalence(oracle, name):
    """csr2csc(csc2csr(A)) == A, and the row-sequential CSR sum over the STABLE transpose is
    bit-identical to the CSC loop (the property strict mode relies on)."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    dtype = H.native_dtype(name)
    vals = H.derive_values(m, dtype, dtype)
    rp, ci, v = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    cp2, ri2, v2 = oracle.transpose(m.rows, m.cols, rp, ci, v)
    assert (cp2 == m.colptr).all() and (ri2 == m.rowind).all() and v2.tobytes() == vals.tobytes()
    x = H.make_x(m.cols, dtype, "rand", seed=99)
    y1 = np.zeros(m.rows, dtype=vals.dtype)
    y2 = np.zeros(m.rows, dtype=vals.dtype)
    oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y1)
    oracle.spmv_csr(dtype, m.rows, rp, ci, v, x, y2)
    assert y1.tobytes() == y2.tobytes()


def test_index_mask_strips_marker_bits(oracle):
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    vals = m.data_raw.view(np.float64)
    inds = m.rowind.copy()
    oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(inds), 0, 31)
    assert int((inds >> 31).sum()) == len(np.unique(m.rowind))
    y = np.zeros(m.rows)
    oracle.spmv_csc("f64", m.rows, m.cols, m.colptr, inds, vals, np.ones(m.cols), y, mask=0x3FFFFFFF)
    golden = np.fromfile(os.path.join(H.MATRICES, "circuit204", "golden.bin"), dtype="<f8")
    assert y.tobytes() == golden.tobytes()
    oracle.lib.oracle_clear_row_markings(H.C.c_uint32(m.nnz), H.ptr(inds), H.C.c_uint32(0x3FFFFFFF))
    assert (inds == m.rowind).all()


# ---- against the compiled reference itself (build container only) ----

@pytest.mark.parametrize("name", H.ALL_DIRS)
@pytest.mark.parametrize("dtype", ["f64", "u64"])
def test_oracle_vs_compiled_reference_random_values(oracle, reflib, name, dtype):
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.rand_values(m.nnz, dtype, seed=777)
    x = H.make_x(m.cols, dtype, "rand", seed=31337)
    y0 = H.rand_values(m.rows, dtype, seed=5)  # exercises the += (non-zero y)
    y_ref, y_or = y0.copy(), y0.copy()
    reflib.spmv(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y_ref)
    oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y_or)
    assert y_ref.tobytes() == y_or.tobytes()


def test_reference_default_vectors(reflib):
    """SoftwareSpMV allocates x = 1, y = 0 when given null (SoftwareSpMV.cpp:23-39)."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    y = reflib.spmv_default("f64", m.rows, m.cols, m.colptr, m.rowind, m.data_raw.view(np.float64))
    golden = np.fromfile(os.path.join(H.MATRICES, "circuit204", "golden.bin"), dtype="<f8")
    assert y.tobytes() == golden.tobytes()
    assert reflib.f64.ref_f64_sizeof_data() == 8 and reflib.u64.ref_u64_sizeof_data() == 8


@pytest.mark.parametrize("name", ["circuit204", "row64k", "i1k", "dia64-uint64"])
def test_row_marking_vs_reference(oracle, reflib, name):
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    for reverse, shift in ((0, 31), (1, 30)):
        a, b = m.rowind.copy(), m.rowind.copy()
        oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(a), reverse, shift)
        reflib.f64.ref_f64_mark_row_starts(H.C.c_uint(m.rows), H.C.c_uint(m.nnz), H.ptr(b), reverse, shift)
        assert (a == b).all()
    a, b = m.rowind.copy(), m.rowind.copy()
    ma = oracle.lib.oracle_max_alive(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(a))
    mb = reflib.f64.ref_f64_max_alive(H.C.c_uint(m.rows), H.C.c_uint(m.nnz), H.ptr(b))
    assert ma == mb and (a == b).all()
    if name != "dia64-uint64":  # empty column 0: the reference reads inds[-1] there
        sa = oracle.lib.oracle_max_col_span(H.C.c_uint32(m.cols), H.C.c_uint32(m.nnz), H.ptr(m.colptr), H.ptr(m.rowind))
        sb = reflib.f64.ref_f64_max_col_span(H.C.c_uint(m.cols), H.C.c_uint(m.nnz), H.ptr(m.colptr), H.ptr(m.rowind))
        assert sa == sb


def test_transpose_vs_reference_csr2csc(oracle, reflib):
    """software/csr2csc.c run on the CSR our transpose produced must give back the CSC files."""
    m = H.load_matrix_dir(os.path.join(H.MATRICES, "circuit204"))
    vals = m.data_raw.view(np.float64).copy()
    rp, ci, v = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    cs, ri, ca = reflib.csr2csc(m.rows, m.cols, v, ci.astype(np.int32), rp.astype(np.int32))
    assert (cs.astype(np.uint32) == m.colptr).all() and (ri.astype(np.uint32) == m.rowind).all()
    assert ca.tobytes() == vals.tobytes()
    # and the other direction: reference csr2csc used as CSC->CSR (roles swapped)
    rs, cidx, a = reflib.csr2csc(m.cols, m.rows, vals, m.rowind.astype(np.int32), m.colptr.astype(np.int32))
    assert (rs.astype(np.uint32) == rp).all() and (cidx.astype(np.uint32) == ci).all()
    assert a.tobytes() == v.tobytes()


def test_synth_host_generators(oracle):
    for kind, npr in ((0, 8), (2, 5)):
        rp, ci, v = oracle.synth_csr(kind, 0x5EED5EED, 300, 300, npr, "f64")
        assert rp[-1] == ci.size
        for r in range(300):
            seg = ci[rp[r]:rp[r + 1]]
            assert (np.diff(seg.astype(np.int64)) >= 0).all() and (seg < 300).all()
        if kind == 2:
            assert rp[1] - rp[0] == 6 + npr and rp[151] - rp[150] == 11 + npr
    r, c, v = oracle.synth_rmat_coo(7, 10, 4096, "f64")
    assert r.max() < 1024 and c.max() < 1024 and (np.abs(v) <= 1).all()
    # power-law skew: top-left quadrant holds ~57% of the edges
    frac = ((r < 512) & (c < 512)).mean()
    assert 0.5 < frac < 0.64
