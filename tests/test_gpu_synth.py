"""GPU tests on the synthetic matrices of BASELINE.json's configs: small sizes against the
oracle on the host mirror of the generator; full size through size-independent properties."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu
SEED = 0x5EED5EED
EPS = 2.0 ** -53


@pytest.fixture(scope="module")
def b2():
    import b200spmv
    return b200spmv


def DT(b2, dtype):
    return b2.F64 if dtype == "f64" else b2.U64


@pytest.mark.parametrize("kind,rows,npr", [(0, 5000, 32), (0, 777, 1), (2, 4099, 5), (1, 4096, 8), (1, 1 << 14, 32), (3, 3000, 16), (3, 129, 7)])
@pytest.mark.parametrize("dtype", ["f64", "u64"])
@pytest.mark.parametrize("slabs", [1, 3])
def test_synthetic_small_vs_oracle(b2, oracle, kind, rows, npr, dtype, slabs):
    e = b2.Engine(0)
    e.set_option("keep_csc", 1)
    e.set_option("slabs", slabs)
    e.setup_synthetic(kind, rows, rows, npr, SEED, DT(b2, dtype))
    # (1) the device generator + CSR->CSC transpose equal the host mirror, bit for bit
    cp, ri, v = e.get_csc()
    scale = int(np.log2(rows)) if kind == 1 else None
    cp_h, ri_h, v_h = oracle.synth_csc(kind, SEED, rows, rows, npr, dtype, scale=scale)
    assert (cp == cp_h).all() and (ri == ri_h).all() and v.tobytes() == v_h.tobytes()
    # (2) the CSR used by the kernels is the stable transpose of that CSC
    rp, ci, vv = e.get_csr()
    rp_o, ci_o, v_o = oracle.transpose(rows, rows, cp, ri, v)
    assert (rp == rp_o).all() and (ci == ci_o).all() and vv.tobytes() == v_o.tobytes()
    # (3) SpMV on the device-resident synthetic x against the reference loop
    e.fill_x_synthetic(SEED + 1)
    x = e.get_x()
    assert x.tobytes() == oracle.synth_x(SEED + 1, rows, dtype).tobytes()
    want = oracle.spmv_csc(dtype, rows, rows, cp, ri, v, x, np.zeros(rows, dtype=v.dtype))
    e.exec_device(1, b2.STRICT_ORDER)
    assert e.get_y().tobytes() == want.tobytes()
    e.exec_device(2, 0)  # y = A*x twice: idempotent without ACCUMULATE
    y = e.get_y()
    if dtype == "u64":
        assert y.tobytes() == want.tobytes()
    else:
        abs_sum = np.zeros(rows)
        row_nnz = np.zeros(rows, dtype=np.uint32)
        oracle.lib.oracle_row_abs_sums_f64(H.C.c_uint32(rows), H.C.c_uint32(rows), H.ptr(cp), H.ptr(ri), H.ptr(v),
                                           H.ptr(x), H.ptr(abs_sum), H.ptr(row_nnz))
        assert (np.abs(y - want) <= 2.0 * row_nnz * EPS * abs_sum).all()
    # (4) accumulate on the device: y <- y + A*x
    e.exec_device(1, b2.ACCUMULATE | b2.STRICT_ORDER)
    want2 = oracle.spmv_csc(dtype, rows, rows, cp, ri, v, x, y.copy())
    assert e.get_y().tobytes() == want2.tobytes()
    e.close()


@pytest.mark.parametrize("dtype,kernel", [("u64", 3), ("f64", 3), ("u64", 6)])
def test_headline_config_full_size_properties(b2, oracle, dtype, kernel):
    """configs[2]: uniform 16M x 16M, 32 nnz/row.  Too large for the host oracle, so:
    rows regenerated on the host from the counter-based definition (spot check, exact for
    u64 / strict fp64), merge-path == strict kernel (u64 bit-exact), and linearity."""
    rows, npr = 1 << 24, 32
    e = b2.Engine(0)
    e.set_option("kernel", kernel)
    e.setup_synthetic(0, rows, rows, npr, SEED, DT(b2, dtype))
    assert e.stat("nz") == rows * npr
    e.fill_x_synthetic(SEED + 1)
    e.exec_device(1, b2.STRICT_ORDER)
    y_strict = e.get_y()
    rng = np.random.default_rng(7)
    picks = np.concatenate([[0, 1, rows - 1], rng.integers(0, rows, 2000)])
    for r in picks:
        bits, mag = oracle.synth_row_dot(0, SEED, int(r), rows, rows, npr, dtype, SEED + 1)
        assert int(y_strict.view(np.uint64)[r]) == bits
    e.exec_device(1, 0)
    y_merge = e.get_y()
    if dtype == "u64":
        assert y_merge.tobytes() == y_strict.tobytes()
        # linearity mod 2^64: A*(2x) = 2*(A*x); accumulate: y + A*x = 2*A*x
        e.exec_device(1, b2.ACCUMULATE)
        assert e.get_y().tobytes() == (y_strict * np.uint64(2)).tobytes()
    else:
        for r in picks:
            bits, mag = oracle.synth_row_dot(0, SEED, int(r), rows, rows, npr, dtype, SEED + 1)
            ref = np.array([bits], dtype=np.uint64).view(np.float64)[0]
            assert abs(y_merge[r] - ref) <= 2 * npr * EPS * mag
        assert np.allclose(y_merge, y_strict, rtol=0, atol=2 * npr * EPS * 32)
    e.close()


@pytest.mark.parametrize("name", ["circuit204", "row64k", "i1k", "dia64-uint64", "rowvec64-uint64"])
def test_row_marking_on_device(b2, oracle, name):
    import os
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    e = b2.Engine(0)
    for reverse, shift in ((0, 31), (1, 30)):
        a, b = m.rowind.copy(), m.rowind.copy()
        oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(a), reverse, shift)
        e.mark_row_starts(m.rows, b, reverse, shift)
        assert (a == b).all()
    a, b = m.rowind.copy(), m.rowind.copy()
    want = oracle.lib.oracle_max_alive(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(a))
    assert e.max_alive(m.rows, b) == want and (a == b).all()
    want = oracle.lib.oracle_max_col_span(H.C.c_uint32(m.cols), H.C.c_uint32(m.nnz), H.ptr(m.colptr), H.ptr(m.rowind))
    assert e.max_col_span(m.cols, m.colptr, m.rowind) == want
    e.close()


@pytest.mark.parametrize("dtype", ["f64", "u64"])
def test_power_iteration_single_gpu(b2, oracle, dtype):
    n, npr, steps = 20000, 5, 6
    e = b2.Engine(0)
    e.set_option("keep_csc", 1)
    e.setup_synthetic(2, n, n, npr, SEED, DT(b2, dtype))
    cp, ri, v = e.get_csc()
    ms, nrm = e.power_iterate(steps, False, b2.STRICT_ORDER)
    x_gpu = e.get_x()
    if dtype == "u64":
        x = np.ones(n, dtype=np.uint64)
        y = np.zeros(n, dtype=np.uint64)
        oracle.lib.oracle_power_iterate_u64(H.C.c_uint32(n), H.ptr(cp), H.ptr(ri), H.ptr(v), H.ptr(x), H.ptr(y), steps)
        assert x_gpu.tobytes() == x.tobytes()
    else:
        x = np.full(n, 1.0 / np.sqrt(n))
        y = np.zeros(n)
        want = oracle.lib.oracle_power_iterate_f64(H.C.c_uint32(n), H.ptr(cp), H.ptr(ri), H.ptr(v), H.ptr(x), H.ptr(y), steps)
        # the norm is summed in a different order on the device: tolerance, not bits
        assert abs(nrm - want) <= 1e-12 * want
        assert np.allclose(x_gpu, x, rtol=1e-11, atol=1e-15)
    assert ms > 0
    e.close()


@pytest.mark.parametrize("dtype", ["u64", "f64"])
def test_exec_accumulates_onto_a_nonzero_host_y(b2, dtype):
    """b200spmv_exec with ACCUMULATE on vectors large enough (>= 1 MB) for the overlapped upload of y:
    y_out = y_in + A*x.  u64: bit-equal to the strict kernel; f64: y_in + t within one rounding of it."""
    rows = 1 << 18
    e = b2.Engine(0)
    e.setup_synthetic(0, rows, rows, 16, SEED, DT(b2, dtype))
    e.fill_x_synthetic(SEED + 1)
    x = e.get_x()
    npdt = np.uint64 if dtype == "u64" else np.float64
    y_in = H.make_x(rows, dtype, "rand", 77)
    y_strict = y_in.copy()
    e.exec(x, y_strict, b2.ACCUMULATE | b2.STRICT_ORDER)
    y = y_in.copy()
    e.exec(x, y, b2.ACCUMULATE)
    if dtype == "u64":
        assert y.tobytes() == y_strict.tobytes()
    else:
        t = np.zeros(rows, dtype=npdt)
        e.exec(x, t, 0)
        assert (y == y_in + t).all()
        assert np.allclose(y, y_strict, rtol=0, atol=2 * 17 * EPS * (np.abs(y_in).max() + 16.0))
    e.close()


@pytest.mark.parametrize("kind,rows,npr,slabs", [(1, 1 << 14, 32, 4), (1, 1 << 12, 8, 2), (2, 4099, 5, 3), (0, 5000, 3, 5)])
def test_slab_statistics_match_host_counts(b2, kind, rows, npr, slabs):
    """What the slab heuristics are fed with (non-empty rows of the block, longest row piece and non-empty rows of
    every slab) is counted on the device at setup: compare with numpy on the CSR the kernels use.  A slab with a
    long row piece must not go to the thread-per-row kernel whatever its average (dcsr_maxrow): same result."""
    e = b2.Engine(0)
    e.set_option("slabs", slabs)
    e.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)
    rp, ci, _ = e.get_csr()
    assert e.stat("nonempty_rows") == int((np.diff(rp.astype(np.int64)) > 0).sum())
    S = e.stat("slabs")
    assert S == slabs
    width = -(-rows // S)
    width = (width + 31) // 32 * 32
    rows_of = np.repeat(np.arange(rows), np.diff(rp.astype(np.int64)))
    for sl in range(S):
        inside = (ci >= sl * width) & (ci < (sl + 1) * width)
        per_row = np.bincount(rows_of[inside], minlength=rows)
        assert e.stat("slab_nnz_%d" % sl) == int(inside.sum())
        assert e.stat("slab_rows_%d" % sl) == int((per_row > 0).sum())
        assert e.stat("slab_maxrow_%d" % sl) == int(per_row.max())
    e.fill_x_synthetic(SEED + 1)
    e.exec_device(1, 0)
    want = e.get_y()
    for maxrow in (1, 1 << 30):  # never / whenever the average allows: the two kernels agree bit for bit on u64
        f = b2.Engine(0)
        f.set_option("slabs", slabs)
        f.set_option("dcsr_maxrow", maxrow)
        f.set_option("dcsr_x16", 1 << 20)
        f.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)
        f.fill_x_synthetic(SEED + 1)
        f.exec_device(1, 0)
        assert f.get_y().tobytes() == want.tobytes()
        f.close()
    e.close()


@pytest.mark.parametrize("kind,rows,npr,slabs", [(0, 1 << 18, 16, 1), (0, 1 << 18, 16, 3), (1, 1 << 18, 16, 2), (2, (1 << 18) + 4099, 5, 2),
                                                 (0, 1 << 17, 1, 1)])
@pytest.mark.parametrize("dtype", ["u64", "f64"])
@pytest.mark.parametrize("kernel", [3, 6])
def test_exec_download_pipeline(b2, kind, rows, npr, slabs, dtype, kernel):
    """b200spmv_exec on vectors >= 1 MB launches the last slab's kernel in pieces of tiles and downloads the rows
    each piece completes while the later pieces run (R-MAT: hub rows span several pieces; one entry per row: a
    piece boundary on every kind of row boundary).  u64: bit-equal to the strict kernel, with and without
    ACCUMULATE, several times in a row; f64: equal to the device-resident result of the same kernels (kernel 3;
    with kernel 6 selected the pieces of the last slab still run kernel 3 while the device-resident step runs
    kernel 6 throughout: two re-associations of the same sums, compared to 1e-9 absolute on O(1) values)."""
    e = b2.Engine(0)
    e.set_option("slabs", slabs)
    e.set_option("kernel", kernel)
    e.set_option("dcsr_x16", 0)  # the flagged kernels on every slab, whatever its density
    e.setup_synthetic(kind, rows, rows, npr, SEED, DT(b2, dtype))
    e.fill_x_synthetic(SEED + 1)
    x = e.get_x()
    npdt = np.uint64 if dtype == "u64" else np.float64
    want = np.zeros(rows, dtype=npdt)
    e.exec(x, want, b2.STRICT_ORDER)
    e.exec_device(1, 0)
    dev = e.get_y()
    for rep in range(3):
        y = np.full(rows, 7, dtype=npdt)  # overwritten, not accumulated
        e.exec(x, y, 0)
        if dtype == "u64" or kernel == 3:
            assert y.tobytes() == (want if dtype == "u64" else dev).tobytes()
        else:
            assert np.allclose(y, dev, rtol=0, atol=1e-9)
    y_in = H.make_x(rows, dtype, "rand", 91)
    y = y_in.copy()
    e.exec(x, y, b2.ACCUMULATE)
    if dtype == "u64":
        assert y.tobytes() == (y_in + want).tobytes()
    else:
        if kernel == 3:
            assert (y == y_in + dev).all()
        else:
            assert np.allclose(y, y_in + dev, rtol=0, atol=1e-9)
        assert np.allclose(dev, want, rtol=0, atol=1e-9)
    e.close()


@pytest.mark.parametrize("dtype", ["u64", "f64"])
@pytest.mark.parametrize("pin", [0, 1])
def test_exec_sliced_x_upload(b2, dtype, pin):
    """b200spmv_exec with x >= 8 MB and several slabs uploads x slab slice by slab slice on the copy stream and the
    kernels of slab s wait only for slice s (api.cu, ev_slab) -- the path bench.py's e2e number runs through
    (ADVICE round 1).  2^21 columns, 3 forced slabs, pageable and page-locked-in-place host vectors, flags 0 /
    ACCUMULATE / NO_PERSIST, a DIFFERENT x every repetition (a kernel that started before its slice had landed
    would read the previous x): u64 bit-equal to the strict device-resident result, f64 within the per-row bound
    2 n 2^-53 sum|a x| (+ the starting y), n = 4."""
    rows = 1 << 21
    e = b2.Engine(0)
    e.set_option("slabs", 3)
    e.set_option("pin_host", pin)
    e.setup_synthetic(0, rows, rows, 4, SEED, DT(b2, dtype))
    assert e.stat("slabs") == 3
    npdt = np.uint64 if dtype == "u64" else np.float64
    # the SAME host buffers for every repetition, refilled in place: option pin_host page-locks them where they lie
    # and requires them to outlive the context's use (SpMV objects keep their vectors for life, software/SpMV.h:31-33)
    x = np.zeros(rows, dtype=npdt)
    y = np.zeros(rows, dtype=npdt)
    for rep, flags in enumerate((0, b2.ACCUMULATE, b2.NO_PERSIST, b2.ACCUMULATE | b2.NO_PERSIST, 0, b2.ACCUMULATE)):
        x[:] = H.make_x(rows, dtype, "rand", 500 + rep)
        y0 = H.make_x(rows, dtype, "rand", 600 + rep)
        # oracle for this x: the strict kernel on device-resident vectors (itself pinned against the reference)
        e.set_x(x)
        e.set_y(y0 if flags & b2.ACCUMULATE else np.zeros(rows, dtype=npdt))
        e.exec_device(1, b2.STRICT_ORDER | (flags & b2.ACCUMULATE))
        want = e.get_y()
        e.set_x(np.zeros(rows, dtype=npdt))  # exec must bring its own x
        y[:] = y0
        e.exec(x, y, flags)
        if dtype == "u64":
            assert y.tobytes() == want.tobytes(), (rep, flags)
        else:
            # |a| < 1, |x| < 1, 4 entries per row: sum|a x| < 4
            bound = 2.0 * 5 * EPS * (4.0 + np.abs(y0))
            assert (np.abs(y - want) <= bound).all(), (rep, flags, float(np.abs(y - want).max()))
    e.close()


@pytest.mark.parametrize("kind,rows,npr", [(0, 5000, 32), (2, 4099, 5), (1, 1 << 14, 32), (3, 3000, 16)])
def test_generator_cross_check_small(b2, oracle, kind, rows, npr):
    """b200spmv_check_synthetic_u64 (A*x straight from the generator's definition, no CSR and no SpMV kernel) agrees
    with the oracle on the host-materialised matrix, and reports a deliberately corrupted y."""
    e = b2.Engine(0)
    e.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)
    e.fill_x_synthetic(SEED + 1)
    e.exec_device(1, 0)
    assert e.check_synthetic_u64(kind, npr, SEED) == 0
    scale = int(np.log2(rows)) if kind == 1 else None
    cp, ri, v = oracle.synth_csc(kind, SEED, rows, rows, npr, "u64", scale=scale)
    want = oracle.spmv_csc("u64", rows, rows, cp, ri, v, e.get_x(), np.zeros(rows, dtype=np.uint64))
    assert e.get_y().tobytes() == want.tobytes()
    bad = want.copy()
    bad[[3, rows // 2, rows - 1]] += np.uint64(1)
    e.set_y(bad)
    assert e.check_synthetic_u64(kind, npr, SEED) == 3
    e.close()


@pytest.mark.parametrize("config", ["C4 R-MAT scale 26", "C5 banded(11)+5, 128M rows"])
def test_largest_configs_full_size_parity(b2, oracle, config):
    """configs[3] and configs[4] at FULL size on one B200 (2^31 stored entries each: the u32 entry-offset edge,
    SURVEY 7.2-2), u64 so that every check is bit-exact (HardwareSpMV::compareGolden contract,
    software/HardwareSpMV.cpp:37-39):
      (i)   throughput kernels == strict row-sequential kernel on every row, y = A*x and y += A*x;
      (ii)  y == an independent device formulation straight off the generator (b200spmv_check_synthetic_u64: R-MAT
            one wrapping atomic add per edge; banded one regenerated row per thread);
      (iii) banded: 2000 rows regenerated on the HOST from the counter-based definition;
      (iv)  banded: 3 power-iteration steps (x <- A*x, wrapping) == three explicit SpMVs."""
    kind, rows, npr = (1, 1 << 26, 32) if config.startswith("C4") else (2, 1 << 27, 5)
    e = b2.Engine(0)
    e.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)
    assert e.stat("nz") == (rows * npr if kind == 1 else rows * 16 - 30)
    e.fill_x_synthetic(SEED + 1)
    e.fill_y_zero()
    e.exec_device(1, 0)
    assert e.check_synthetic_u64(kind, npr, SEED) == 0          # (ii)
    y = e.get_y()
    e.exec_device(1, b2.ACCUMULATE)
    y2 = e.get_y()
    e.exec_device(1, b2.STRICT_ORDER)
    assert e.check_synthetic_u64(kind, npr, SEED) == 0
    ys = e.get_y()
    assert y.tobytes() == ys.tobytes() and y2.tobytes() == (ys + ys).tobytes()  # (i)
    del y, y2
    if kind == 2:
        rng = np.random.default_rng(11)
        for r in np.concatenate([[0, 1, 4, 5, 6, rows - 6, rows - 5, rows - 1], rng.integers(0, rows, 2000)]):   # (iii)
            bits, _ = oracle.synth_row_dot(kind, SEED, int(r), rows, rows, npr, "u64", SEED + 1)
            assert int(ys[r]) == bits, r
        e.power_iterate(3, False, 0)                                  # (iv): x0 = 1, x <- A*x three times
        xp = e.get_x()
        e.set_x(np.ones(rows, dtype=np.uint64))
        for _ in range(3):
            e.exec_device(1, 0)
            e.set_x(e.get_y())
        assert xp.tobytes() == e.get_x().tobytes()
    e.close()


def test_band_matrix_is_not_cut_into_slabs(b2):
    """a matrix banded around the diagonal keeps one column slab however large x is (its gathers follow the row sweep);
    the same size with random entries is cut (api.cu choose_slabs, transpose.cu b2_count_band)"""
    rows = 1 << 24  # x = 134 MB > L2
    e = b2.Engine(0)
    e.setup_synthetic(b2.SYNTH_BAND, rows, rows, 16, SEED, b2.U64)
    assert e.stat("band_nz") == e.stat("nz") and e.stat("slabs") == 1
    e.close()
    e = b2.Engine(0)
    e.setup_synthetic(2, rows, rows, 5, SEED, b2.U64)  # banded(11) + 5 random per row
    nz, band = e.stat("nz"), e.stat("band_nz")
    assert 0.68 * nz < band < 0.80 * nz and e.stat("slabs") > 1
    e.close()
