"""b200spmv_exec_noalloc: the SpMV entry point that allocates nothing (tile carries in a static device array inside the
library; no counterpart in the reference).  Through the C ABI, on device
vectors, against the existing device-resident entry point b200spmv_exec_device on the same context, two calls in a row.
Collected last on purpose (pytest -x: a failure here hides nothing)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x5EED5EED


@pytest.fixture(scope="module")
def b2():
    import b200spmv
    return b200spmv


@pytest.fixture(scope="module")
def torch():
    return pytest.importorskip("torch")


def to_dev(torch, a):
    """numpy u64 / f64 vector -> device tensor of the same bits (torch has no full uint64 support: int64 view)"""
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.int64).copy()).cuda()
    torch.cuda.synchronize()
    return t


def to_host(torch, t, npdt):
    torch.cuda.synchronize()
    return t.cpu().numpy().view(npdt)


# (kind, rows, nnz per row, slabs): uniform, R-MAT (empty rows, long rows), banded + random, pure band; 1 and 3 slabs
CASES = [(0, 1 << 18, 16, 3), (1, 1 << 16, 16, 2), (2, (1 << 17) + 4099, 5, 3), (3, 50000, 16, 1), (0, 3000, 1, 1)]


@pytest.mark.parametrize("kind,rows,npr,slabs", CASES)
@pytest.mark.parametrize("dtype", ["u64", "f64"])
def test_no_alloc_entry_point_equals_exec_device(b2, torch, kind, rows, npr, slabs, dtype):
    npdt = np.uint64 if dtype == "u64" else np.float64
    e = b2.Engine(0)
    e.set_option("slabs", slabs)
    e.set_option("kernel", 3)
    e.set_option("dcsr_x16", 0)  # the existing entry point runs kernel 3 on every slab too: same kernel, same tiles
    e.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64 if dtype == "u64" else b2.F64)
    e.fill_x_synthetic(SEED + 1)
    x = e.get_x()
    rng = np.random.default_rng(7)
    y0 = rng.integers(0, 1 << 62, rows, dtype=np.uint64) if dtype == "u64" else rng.standard_normal(rows)
    x_d = to_dev(torch, x)
    for flags in (0, b2.ACCUMULATE):
        # the existing entry point, two steps in a row from y0 (flags 0 overwrites: y0 must not show through)
        e.set_y(y0)
        e.exec_device(1, flags)
        want1 = e.get_y()
        e.exec_device(1, flags)
        want2 = e.get_y()
        # the new one, two calls in a row on the caller's device vectors (the scratch keeps the first call's carries)
        y_d = to_dev(torch, y0)
        e.exec_noalloc(x_d.data_ptr(), y_d.data_ptr(), flags)
        got1 = to_host(torch, y_d, npdt).copy()
        e.exec_noalloc(x_d.data_ptr(), y_d.data_ptr(), flags)
        got2 = to_host(torch, y_d, npdt).copy()
        assert got1.tobytes() == want1.tobytes()
        assert got2.tobytes() == want2.tobytes()
        if flags == 0:
            assert got2.tobytes() == got1.tobytes()
    # ... and inside the tolerance of the other tests against the context's shipped kernel choice (measured per slab)
    e.set_option("dcsr_x16", -1)
    e.set_y(y0)
    e.exec_device(1, b2.ACCUMULATE)
    ref = e.get_y()
    y_d = to_dev(torch, y0)
    e.exec_noalloc(x_d.data_ptr(), y_d.data_ptr(), b2.ACCUMULATE)
    got = to_host(torch, y_d, npdt)
    if dtype == "u64":
        assert got.tobytes() == ref.tobytes()
    else:
        assert np.allclose(got, ref, rtol=1e-12, atol=1e-12 * float(np.max(np.abs(ref))))
    e.close()


def test_no_alloc_on_the_callers_stream(b2, torch):
    """an explicit stream: the call only enqueues, the caller waits"""
    rows = 1 << 17
    e = b2.Engine(0)
    e.set_option("slabs", 2)
    e.set_option("dcsr_x16", 0)
    e.setup_synthetic(0, rows, rows, 8, SEED, b2.U64)
    e.fill_x_synthetic(SEED + 1)
    e.fill_y_zero()
    e.exec_device(1, 0)
    want = e.get_y()
    x_d = to_dev(torch, e.get_x())
    y_d = to_dev(torch, np.full(rows, 5, dtype=np.uint64))
    st = torch.cuda.Stream()
    for _ in range(2):
        e.exec_noalloc(x_d.data_ptr(), y_d.data_ptr(), 0, st.cuda_stream)
        st.synchronize()
        assert to_host(torch, y_d, np.uint64).tobytes() == want.tobytes()
    e.close()


def test_no_alloc_rejects_what_it_cannot_do(b2, torch):
    rows = 1 << 12
    e = b2.Engine(0)
    e.setup_synthetic(0, rows, rows, 4, SEED, b2.F64)
    x_d = to_dev(torch, np.ones(rows))
    y_d = to_dev(torch, np.zeros(rows))
    with pytest.raises(b2.B200Error, match="ACCUMULATE"):
        e.exec_noalloc(x_d.data_ptr(), y_d.data_ptr(), b2.STRICT_ORDER)
    with pytest.raises(b2.B200Error, match="null"):
        e.exec_noalloc(0, y_d.data_ptr(), 0)
    e.close()
    # beyond the static scratch: the 16M x 16M x 32 matrix in ONE column slab is 2^18 + 1 tiles
    big = b2.Engine(0)
    big.set_option("slabs", 1)
    big.setup_synthetic(0, 1 << 24, 1 << 24, 32, SEED, b2.F64)
    xb = torch.zeros(1 << 24, dtype=torch.float64, device="cuda")
    yb = torch.zeros(1 << 24, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    with pytest.raises(b2.B200Error, match="static scratch holds 262144"):
        big.exec_noalloc(xb.data_ptr(), yb.data_ptr(), 0)
    big.close()
    # the same matrix as shipped (3 slabs) fits
    ok = b2.Engine(0)
    ok.setup_synthetic(0, 1 << 24, 1 << 24, 32, SEED, b2.F64)
    ok.exec_noalloc(xb.data_ptr(), yb.data_ptr(), 0)
    assert float(yb.abs().max().item()) == 0.0  # x = 0
    ok.close()
