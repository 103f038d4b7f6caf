"""Row-block sharding on ONE GPU: every rank of a split takes its place in turn (options "world" /
"rank", no communicator -- a single SpMV has no collective) and the assembled y must be
the full product; block boundaries must be the nnz-balanced split of b200spmv.row_splits."""
import os

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu
SEED = 0x5EED5EED


@pytest.mark.parametrize("name,dtype", [("circuit204", "u64"), ("circuit204", "f64"), ("row64k", "f64"), ("dia64-uint64", "u64")])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_loaded_matrix_sharded_equals_whole(oracle, name, dtype, world):
    import b200spmv as b2
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.rand_values(m.nnz, dtype, 21)
    x = H.make_x(m.cols, dtype, "rand", 22)
    npdt = np.uint64 if dtype == "u64" else np.float64
    want = np.zeros(m.rows, dtype=npdt)
    oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, want)
    rowptr, _, _ = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    splits = b2.row_splits(rowptr, world)
    y = np.zeros(m.rows, dtype=npdt)
    total_nz = 0
    for r in range(world):
        e = b2.Engine(0)
        e.set_option("world", world)
        e.set_option("rank", r)
        e.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, b2.U64 if dtype == "u64" else b2.F64)
        assert (e.stat("row_begin"), e.stat("row_end")) == (splits[r], splits[r + 1])
        assert e.stat("gpus") == world and e.stat("rank") == r
        total_nz += e.stat("local_nz")
        e.exec(x, y, b2.ACCUMULATE | b2.STRICT_ORDER)  # writes only [row_begin, row_end) of y
        e.close()
    assert total_nz == m.nnz
    assert y.tobytes() == want.tobytes()


@pytest.mark.parametrize("kind,rows,npr", [(0, 1 << 16, 32), (1, 1 << 14, 16), (2, 50000, 5)])
def test_synthetic_sharded_equals_whole(kind, rows, npr):
    import b200spmv as b2
    whole = b2.Engine(0)
    whole.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)
    whole.fill_x_synthetic(SEED + 1)
    whole.exec_device(1, 0)
    want = whole.get_y()
    nnz = whole.stat("nz")
    whole.close()
    world = 4
    y = np.zeros(rows, dtype=np.uint64)
    per = []
    for r in range(world):
        e = b2.Engine(0)
        e.set_option("world", world)
        e.set_option("rank", r)
        e.setup_synthetic(kind, rows, rows, npr, SEED, b2.U64)  # generates only this rank's rows
        e.fill_x_synthetic(SEED + 1)
        e.exec_device(1, 0)
        yr = e.get_y()
        rb, re = e.stat("row_begin"), e.stat("row_end")
        y[rb:re] = yr[rb:re]
        per.append(e.stat("local_nz"))
        assert e.stat("nz") == nnz
        with pytest.raises(b2.B200Error):
            e.power_iterate(1)  # several ranks, no communicator: refused, not silently wrong
        e.close()
    assert sum(per) == nnz
    assert y.tobytes() == want.tobytes()


@pytest.mark.parametrize("name,dtype,world,slabs", [("circuit204", "u64", 3, 0), ("circuit204", "f64", 2, 4), ("i64k", "u64", 8, 0),
                                                    ("i64k", "f64", 4, 6)])
@pytest.mark.parametrize("layout", [1, 2])
def test_slabs_aligned_with_the_ranks_blocks(oracle, name, dtype, world, slabs, layout):
    """option own_slabs: the column slabs follow the ranks' blocks of x (what lets the iterated SpMV overlap the
    exchange; layout 2: one slab each for everything left and right of the own block); a single SpMV over such
    slabs must still be the reference product, strict order included."""
    import b200spmv as b2
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.rand_values(m.nnz, dtype, 31)
    x = H.make_x(m.cols, dtype, "rand", 32)
    npdt = np.uint64 if dtype == "u64" else np.float64
    want = np.zeros(m.rows, dtype=npdt)
    oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, want)
    for flags in (b2.ACCUMULATE | b2.STRICT_ORDER, b2.ACCUMULATE):
        y = np.zeros(m.rows, dtype=npdt)
        for r in range(world):
            e = b2.Engine(0)
            e.set_option("world", world)
            e.set_option("rank", r)
            e.set_option("own_slabs", layout)
            if slabs:
                e.set_option("slabs", slabs)
            e.setup_csc(m.rows, m.cols, m.colptr, m.rowind, vals, b2.U64 if dtype == "u64" else b2.F64)
            if layout == 1:
                assert e.stat("slabs") % world == 0
            else:
                per = max(1, -(-slabs // world)) if slabs else None
                assert per is None or e.stat("slabs") == per + (r > 0) + (r < world - 1)
            e.exec(x, y, flags)
            e.close()
        if dtype == "u64" or flags & b2.STRICT_ORDER:
            assert y.tobytes() == want.tobytes()
        else:
            assert np.allclose(y, want, rtol=1e-12, atol=1e-12 * float(np.max(np.abs(want))))
