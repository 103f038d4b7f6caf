"""The reference's namesake as an experiment: the direct-mapped write-back vector cache of
chisel/cache-new/NoWMVectorCache.scala as a functional model over the CSC row-index stream, on the GPU
(csrc/vcache.cu, per-set parallel) against the sequential oracle restatement; counters named as
software/HardwareSpMVNewCache.cpp:189-204 reports them."""
import os

import numpy as np
import pytest

import helpers as H

KEYS = ("reads", "readMisses", "conflictEvictions", "coldMissSkips", "ddrReads", "flushWrites", "hits")


def oracle_model(oracle, inds, depth_log2, cms):
    out = np.zeros(7, dtype=np.uint64)
    oracle.lib.oracle_vector_cache_model(H.C.c_uint32(inds.size), H.ptr(inds), H.C.c_int(depth_log2), H.C.c_int(int(cms)), H.ptr(out))
    return dict(zip(KEYS, (int(v) for v in out)))


def test_known_answers(oracle):
    seq = np.array([0, 4, 0, 4], dtype=np.uint32)
    r = oracle_model(oracle, seq, 2, 0)  # 4 lines: rows 0 and 4 fight over line 0
    assert (r["readMisses"], r["conflictEvictions"], r["hits"], r["flushWrites"], r["ddrReads"]) == (4, 3, 0, 1, 4)
    r = oracle_model(oracle, seq, 3, 0)  # 8 lines: both fit
    assert (r["readMisses"], r["conflictEvictions"], r["hits"], r["flushWrites"]) == (2, 0, 2, 2)
    # cold-miss skip: the first touch of a row carries bit 31 and reads no DRAM; a re-fetch after eviction does
    marked = np.array([0 | 1 << 31, 4 | 1 << 31, 0, 4], dtype=np.uint32)
    r = oracle_model(oracle, marked, 2, 1)
    assert (r["readMisses"], r["coldMissSkips"], r["ddrReads"], r["conflictEvictions"]) == (4, 2, 2, 3)
    # identity matrix: every access is a first touch; a cache as deep as the vector never evicts
    ident = np.arange(64, dtype=np.uint32)
    r = oracle_model(oracle, ident, 6, 0)
    assert (r["readMisses"], r["conflictEvictions"], r["flushWrites"]) == (64, 0, 64)
    r = oracle_model(oracle, ident, 4, 0)
    assert (r["readMisses"], r["conflictEvictions"], r["flushWrites"]) == (64, 48, 16)


def test_every_row_is_written_back_exactly_once_per_residency(oracle):
    """DRAM writes = conflict evictions + flush = number of residencies = read misses (write-back cache)."""
    for name in ("circuit204", "row64k", "i1k"):
        m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
        for d in (0, 3, 8, 12):
            r = oracle_model(oracle, m.rowind, d, 0)
            assert r["conflictEvictions"] + r["flushWrites"] == r["readMisses"]
            assert r["hits"] + r["readMisses"] == m.nnz


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["circuit204", "row64k", "i64k", "dia64-uint64", "rowvec64-uint64"])
def test_gpu_model_equals_oracle(oracle, name):
    import b200spmv as b2
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    e = b2.Engine(0)
    marked = m.rowind.copy()
    oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(marked), 0, 31)
    for depth_log2 in (0, 1, 4, 8, 10, 16):
        assert e.vector_cache_model(m.rowind, depth_log2, False) == oracle_model(oracle, m.rowind, depth_log2, False)
        got = e.vector_cache_model(marked, depth_log2, True)
        assert got == oracle_model(oracle, marked, depth_log2, True)
        assert got["coldMissSkips"] == len(np.unique(m.rowind))  # one first touch per row that has entries
    e.close()


@pytest.mark.gpu
def test_gpu_model_on_a_synthetic_stream(oracle):
    import b200spmv as b2
    colptr, rowind, vals = oracle.synth_csc(0, 0x5EED5EED, 1 << 15, 1 << 15, 8, "u64")
    e = b2.Engine(0)
    for depth_log2 in (6, 10, 13, 15):
        assert e.vector_cache_model(rowind, depth_log2, False) == oracle_model(oracle, rowind, depth_log2, False)
    e.close()
