"""Shared test plumbing: matrix-directory reader, ctypes views of the checker libraries.

The oracle (oracle/liboracle.so) and the compiled reference (oracle/_ref/*.so) are TEST
infrastructure; the product is reached only through the C ABI in include/b200spmv.h.
"""
import ctypes as C
import os
from collections import namedtuple

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
MATRICES = os.path.join(GOLDEN, "matrices")

FP64_DIRS = ["circuit204", "i64", "i1k", "i64k", "row64k"]
U64_DIRS = ["circuit204-uint64", "dia64-uint64", "i64-uint64", "i1024-uint64", "rowvec64-uint64"]
ALL_DIRS = FP64_DIRS + U64_DIRS

Matrix = namedtuple("Matrix", "name rows cols nnz colptr rowind data_raw")

u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
f64p = C.POINTER(C.c_double)


def ptr(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


def load_matrix_dir(path):
    """<name>/<name>-{meta,indptr,inds,data}.bin (software/main.cpp:26-37 naming;
    layout matrices/matrixutils.py:187-260).  The three *Base address fields of the
    metadata are Zynq DRAM addresses and are ignored on a 64-bit host."""
    name = os.path.basename(os.path.normpath(path))
    meta = np.fromfile(os.path.join(path, name + "-meta.bin"), dtype="<u4")
    assert meta.size == 7, "metadata is 7 x u32 (SparseMatrix.h:8-16)"
    rows, cols, nnz = int(meta[0]), int(meta[1]), int(meta[2])
    colptr = np.fromfile(os.path.join(path, name + "-indptr.bin"), dtype="<u4")
    rowind = np.fromfile(os.path.join(path, name + "-inds.bin"), dtype="<u4")
    data = np.fromfile(os.path.join(path, name + "-data.bin"), dtype="<u8")
    assert colptr.size == cols + 1 and rowind.size == nnz and data.size == nnz
    return Matrix(name, rows, cols, nnz, colptr, rowind, data)


def native_dtype(name):
    return "u64" if name.endswith("uint64") else "f64"


def derive_values(m, native, dtype):
    """Values of matrix m in `dtype`.  Native dtype: the file's words.  fp64->u64: every
    value becomes uint64(1) (matrixutils.py:100-103); u64->fp64: numeric conversion.
    Bit patterns are never reinterpreted across dtypes."""
    if native == dtype:
        return m.data_raw.view(np.float64).copy() if dtype == "f64" else m.data_raw.copy()
    if dtype == "u64":
        return np.ones(m.nnz, dtype=np.uint64)
    return m.data_raw.astype(np.float64)


def splitmix64(seed, n):
    z = (np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)) + np.uint64(seed)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def make_x(n, dtype, kind, seed=1234):
    if kind == "ones":
        return np.ones(n, dtype=np.float64 if dtype == "f64" else np.uint64)
    bits = splitmix64(seed, n)
    if dtype == "u64":
        return bits  # full-range words: products and sums wrap mod 2^64
    return (bits >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0) * 2.0 - 1.0


def rand_values(n, dtype, seed):
    bits = splitmix64(seed, n)
    if dtype == "u64":
        return bits
    return (bits >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0) * 2.0 - 1.0


def hash_words(y):
    """h = h*1099511628211 XOR w over the 8-byte words."""
    h = 0
    for w in np.ascontiguousarray(y).view(np.uint64).tolist():
        h = ((h * 1099511628211) & 0xFFFFFFFFFFFFFFFF) ^ w
    return h


class OracleLib:
    """oracle/liboracle.so -- our C restatement (spmv_oracle.c, synth_host.c)."""

    def __init__(self):
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/liboracle.so missing: run `make -C oracle oracle`")
        self.lib = L = C.CDLL(path)
        L.oracle_hash_words.restype = C.c_uint64
        L.oracle_max_alive.restype = C.c_uint32
        L.oracle_max_col_span.restype = C.c_uint32
        L.oracle_time_spmv_csc_f64.restype = C.c_double
        L.oracle_power_iterate_f64.restype = C.c_double
        L.synth_host_nnz.restype = C.c_uint64
        L.synth_host_row_dot.restype = C.c_uint64

    def spmv_csc(self, dtype, rows, cols, colptr, rowind, vals, x, y, mask=0xFFFFFFFF):
        fn = self.lib.oracle_spmv_csc_f64 if dtype == "f64" else self.lib.oracle_spmv_csc_u64
        fn(C.c_uint32(rows), C.c_uint32(cols), ptr(colptr), ptr(rowind), ptr(vals), ptr(x), ptr(y),
           C.c_uint32(mask))
        return y

    def spmv_csr(self, dtype, rows, rowptr, colind, vals, x, y):
        fn = self.lib.oracle_spmv_csr_f64 if dtype == "f64" else self.lib.oracle_spmv_csr_u64
        fn(C.c_uint32(rows), ptr(rowptr), ptr(colind), ptr(vals), ptr(x), ptr(y))
        return y

    def transpose(self, src_major, dst_major, src_ptr, src_ind, src_val, mask=0xFFFFFFFF):
        nnz = int(src_ind.size)
        dptr = np.zeros(dst_major + 1, dtype=np.uint32)
        dind = np.zeros(nnz, dtype=np.uint32)
        dval = np.zeros(nnz, dtype=np.uint64)
        sv = np.ascontiguousarray(src_val).view(np.uint64)
        self.lib.oracle_transpose(C.c_uint32(src_major), C.c_uint32(dst_major), C.c_uint32(nnz),
                                  ptr(src_ptr), ptr(src_ind), ptr(sv), ptr(dptr), ptr(dind),
                                  ptr(dval), C.c_uint32(mask))
        return dptr, dind, dval.view(src_val.dtype)

    def synth_csr(self, kind, seed, rows, cols, npr, dtype):
        u = 1 if dtype == "u64" else 0
        nnz = int(self.lib.synth_host_nnz(C.c_int(kind), C.c_uint32(rows), C.c_uint32(npr)))
        rowptr = np.zeros(rows + 1, dtype=np.uint32)
        colind = np.zeros(nnz, dtype=np.uint32)
        vals = np.zeros(nnz, dtype=np.uint64)
        self.lib.synth_host_csr(C.c_int(kind), C.c_uint64(seed), C.c_uint32(rows), C.c_uint32(cols),
                                C.c_uint32(npr), C.c_int(u), ptr(rowptr), ptr(colind), ptr(vals))
        return rowptr, colind, vals.view(np.float64 if dtype == "f64" else np.uint64)

    def synth_rmat_coo(self, seed, scale, nedges, dtype):
        u = 1 if dtype == "u64" else 0
        row = np.zeros(nedges, dtype=np.uint32)
        col = np.zeros(nedges, dtype=np.uint32)
        vals = np.zeros(nedges, dtype=np.uint64)
        self.lib.synth_host_rmat_coo(C.c_uint64(seed), C.c_int(scale), C.c_uint64(nedges), C.c_int(u),
                                     ptr(row), ptr(col), ptr(vals))
        return row, col, vals.view(np.float64 if dtype == "f64" else np.uint64)

    def coo_to_compressed(self, nkeys, key, other, vals):
        nnz = int(key.size)
        p = np.zeros(nkeys + 1, dtype=np.uint32)
        ind = np.zeros(nnz, dtype=np.uint32)
        out = np.zeros(nnz, dtype=np.uint64)
        self.lib.synth_host_coo_to_compressed(C.c_uint32(nkeys), C.c_uint64(nnz), ptr(key), ptr(other),
                                              ptr(np.ascontiguousarray(vals).view(np.uint64)), ptr(p),
                                              ptr(ind), ptr(out))
        return p, ind, out.view(vals.dtype)

    def synth_x(self, seed, n, dtype):
        x = np.zeros(n, dtype=np.uint64)
        self.lib.synth_host_x(C.c_uint64(seed), C.c_uint64(n), C.c_int(1 if dtype == "u64" else 0), ptr(x))
        return x.view(np.float64 if dtype == "f64" else np.uint64)

    def synth_row_dot(self, kind, seed, r, rows, cols, npr, dtype, xseed):
        mag = C.c_double(0.0)
        bits = self.lib.synth_host_row_dot(C.c_int(kind), C.c_uint64(seed), C.c_uint32(r), C.c_uint32(rows),
                                           C.c_uint32(cols), C.c_uint32(npr), C.c_int(1 if dtype == "u64" else 0),
                                           C.c_uint64(xseed), C.byref(mag))
        return int(bits), mag.value

    def synth_csc(self, kind, seed, rows, cols, npr, dtype, scale=None):
        """Host CSC (the reference's layout) of a synthetic matrix, via CSR/COO -> counting sort."""
        if kind == 1:
            r, c, v = self.synth_rmat_coo(seed, scale, rows * npr, dtype)
            return self.coo_to_compressed(cols, c, r, v)
        rowptr, colind, vals = self.synth_csr(kind, seed, rows, cols, npr, dtype)
        return self.transpose(rows, cols, rowptr, colind, vals)


class RefLib:
    """oracle/_ref/libref_{f64,u64}.so -- the UNMODIFIED reference sources, compiled."""

    def __init__(self):
        d = os.path.join(ROOT, "oracle", "_ref")
        self.f64 = C.CDLL(os.path.join(d, "libref_f64.so"))
        self.u64 = C.CDLL(os.path.join(d, "libref_u64.so"))
        self.csr2csc_lib = C.CDLL(os.path.join(d, "libref_csr2csc.so"))
        for L, p in ((self.f64, "ref_f64_"), (self.u64, "ref_u64_")):
            getattr(L, p + "spmv").restype = C.c_double
            getattr(L, p + "max_alive").restype = C.c_uint32
            getattr(L, p + "max_col_span").restype = C.c_uint32
            getattr(L, p + "sizeof_data").restype = C.c_uint32

    @staticmethod
    def available():
        d = os.path.join(ROOT, "oracle", "_ref")
        return all(os.path.exists(os.path.join(d, f))
                   for f in ("libref_f64.so", "libref_u64.so", "libref_csr2csc.so"))

    def _fn(self, dtype, name):
        return getattr(self.f64 if dtype == "f64" else self.u64, "ref_%s_%s" % (dtype, name))

    def spmv(self, dtype, rows, cols, colptr, rowind, vals, x, y):
        return self._fn(dtype, "spmv")(C.c_uint(rows), C.c_uint(cols), C.c_uint(int(rowind.size)),
                                       ptr(colptr), ptr(rowind), ptr(vals), ptr(x), ptr(y))

    def spmv_default(self, dtype, rows, cols, colptr, rowind, vals):
        y = np.zeros(rows, dtype=np.float64 if dtype == "f64" else np.uint64)
        self._fn(dtype, "spmv_default")(C.c_uint(rows), C.c_uint(cols), C.c_uint(int(rowind.size)),
                                        ptr(colptr), ptr(rowind), ptr(vals), ptr(y))
        return y

    def csr2csc(self, n, m, a, col_idx, row_start):
        nz = int(col_idx.size)
        csc_a = np.zeros(nz, dtype=np.float64)
        row_idx = np.zeros(nz, dtype=np.int32)
        col_start = np.zeros(m + 1, dtype=np.int32)
        self.csr2csc_lib.csr2csc(C.c_int(n), C.c_int(m), C.c_int(nz), ptr(a), ptr(col_idx),
                                 ptr(row_start), ptr(csc_a), ptr(row_idx), ptr(col_start))
        return col_start, row_idx, csc_a
