"""ctypes view of libb200spmv.so (include/b200spmv.h) for the tests and bench.py.

Plumbing only: every call goes straight to the C ABI.  There is no fallback -- if the
library or a GPU is missing this raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# B200SPMV_LIB: a measurement build of the same library (make ablate), experiments only
LIB_PATH = os.environ.get("B200SPMV_LIB") or os.path.join(HERE, "libb200spmv.so")

F64, U64 = 0, 1
ACCUMULATE, STRICT_ORDER, NO_PERSIST = 1, 2, 4
SYNTH_UNIFORM, SYNTH_RMAT, SYNTH_BANDED, SYNTH_BAND = 0, 1, 2, 3

EXPORTS = [
    "b200spmv_create", "b200spmv_destroy", "b200spmv_last_error", "b200spmv_set_option",
    "b200spmv_setup_csc", "b200spmv_setup_synthetic", "b200spmv_exec", "b200spmv_exec_device",
    "b200spmv_set_x", "b200spmv_set_y", "b200spmv_fill_x_synthetic", "b200spmv_fill_y_zero",
    "b200spmv_get_x", "b200spmv_get_y", "b200spmv_get_csr", "b200spmv_get_csc", "b200spmv_stat_u64",
    "b200spmv_comm_unique_id", "b200spmv_comm_init", "b200spmv_power_iterate",
    "b200spmv_check_synthetic_u64", "b200spmv_mark_row_starts", "b200spmv_max_alive", "b200spmv_max_col_span", "b200spmv_vector_cache_model",
    "b200spmv_exec_noalloc",
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(LIB_PATH + " missing: run `make -C spmv-vector-cache_b200` "
                               "(or __graft_entry__.build()); there is no fallback path")
        _lib = C.CDLL(LIB_PATH)
        _lib.b200spmv_last_error.restype = C.c_char_p
        _lib.b200spmv_last_error.argtypes = [C.c_void_p]
    return _lib


class B200Error(RuntimeError):
    pass


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def np_dtype(dtype):
    return np.float64 if dtype == F64 else np.uint64


class Engine:
    """One b200spmv_ctx (one GPU)."""

    def __init__(self, device=0):
        self.L = lib()
        self.ctx = C.c_void_p()
        rc = self.L.b200spmv_create(C.byref(self.ctx), C.c_int(device))
        if rc != 0:
            raise B200Error("b200spmv_create: %d %s" % (rc, self.L.b200spmv_last_error(None).decode()))
        self.dtype = F64

    def close(self):
        if self.ctx:
            self.L.b200spmv_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise B200Error("%s: %d %s" % (what, rc, self.L.b200spmv_last_error(self.ctx).decode()))

    def rc_of(self, fn, *args):
        return fn(self.ctx, *args)

    def set_option(self, key, value):
        self._ck(self.L.b200spmv_set_option(self.ctx, key.encode(), C.c_int64(value)), "set_option")

    def stat(self, key):
        out = C.c_uint64()
        self._ck(self.L.b200spmv_stat_u64(self.ctx, key.encode(), C.byref(out)), "stat " + key)
        return out.value

    def setup_csc(self, rows, cols, colptr, rowind, vals, dtype, index_mask=0xFFFFFFFF):
        self.dtype = dtype
        self._keep = (np.ascontiguousarray(colptr, np.uint32), np.ascontiguousarray(rowind, np.uint32),
                      np.ascontiguousarray(vals))
        self._ck(self.L.b200spmv_setup_csc(self.ctx, C.c_uint32(rows), C.c_uint32(cols),
                                           C.c_uint32(int(self._keep[1].size)), _p(self._keep[0]),
                                           _p(self._keep[1]), _p(self._keep[2]), C.c_int(dtype),
                                           C.c_uint32(index_mask)), "setup_csc")

    def setup_synthetic(self, kind, rows, cols, npr, seed, dtype):
        self.dtype = dtype
        self._ck(self.L.b200spmv_setup_synthetic(self.ctx, C.c_int(kind), C.c_uint32(rows), C.c_uint32(cols),
                                                 C.c_uint32(npr), C.c_uint64(seed), C.c_int(dtype)),
                 "setup_synthetic")

    def check_synthetic_u64(self, kind, npr, seed):
        """rows of this rank's block where the y in HBM differs from A*x recomputed from the generator (u64)"""
        out = C.c_uint64()
        self._ck(self.L.b200spmv_check_synthetic_u64(self.ctx, C.c_int(kind), C.c_uint32(npr), C.c_uint64(seed), C.byref(out)),
                 "check_synthetic_u64")
        return out.value

    def exec(self, x, y, flags=ACCUMULATE):
        """y (+)= A*x on host vectors (numpy arrays or raw addresses), in place."""
        xp = _p(x) if isinstance(x, np.ndarray) else C.c_void_p(x)
        yp = _p(y) if isinstance(y, np.ndarray) else C.c_void_p(y)
        self._ck(self.L.b200spmv_exec(self.ctx, xp, yp, C.c_int(flags)), "exec")
        return y

    def exec_device(self, iters, flags=0):
        ms = np.zeros(iters, dtype=np.float32)
        self._ck(self.L.b200spmv_exec_device(self.ctx, C.c_int(iters), C.c_int(flags), _p(ms)), "exec_device")
        return ms

    def exec_noalloc(self, x_dev_ptr, y_dev_ptr, flags=0, stream=None):
        """y (+)= A*x on device vectors given by address (ints), without any allocation inside the library"""
        self._ck(self.L.b200spmv_exec_noalloc(self.ctx, C.c_void_p(x_dev_ptr), C.c_void_p(y_dev_ptr), C.c_int(flags),
                                              C.c_void_p(stream)), "exec_noalloc")

    def set_x(self, x):
        self._ck(self.L.b200spmv_set_x(self.ctx, _p(np.ascontiguousarray(x))), "set_x")

    def set_y(self, y):
        self._ck(self.L.b200spmv_set_y(self.ctx, _p(np.ascontiguousarray(y))), "set_y")

    def fill_x_synthetic(self, seed):
        self._ck(self.L.b200spmv_fill_x_synthetic(self.ctx, C.c_uint64(seed)), "fill_x_synthetic")

    def fill_y_zero(self):
        self._ck(self.L.b200spmv_fill_y_zero(self.ctx), "fill_y_zero")

    def get_x(self):
        x = np.zeros(self.stat("cols"), dtype=np_dtype(self.dtype))
        self._ck(self.L.b200spmv_get_x(self.ctx, _p(x)), "get_x")
        return x

    def get_y(self):
        y = np.zeros(self.stat("rows"), dtype=np_dtype(self.dtype))
        self._ck(self.L.b200spmv_get_y(self.ctx, _p(y)), "get_y")
        return y

    def get_csr(self):
        rows, nnz = self.stat("local_rows"), self.stat("local_nz")
        rp = np.zeros(rows + 1, np.uint32)
        ci = np.zeros(nnz, np.uint32)
        v = np.zeros(nnz, np_dtype(self.dtype))
        self._ck(self.L.b200spmv_get_csr(self.ctx, _p(rp), _p(ci), _p(v)), "get_csr")
        return rp, ci, v

    def get_csc(self):
        cols, nnz = self.stat("cols"), self.stat("local_nz")
        cp = np.zeros(cols + 1, np.uint32)
        ri = np.zeros(nnz, np.uint32)
        v = np.zeros(nnz, np_dtype(self.dtype))
        self._ck(self.L.b200spmv_get_csc(self.ctx, _p(cp), _p(ri), _p(v)), "get_csc")
        return cp, ri, v

    def comm_init(self, rank, world, uid_bytes):
        buf = C.create_string_buffer(bytes(uid_bytes), 128)
        self._ck(self.L.b200spmv_comm_init(self.ctx, C.c_int(rank), C.c_int(world), buf), "comm_init")

    def power_iterate(self, steps, use_current_x=False, flags=0):
        ms = C.c_float()
        nrm = C.c_double()
        self._ck(self.L.b200spmv_power_iterate(self.ctx, C.c_int(steps), C.c_int(1 if use_current_x else 0),
                                               C.c_int(flags), C.byref(ms), C.byref(nrm)), "power_iterate")
        return ms.value, nrm.value

    def mark_row_starts(self, rows, inds, reverse, shift):
        self._ck(self.L.b200spmv_mark_row_starts(self.ctx, C.c_uint32(rows), C.c_uint32(int(inds.size)), _p(inds),
                                                 C.c_int(reverse), C.c_int(shift)), "mark_row_starts")
        return inds

    def max_alive(self, rows, inds):
        out = C.c_uint32()
        self._ck(self.L.b200spmv_max_alive(self.ctx, C.c_uint32(rows), C.c_uint32(int(inds.size)), _p(inds),
                                           C.byref(out)), "max_alive")
        return out.value

    def vector_cache_model(self, inds, depth_log2, cms):
        """-> dict of the NewCache counters for the row-index stream `inds` (uint32, storage order)."""
        out = np.zeros(7, dtype=np.uint64)
        self._ck(self.L.b200spmv_vector_cache_model(self.ctx, C.c_uint32(inds.size), _p(inds), C.c_int(depth_log2),
                                                    C.c_int(int(cms)), _p(out)), "vector_cache_model")
        keys = ("reads", "readMisses", "conflictEvictions", "coldMissSkips", "ddrReads", "flushWrites", "hits")
        return dict(zip(keys, (int(v) for v in out)))

    def max_col_span(self, cols, colptr, inds):
        out = C.c_uint32()
        self._ck(self.L.b200spmv_max_col_span(self.ctx, C.c_uint32(cols), C.c_uint32(int(inds.size)), _p(colptr),
                                              _p(inds), C.byref(out)), "max_col_span")
        return out.value


def row_splits(rowptr, world):
    """Host statement of the nnz-balanced row-block split (csrc/multi.cu splits_kernel):
    split k = first row whose prefix of stored entries reaches k*nnz/world.  rowptr: CSR row pointer."""
    rowptr = np.asarray(rowptr, dtype=np.uint64)
    rows, nnz = rowptr.size - 1, int(rowptr[-1])
    out = [0]
    for k in range(1, world):
        out.append(int(np.searchsorted(rowptr[:rows], np.uint64(nnz * k // world), side="left")))
    out.append(rows)
    if world > 1 and rows % world == 0:  # equal row counts when they are balanced to 2 % anyway (common.cuh b2_row_splits)
        per = rows // world
        worst = max(int(rowptr[(r + 1) * per] - rowptr[r * per]) for r in range(world))
        if worst * 100 * world <= 102 * nnz + 100 * world:
            return [r * per for r in range(world + 1)]
    return out


def exec_x_blocks(cols, world):
    """Host statement of csrc/multi.cu b2_exec_x_block: the columns of x rank r uploads in b200spmv_exec when a
    communicator is attached ([lo_r, hi_r), boundaries cut to whole 32-byte units); the rest arrives over NVLink."""
    lo = [(cols * r // world) & ~3 for r in range(world)]
    return [(lo[r], cols if r + 1 == world else lo[r + 1]) for r in range(world)]


def comm_unique_id():
    buf = C.create_string_buffer(128)
    rc = lib().b200spmv_comm_unique_id(buf)
    if rc != 0:
        raise B200Error("b200spmv_comm_unique_id failed: %d" % rc)
    return buf.raw
