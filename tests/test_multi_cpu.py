"""The N > 1 decomposition on CPU, world_size 2 over gloo (no GPU): nnz-balanced row blocks with a
replicated x need NO collective for one SpMV (each rank owns its slice of y), and the iterated SpMV
needs exactly one exchange per step -- the all-gather of the ranks' unequal row blocks into x (plus
one scalar all-reduce for the norm).  Each rank computes its block with the oracle; the assembled
result must equal the single-process oracle bit for bit (u64) / in the same order of operations (f64:
a row is never split, so every y_i is computed by exactly the reference's sequence of operations).
Also: `bench.py --impl reference` on a non-zero rank prints nothing and exits 0."""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, name, dtype, steps, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "spmv-vector-cache_b200"))
    import b200spmv as b2
    import helpers as HH
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    oracle = HH.OracleLib()
    m = HH.load_matrix_dir(os.path.join(HH.MATRICES, name))
    vals = HH.rand_values(m.nnz, dtype, 11)
    npdt = np.uint64 if dtype == "u64" else np.float64
    rowptr, colind, v = oracle.transpose(m.cols, m.rows, m.colptr, m.rowind, vals)
    splits = b2.row_splits(rowptr, world)
    rb, re = splits[rank], splits[rank + 1]
    # this rank's block as its own CSR with global column ids
    lptr = (rowptr[rb:re + 1] - rowptr[rb]).astype(np.uint32)
    lcol = colind[rowptr[rb]:rowptr[re]].copy()
    lval = v[rowptr[rb]:rowptr[re]].copy()
    x = HH.make_x(m.cols, dtype, "rand", 5) if dtype == "u64" else np.full(m.cols, 1.0 / np.sqrt(m.cols))
    for _ in range(steps):
        yb = np.zeros(re - rb, dtype=npdt)
        oracle.spmv_csr(dtype, re - rb, lptr, lcol, lval, x, yb)          # no collective: y slice is mine
        if dtype == "f64":
            s = torch.tensor([float(np.dot(yb, yb))], dtype=torch.float64)
            dist.all_reduce(s)                                             # one scalar
            yb = yb / np.sqrt(s.item())
        parts = [None] * world
        dist.all_gather_object(parts, yb)                                  # unequal blocks -> replicated x
        x = np.concatenate(parts).astype(npdt)
    if rank == 0:
        q.put((splits, x))
    dist.destroy_process_group()


@pytest.mark.parametrize("name,dtype", [("circuit204", "u64"), ("circuit204", "f64"), ("row64k", "u64")])
def test_row_block_decomposition_world2_gloo(oracle, name, dtype):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 300)
    steps = 3
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, dtype, steps, q)) for r in range(2)]
    for p in procs:
        p.start()
    splits, x_dist = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # single-process statement of the same iteration
    m = H.load_matrix_dir(os.path.join(H.MATRICES, name))
    vals = H.rand_values(m.nnz, dtype, 11)
    npdt = np.uint64 if dtype == "u64" else np.float64
    x = H.make_x(m.cols, dtype, "rand", 5) if dtype == "u64" else np.full(m.cols, 1.0 / np.sqrt(m.cols))
    for _ in range(steps):
        y = np.zeros(m.rows, dtype=npdt)
        oracle.spmv_csc(dtype, m.rows, m.cols, m.colptr, m.rowind, vals, x, y)
        x = y if dtype == "u64" else y / np.sqrt(float(np.dot(y, y)))
    assert splits[0] == 0 and splits[-1] == m.rows and splits == sorted(splits)
    if dtype == "u64":
        assert x_dist.tobytes() == x.tobytes()
    else:  # the norm is summed in two parts: tolerance on the scaling only
        assert np.allclose(x_dist, x, rtol=1e-13, atol=0)


def test_row_splits_are_nnz_balanced():
    import b200spmv as b2
    rng = np.random.default_rng(3)
    lens = rng.integers(0, 40, 5000)
    lens[17] = 20000  # a hub row bounds the balance
    rowptr = np.concatenate([[0], np.cumsum(lens)])
    for world in (1, 2, 4, 8):
        s = b2.row_splits(rowptr, world)
        assert len(s) == world + 1 and s[0] == 0 and s[-1] == 5000
        per = [int(rowptr[s[k + 1]] - rowptr[s[k]]) for k in range(world)]
        assert sum(per) == int(rowptr[-1])
        assert max(per) <= int(rowptr[-1]) / world + lens.max()


def test_reference_arm_nonzero_rank_is_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def _gather_worker(rank, world, port, cols, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "spmv-vector-cache_b200"))
    import b200spmv as b2
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    host_x = np.arange(cols, dtype=np.float64) * 0.5 + 1.0       # the same host vector on every rank
    dev_x = np.full(cols, np.nan)                                  # this rank's device copy: nothing uploaded yet
    blocks = b2.exec_x_blocks(cols, world)
    lo, hi = blocks[rank]
    dev_x[lo:hi] = host_x[lo:hi]                                   # the only host->device copy of x on this rank
    for r, (a, b) in enumerate(blocks):                            # grouped broadcasts = variable-size all-gather
        t = torch.from_numpy(dev_x[a:b].copy())
        dist.broadcast(t, r)
        dev_x[a:b] = t.numpy()
    q.put((rank, hi - lo, bool(np.array_equal(dev_x, host_x))))
    dist.destroy_process_group()


@pytest.mark.parametrize("cols", [1 << 16, (1 << 16) + 1234, 37])
def test_exec_x_upload_split_world2_gloo(cols):
    """b200spmv_exec over several ranks (csrc/api.cu, multi.cu b2_exec_gather_x): every rank uploads only its block of
    the host x and the blocks are exchanged -- here the same schedule over gloo with world_size 2.  The blocks tile
    [0, cols) exactly, their boundaries are whole 32-byte units, every rank ends with the full x, and no rank moved
    more than its share (+ 3 elements of rounding) over its host link."""
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "spmv-vector-cache_b200"))
    import b200spmv as b2
    for world in (2, 3, 8):
        bl = b2.exec_x_blocks(cols, world)
        assert bl[0][0] == 0 and bl[-1][1] == cols and all(bl[r][1] == bl[r + 1][0] for r in range(world - 1))
        assert all(lo % 4 == 0 for lo, _ in bl)
        assert max(hi - lo for lo, hi in bl) <= cols // world + 4 + (cols % world > 0) + 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + (os.getpid() % 90)
    ps = [ctx.Process(target=_gather_worker, args=(r, 2, port, cols, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(60)
    assert all(ok for _, _, ok in got)
    assert sum(n for _, n, _ in got) == cols
