#!/usr/bin/env python
"""Iterated SpMV over several GPUs against the same iteration on one GPU (run under torchrun, >= 2 ranks).

Every rank builds its row block of a banded+random matrix (a) with the slabs aligned to the ranks' x blocks and
the other blocks pulled from the peers block by block under the slabs (exchange 2; exchange 3: the same with PACKED
blocks -- only the columns the pulling rank gathers), (b) the same with one NCCL
all-gather under the own-block slabs (exchange 1), (c) with the all-gather after the step, and (d) the whole matrix
alone.  After K steps the replicated x of (a)-(c) must equal the x of (d): bit for bit for uint64, to 1e-12 relative for fp64 (the slabs are summed in another order).
"""
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200spmv", os.path.join(ROOT, "spmv-vector-cache_b200", "b200spmv.py"))
b2 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b2)


def make(local, rank, world, dtype, rows, npr, slabs, own, exchange=2):
    e = b2.Engine(local)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(b2.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        e.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))
        if own:
            e.set_option("own_slabs", int(own))  # 1: every block its own slabs; 2: own block + one slab left, one right
            e.set_option("exchange", exchange)
    if slabs:
        e.set_option("slabs", slabs)
    e.setup_synthetic(2, rows, rows, npr, 12345, dtype)
    return e


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bad = 0
    quick = "quick" in sys.argv[1:]  # one unequal-block case and one large one (8-GPU runs are charged 8x)
    cases = (((1 << 20) + 8 * 977, 3), (1 << 22, 0)) if quick else ((1 << 20, 0), (1 << 20, 4), ((1 << 20) + 8 * 977, 3), (1 << 22, 0))
    for dtype, name in ((b2.F64, "f64"), (b2.U64, "u64")):
        for rows, slabs in cases:
            steps = 6 if dtype == b2.F64 else 3
            single = make(local, 0, 1, dtype, rows, 5, slabs, False)
            single.power_iterate(steps, False, 0)
            want = single.get_x()
            del single
            for own, exchange in ((2, 3), (1, 2)) if quick else ((2, 3), (1, 3), (2, 2), (1, 2), (2, 1), (1, 1), (0, 0)):
                e = make(local, rank, world, dtype, rows, 5, slabs, own, exchange)
                # two calls: the second continues from the replicated x (what bench.py does after its warm-up)
                e.power_iterate(steps // 2, False, 0)
                e.power_iterate(steps - steps // 2, True, 0)
                assert e.stat("exchange_mode") == exchange, e.stat("exchange_mode")
                got = e.get_x()
                if dtype == b2.U64:
                    ok = np.array_equal(got, want)
                    err = int((got != want).sum())
                else:
                    err = float(np.max(np.abs(got - want)) / np.max(np.abs(want)))
                    ok = err < 1e-12
                print("rank %d %s rows %d slabs opt %d -> %d own_slabs %d exchange %d: %s (%s)" %
                      (rank, name, rows, slabs, e.stat("slabs"), int(own), exchange, "ok" if ok else "MISMATCH", err), flush=True)
                bad += not ok
                del e
    # b200spmv_exec over several ranks: each rank uploads 1/world of the host x and the rest arrives over NVLink
    # (collective call); every rank's block of y must equal the single-GPU result bit for bit (u64), with the
    # gather on and off, overwrite and accumulate
    for rows in ((1 << 20) + 8 * 977, 1 << 21):
        single = make(local, 0, 1, b2.U64, rows, 5, 3, False)
        rng = np.random.default_rng(99)
        x = rng.integers(0, 1 << 63, rows, dtype=np.uint64)
        y0 = rng.integers(0, 1 << 63, rows, dtype=np.uint64)
        want = {fl: single.exec(x, y0.copy(), fl) for fl in (0, b2.ACCUMULATE)}
        del single
        for gather in (1, 0):
            e = make(local, rank, world, b2.U64, rows, 5, 3, False)
            e.set_option("exec_gather", gather)
            rb, re = e.stat("row_begin"), e.stat("row_end")
            for fl in (0, b2.ACCUMULATE, 0):
                e.set_x(np.zeros(rows, dtype=np.uint64))
                y = e.exec(x, y0.copy(), fl)
                ok = np.array_equal(y[rb:re], want[fl][rb:re])
                print("rank %d exec rows %d gather %d flags %d h2d %d: %s" % (rank, rows, gather, fl, e.stat("exec_h2d_bytes"), "ok" if ok else "MISMATCH"), flush=True)
                bad += not ok
            if gather:
                bad += not (e.stat("exec_h2d_bytes") <= 8 * (rows // world + 8) + 0)  # overwrite: 1/world of x only
            del e
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print("POWER PARITY %s" % ("OK" if int(t.item()) == 0 else "FAILED"), flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
