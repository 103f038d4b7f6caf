#!/usr/bin/env python
"""bench.py -- SpMV GNNZ/s (+ algorithmic HBM GB/s vs roofline) on 1/2/4/8 B200.

A "step" is one pass of the hot path (y += A*x, software/SoftwareSpMV.cpp:59-64) over the
whole matrix.  Default workload = BASELINE.json configs[2]: synthetic uniform-random
16M x 16M, 32 nnz/row, fp64.  N > 1: the same matrix split into nnz-balanced row blocks, one
rank per GPU, no data-path collective ("strong" scaling; value = total nnz / max-over-ranks time).

  python bench.py [--gpus N --steps K --warmup W] [--workload uniform|rmat|banded-power]
  python bench.py --impl reference ...   # the reference's own CPU loop on a bounded sample

Python/torch here is plumbing (rendezvous, pinned buffers); the timed work is libb200spmv.so.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "spmv-vector-cache_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

SEED = 0x5EED5EED
METRIC = "SpMV GNNZ/s"
WORKLOADS = {
    # name: (kind, rows, nnz_per_row, description)
    "uniform": (0, 1 << 24, 32, "synthetic uniform-random 16M x 16M, 32 nnz/row, fp64"),
    "rmat": (1, 1 << 26, 32, "synthetic R-MAT scale 26 (64M rows, 2^31 nnz), fp64"),
    "banded-power": (2, 1 << 27, 5, "power iteration, 128M-row banded(11)+5 random, fp64"),
    # locality probes: here HBM, not the gather path, should bind
    "band": (3, 1 << 24, 16, "synthetic pure band 16M x 16M, 16 consecutive columns per row, fp64"),
    "banded": (2, 1 << 24, 5, "synthetic banded(11)+5 random 16M x 16M, fp64"),
}


def b_alg(nnz, rows, cols):
    """values + u32 indices + u32 pointers + x read once + y written once."""
    return 12 * nnz + 4 * (rows + 1) + 8 * cols + 8 * rows


KERNEL_NAMES = {0: "spmv_rowseq_kernel", 3: "spmv_flag_kernel", 6: "spmv_stream_kernel"}


def recorded_traffic(workload, rows, cols, slabs, mode, world, flags):
    """DRAM bytes per step of the dominant kernel (ncu dram__bytes_read.sum + dram__bytes_write.sum, summed over
    the kernel's launches of one step) from the committed capture of exactly this configuration, else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if world != 1 or not os.path.exists(p):
        return None, None
    with open(p) as f:
        for rec in json.load(f):
            if (rec["workload"], rec["rows"], rec["cols"], rec["slabs"], rec["mode"], rec["flags"]) == \
                    (workload, rows, cols, slabs, mode, flags):
                return rec["dram_bytes_per_step"], rec["source"]
    return None, None


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi-equivalent clock / throttle-reason samples via NVML during the timed region."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def mem_available_gb():
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    return int(ln.split()[1]) / 1048576.0
    except OSError:
        pass
    return 0.0


def cpu_reference_sample(steps, warmup, budget_s=150.0):
    """The reference's CPU loop (software/SoftwareSpMV.cpp:50-70) on the headline workload itself: the
    16M x 16M x 32 uniform matrix (same seed, the host mirror of the device generator, oracle/synth_host.c)
    materialised on the host in the reference's CSC layout -- 6.8 GB, needs ~16 GB while it is built.  The
    sample is bounded in STEPS (one SpMV is ~4 s on one core), not in size.  Only when the host has less than
    24 GB available does it fall back to the same shape at 2 nnz/row (same x / y footprints, 1/16 of the
    entries) and say so.  Uses oracle/_ref (the unmodified reference, kind "reference") when present, else
    the C port."""
    import numpy as np
    import helpers as H
    oracle = H.OracleLib()
    rows = cols = 1 << 24
    full = mem_available_gb() >= 24.0
    npr = 32 if full else 2
    t_gen = time.perf_counter()
    colptr, rowind, vals = oracle.synth_csc(0, SEED, rows, cols, npr, "f64")
    x = oracle.synth_x(SEED + 1, cols, "f64")
    t_gen = time.perf_counter() - t_gen
    y = np.zeros(rows)
    nnz = int(rowind.size)
    use_ref = H.RefLib.available()
    ref = H.RefLib() if use_ref else None
    times = []
    t_begin = time.perf_counter()
    for i in range(warmup + steps):
        if i > warmup and time.perf_counter() - t_begin > budget_s:
            break  # bounded: keep the whole arm within a few minutes whatever K is
        y[:] = 0.0
        if use_ref:
            t = ref.spmv("f64", rows, cols, colptr, rowind, vals, x, y)  # loop-only bracket, SoftwareSpMV.cpp:57,66
        else:
            t0 = time.perf_counter()
            oracle.spmv_csc("f64", rows, cols, colptr, rowind, vals, x, y)
            t = time.perf_counter() - t0
        if i >= warmup:
            times.append(t)
    tot = sum(times)
    what = ("the workload itself: uniform 16M x 16M, 32 nnz/row (%d nnz), host-materialised" % nnz) if full else \
           ("uniform 16M x 16M at %d nnz/row (%d nnz = 1/16 of the workload, same x/y footprint; host has < 24 GB free)" % (npr, nnz))
    return {
        "value": nnz * len(times) / tot / 1e9, "unit": "GNNZ/s", "cores": 1,
        "kind": "reference" if use_ref else "port",
        "sample": "%s, %d warm-up + %d timed SpMV, single thread (the reference loop is single-threaded), host has %d cores, "
                  "matrix built in %.0f s" % (what, warmup, len(times), os.cpu_count(), t_gen),
        "ms_per_step": 1e3 * tot / len(times), "nnz": nnz, "rows": rows, "cols": cols, "steps": len(times),
        "same_config": full, "y_checksum": float(np.sum(y)),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    warmup = min(args.warmup, 1)
    r = cpu_reference_sample(args.steps, warmup)
    steps = r["steps"]
    kind, rows, npr, desc = WORKLOADS["uniform"]
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "GNNZ/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "rows": r["rows"], "cols": r["cols"], "nnz": r["nnz"], "seed": SEED, "sample": r["sample"],
                   "same_config_as_gpu_arm": r["same_config"]},
        "cpu_baseline": {"value": r["value"], "unit": "GNNZ/s", "cores": 1, "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": "GNNZ/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gb_s_algorithmic": b_alg(r["nnz"], r["rows"], r["cols"]) / (r["ms_per_step"] * 1e-3) / 1e9,
    }
    print(json.dumps(line))


class Job:
    """Rendezvous + helpers shared by the headline measurement and the other configs of the same launch."""

    def __init__(self, args):
        import numpy as np
        import torch
        import b200spmv as b2
        self.np, self.torch, self.b2, self.args = np, torch, b2, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, v, op="max"):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.MIN)
        return float(t.item())

    def engine(self, comm, exchange=None, shard=True):
        """A context on this rank's GPU.  comm: attach the NCCL communicator (needed by the power iteration);
        otherwise the rank takes its row block by option (a single SpMV needs no collective)."""
        b2, args = self.b2, self.args
        eng = b2.Engine(self.local)
        if self.world > 1 and shard:
            if comm:
                uid = self.torch.zeros(128, dtype=self.torch.uint8, device="cuda")
                if self.rank == 0:
                    uid.copy_(self.torch.frombuffer(bytearray(b2.comm_unique_id()), dtype=self.torch.uint8))
                self.dist.broadcast(uid, 0)
                eng.comm_init(self.rank, self.world, bytes(uid.cpu().numpy().tobytes()))
            else:
                eng.set_option("world", self.world)
                eng.set_option("rank", self.rank)
        if args.shard:  # one rank's row block of a W-way split, alone on this GPU (tuning runs; no collective)
            r_, w_ = (int(v) for v in args.shard.split("/"))
            eng.set_option("world", w_)
            eng.set_option("rank", r_)
        if args.dcsr is not None:
            eng.set_option("dcsr_x16", args.dcsr)
        if args.slabs:
            eng.set_option("slabs", args.slabs)
        if args.kernel:
            eng.set_option("kernel", args.kernel)
        for opt in ("stream_db", "stream_pf", "stream_occ", "stream_vchunks"):
            if getattr(args, opt) is not None:
                eng.set_option(opt, getattr(args, opt))
        if args.prefetch is not None:
            eng.set_option("prefetch_tiles", args.prefetch)
        if args.graph is not None:
            eng.set_option("graph", args.graph)
        if args.own_slabs:
            eng.set_option("own_slabs", 1)
        if args.prefetch_x is not None:
            eng.set_option("prefetch_x", args.prefetch_x)
        if args.persist is not None:
            eng.set_option("persist", args.persist)
        if exchange in ("pull", "overlap", "packed") and self.world > 1 and shard and not args.shard:
            # slabs aligned with the ranks' x blocks: the exchange of the other blocks runs under the own-block slabs
            eng.set_option("own_slabs", args.own_layout)
            eng.set_option("exchange", {"pull": 2, "overlap": 1, "packed": 3}[exchange])
            if args.pull_streams:
                eng.set_option("pull_streams", args.pull_streams)
        return eng

    # ---- parity inside the bench run: no oracle here, the CUDA paths against each other ----
    def parity(self, kind, rows, cols, npr, power, exchange):
        """u64 twin of the timed matrix (same generator, same sparsity structure, same kernels, wrapping integer
        arithmetic so every order of summation gives the same bits): the throughput kernels against the strict
        row-sequential kernel on every row of this rank's block, y = A*x and y += A*x; for the power iteration
        3 distributed steps against 3 steps of a private single-GPU replica of the whole matrix (the replicated x
        must be bit-equal on every rank).  All ranks' verdicts are min-reduced."""
        np, b2 = self.np, self.b2
        ok, checked = True, 0
        e = self.engine(comm=power, exchange=exchange)
        e.setup_synthetic(kind, rows, cols, npr, SEED, b2.U64)
        e.fill_x_synthetic(SEED + 1)
        rb, re = e.stat("row_begin"), e.stat("row_end")
        e.fill_y_zero()
        e.exec_device(1, 0)
        ya = e.get_y()[rb:re].copy()
        e.exec_device(1, b2.ACCUMULATE)          # y += A*x on top: 2 * A*x mod 2^64
        yb = e.get_y()[rb:re].copy()
        e.exec_device(1, b2.STRICT_ORDER)
        ys = e.get_y()[rb:re].copy()
        ok = ok and ya.tobytes() == ys.tobytes() and yb.tobytes() == (ys + ys).tobytes()
        checked += re - rb
        detail = {"spmv_u64_vs_strict": bool(ya.tobytes() == ys.tobytes()), "accumulate": bool(yb.tobytes() == (ys + ys).tobytes())}
        if power and self.world > 1:
            e.power_iterate(3, False, 0)
            xd = e.get_x()
            e.close()
            e = self.engine(comm=False, shard=False)  # the whole matrix on this GPU alone
            e.setup_synthetic(kind, rows, cols, npr, SEED, b2.U64)
            e.power_iterate(3, False, 0)
            same = xd.tobytes() == e.get_x().tobytes()
            detail["power_3_steps_vs_single_gpu"] = bool(same)
            ok = ok and same
        e.close()
        ok_all = self.reduce(1.0 if ok else 0.0, "min") == 1.0
        rows_all = int(self.reduce(float(checked), "max")) if self.world == 1 else rows
        return {"ok": bool(ok_all), "checked_rows": rows_all, "how": "u64 twin of the same matrix: throughput kernels == strict "
                "row-sequential kernel bit for bit on every row (y = A*x and y += A*x)" +
                ("; 3 distributed power-iteration steps == single-GPU replica (replicated x, every rank)" if power and self.world > 1 else ""),
                "rank0": detail}

    # ---- one config, device-resident ---------------------------------------------------------------------------
    def measure(self, workload, steps, warmup, flags, exchange="pull", power=None, rows=None, cols=None, npr=None):
        """-> dict(value, ms_per_step, roofline, ...) for `workload` at this launch's N; leaves nothing allocated."""
        np, b2, args = self.np, self.b2, self.args
        kind, rows_d, npr_d, desc = WORKLOADS[workload]
        rows = rows or rows_d
        cols = cols or rows
        npr = npr or npr_d
        power = (workload == "banded-power" and not args.shard and not args.no_power) if power is None else power
        eng = self.engine(comm=True, exchange=exchange if power else None)
        t0 = time.perf_counter()
        eng.setup_synthetic(kind, rows, cols, npr, SEED, b2.F64)
        setup_s = time.perf_counter() - t0
        nnz, lnnz, lrows = eng.stat("nz"), eng.stat("local_nz"), eng.stat("local_rows")
        eng.fill_x_synthetic(SEED + 1)
        eng.fill_y_zero()
        out = {"desc": desc, "rows": rows, "cols": cols, "nnz": nnz, "setup_s": setup_s, "power": power}
        if power:
            eng.power_iterate(max(warmup, 3), False, flags & ~b2.ACCUMULATE)
        else:
            eng.exec_device(max(warmup, 3), flags)
        self.barrier()
        launches0 = eng.stat("launches")
        if power:
            ms_total, out["last_norm"] = eng.power_iterate(steps, True, flags & ~b2.ACCUMULATE)
        else:
            ms = eng.exec_device(steps, flags)  # CUDA events on the launching stream, one pair per step
            ms_total = float(ms.sum())
            out["step_stats"] = {"ms_best": float(ms.min()), "ms_median": float(np.median(ms))}  # this rank's steps
        self.barrier()
        out["launches"] = eng.stat("launches") - launches0
        t_dev = self.reduce(ms_total) * 1e-3
        out["t_dev"] = t_dev
        out["value"] = (lnnz if args.shard else nnz) * steps / t_dev / 1e9
        out["ms_per_step"] = 1e3 * t_dev / steps
        if power and self.world > 1:
            out["x_exchange_ms_last_step"] = eng.stat("xchgtime_ns") * 1e-6
            out["exchange"] = ["NCCL all-gather after the step", "NCCL all-gather under the own-block slabs",
                               "peer pulls block by block under the slabs",
                               "peer pulls of PACKED blocks (only the columns the puller gathers) under the slabs"][eng.stat("exchange_mode")]
        # the dominant kernel alone (roofline), same workload, events around each launch
        eng.set_option("profile", 1)
        psteps = min(steps, 20)
        eng.exec_device(psteps, flags)
        k_ns, k_n = eng.stat("merge_kernel_ns"), eng.stat("merge_kernel_launches")
        slabs = eng.stat("slabs")
        out["persist_window"] = bool(eng.stat("persist")) and not (flags & b2.NO_PERSIST)
        out["slab_kernels"] = "".join("d" if eng.stat("slab_dcsr_%d" % i) else "f" for i in range(slabs))
        out["tune_ms"] = round(eng.stat("tunetime_ns") * 1e-6, 1)
        if args.per_slab:
            out["per_slab"] = [{"ms": round(eng.stat("slab_ns_%d" % i) * 1e-6 / psteps, 4), "nnz": eng.stat("slab_nnz_%d" % i),
                                "nonempty_rows": eng.stat("slab_rows_%d" % i), "longest_row": eng.stat("slab_maxrow_%d" % i)}
                               for i in range(slabs)]
        eng.set_option("profile", 0)
        mode = eng.stat("mode")
        # the strict kernel has no per-launch bracket: its step time is the kernel time (one launch per slab)
        kernel_ms = k_ns * 1e-6 / k_n * slabs if k_n else 1e3 * t_dev / steps
        peak, peak_src = measured_peak_gbs()
        alg_local = 12 * lnnz + 4 * (lrows + 1) + 8 * cols + 8 * lrows
        achieved = alg_local / (kernel_ms * 1e-3) / 1e9
        traffic, traffic_src = recorded_traffic(workload, rows, cols, slabs, mode, self.world, flags)
        out.update(slabs=slabs, mode=mode, lrows=lrows, lnnz=lnnz)
        out["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                           "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                           "kernel": KERNEL_NAMES.get(mode, "?"), "kernel_ms_per_step": kernel_ms,
                           "kernel_launches_per_step": slabs, "algorithmic_bytes_per_step": alg_local,
                           "frac_of_8TBs_nominal": achieved / 8000.0}
        out["eng"] = eng
        return out


def run_ours(args):
    job = Job(args)
    np, torch, b2 = job.np, job.torch, job.b2
    rank, world, local = job.rank, job.world, job.local
    kind, rows, npr, desc = WORKLOADS[args.workload]
    rows = args.rows or rows
    cols = args.cols or rows
    npr = args.npr or npr
    flags = 0 if args.overwrite else b2.ACCUMULATE
    if args.no_persist:
        flags |= b2.NO_PERSIST
    if args.strict:
        flags |= b2.STRICT_ORDER

    sampler = ClockSampler(local)
    sampler.start()
    m = job.measure(args.workload, args.steps, args.warmup, flags, exchange=args.exchange, rows=rows, cols=cols, npr=npr)
    eng, nnz, lrows, power, t_dev = m["eng"], m["nnz"], m["lrows"], m["power"], m["t_dev"]

    # ---- end to end through b200spmv_exec with pinned HOST vectors --------------------------
    dt = torch.float64
    xh = torch.empty(cols, dtype=dt).pin_memory()
    yh = torch.zeros(rows, dtype=dt).pin_memory()
    xh.copy_(torch.from_numpy(eng.get_x()))
    e2e_steps = min(args.steps, 50)
    for _ in range(3):
        eng.exec(xh.data_ptr(), yh.data_ptr(), flags)
    job.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.exec(xh.data_ptr(), yh.data_ptr(), flags)  # H2D x (+y), kernels, D2H y; blocking
    torch.cuda.synchronize()
    t_e2e = job.reduce(time.perf_counter() - t0)
    timeline = {k: round(eng.stat("exec_%s_ns" % k) * 1e-6, 3) for k in ("x_up", "y_up", "kernels", "done")}
    job.barrier()
    clocks = sampler.stop()
    h2d, d2h = eng.stat("exec_h2d_bytes"), eng.stat("exec_d2h_bytes")  # per rank, counted from the copies b200spmv_exec issued
    # spot-check of the end-to-end path: one more exec from a zeroed host y against the device-resident result of
    # the same x (fp64, two re-associations of the same sums: |a|, |x| < 1, so 1e-9 absolute is ~1e5 ulp of slack
    # on a 32-term sum yet far below any wrong or missing entry)
    rb, re = eng.stat("row_begin"), eng.stat("row_end")
    yh.zero_()
    eng.exec(xh.data_ptr(), yh.data_ptr(), flags & ~b2.ACCUMULATE)
    eng.exec_device(1, flags & ~b2.ACCUMULATE)
    e2e_ok = bool(np.allclose(yh.numpy()[rb:re], eng.get_y()[rb:re], rtol=0, atol=1e-9))
    e2e_ok = job.reduce(1.0 if e2e_ok else 0.0, "min") == 1.0
    eng.close()

    parity = None
    if not args.no_parity:
        parity = job.parity(kind, rows, cols, npr, power, args.exchange)
        parity["e2e_exec_vs_device"] = e2e_ok

    # ---- the other configs BASELINE.json names, at the same N (driver-visible C4 / C5) -------------------------
    others = None
    if args.workload == "uniform" and not args.no_others and not args.shard:
        others = {}
        specs = [("rmat", "rmat", 20, None)]
        for ex in (("nccl", "pull", "packed") if world > 1 else ("pull",)):
            specs.append(("banded_power" + ("_" + ex if world > 1 else ""), "banded-power", 100, ex))
        for key, wl, k_steps, ex in specs:
            try:
                o = job.measure(wl, k_steps, 3, flags, exchange=ex or "pull")
                o.pop("eng").close()
                rec = {"workload": o["desc"], "value": o["value"], "unit": "GNNZ/s", "steps": k_steps, "ms_per_step": o["ms_per_step"],
                       "rows": o["rows"], "nnz": o["nnz"], "slabs": o["slabs"], "persist_window": o["persist_window"], "slab_kernels": o["slab_kernels"],
                       "setup_s": round(o["setup_s"], 2),
                       "roofline": {k: o["roofline"][k] for k in ("achieved", "peak", "frac", "kernel", "kernel_ms_per_step", "algorithmic_bytes_per_step")}}
                for k in ("x_exchange_ms_last_step", "exchange", "last_norm"):
                    if k in o:
                        rec[k] = o[k]
                if not args.no_parity:
                    kd, rw, nr, _ = WORKLOADS[wl]
                    rec["parity"] = job.parity(kd, rw, rw, nr, o["power"], ex or "pull")
                others[key] = rec
            except Exception as exc:  # a config that does not fit or fails must not take the headline line with it
                others[key] = {"error": "%s: %s" % (type(exc).__name__, exc)}

    bad = (parity is not None and not parity["ok"]) or (others and any(not v.get("parity", {"ok": True})["ok"] for v in others.values()))
    if rank == 0:
        line = {
            "metric": METRIC, "value": m["value"], "unit": "GNNZ/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": m["desc"], "rows": rows, "cols": cols, "nnz": nnz, "seed": SEED,
                       "semantics": "y = A*x" if args.overwrite else "y += A*x (reference semantics)",
                       "partition": "nnz-balanced row blocks, no collective" if not power else
                                    "row blocks + x all-gather per step",
                       "l2": "inputs (%.1f GB) larger than L2; no flush" % (b_alg(nnz, rows, cols) / 1e9),
                       "slabs": m["slabs"], "persist_window": m["persist_window"], "persist_window_how": "measured at setup (kept unless off is > 1 % faster)",
                       "slab_kernels": m["slab_kernels"], "setup_measurements_ms": m["tune_ms"], "setup_s": round(m["setup_s"], 2),
                       "setup_amortised_over_steps": int(round(m["setup_s"] / (t_dev / args.steps))), **m.get("step_stats", {}),
                       **({k: m[k] for k in ("x_exchange_ms_last_step", "exchange") if k in m}),
                       **({"norm_after_last_step": m["last_norm"]} if power else {}),
                       **({"per_slab": m["per_slab"]} if "per_slab" in m else {})},
            "gb_s_algorithmic": b_alg(nnz, rows, cols) * args.steps / t_dev / 1e9 if world == 1 else None,
            "roofline": m["roofline"],
            "e2e": {"value": nnz * e2e_steps / t_e2e / 1e9, "unit": "GNNZ/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "bytes_are": "per rank (rank 0)", "steps": e2e_steps, "ms_per_step": 1e3 * t_e2e / e2e_steps,
                    **({"x_exchange": "each rank uploads 1/%d of x over PCIe, ncclAllGather over NVLink" % world} if world > 1 else {}),
                    "last_step_timeline_ms": timeline},
            "gpu_launches": int(m["launches"]), "clocks": clocks,
        }
        if parity is not None:
            line["parity"] = parity
        if others is not None:
            line["other_configs"] = others
        if world == 1 and not args.no_cpu:
            r = cpu_reference_sample(2, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if job.dist:
        job.dist.destroy_process_group()
    if bad:
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="uniform", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override the row count (experiments only)")
    ap.add_argument("--npr", type=int, default=0, help="override entries per row (experiments only)")
    ap.add_argument("--shard", default="", help="R/W: run rank R's row block of a W-way split alone (experiments only)")
    ap.add_argument("--dcsr", type=int, default=None, help="hypersparse-slab threshold x16 (0 = never use spmv_dcsr_kernel)")
    ap.add_argument("--cols", type=int, default=0, help="override the column count (experiments only)")
    ap.add_argument("--slabs", type=int, default=0)
    ap.add_argument("--kernel", type=int, default=0, help="3 flagged segmented sum, one tile per CTA (default); 6 the same by persistent warps")
    ap.add_argument("--stream-db", type=int, default=None, help="kernel 6: register double buffer of the entry streams (default 0)")
    ap.add_argument("--stream-pf", type=int, default=None, help="kernel 6: L2 bulk prefetch distance in warp steps (default 1)")
    ap.add_argument("--stream-occ", type=int, default=None, help="kernel 6: CTAs per SM")
    ap.add_argument("--stream-vchunks", type=int, default=None, help="kernel 6: chunks per warp")
    ap.add_argument("--no-power", action="store_true", help="banded-power: time plain SpMV steps instead of the power iteration")
    ap.add_argument("--prefetch", type=int, default=None, help="kernel 3: L2 bulk prefetch distance in tiles")
    ap.add_argument("--graph", type=int, default=None, help="0: do not replay a CUDA graph in exec_device")
    ap.add_argument("--overwrite", action="store_true", help="y = A*x instead of y += A*x")
    ap.add_argument("--no-persist", action="store_true")
    ap.add_argument("--persist", type=int, default=None, help="L2 window on x: 1 on, 0 off (default: what setup measured)")
    ap.add_argument("--exchange", choices=["pull", "overlap", "nccl", "packed"], default="pull",
                    help="power iteration over several GPUs: how the replicated x is refreshed (DESIGN.md section 7)")
    ap.add_argument("--pull-streams", type=int, default=0)
    ap.add_argument("--own-layout", type=int, default=1, choices=[1, 2],
                    help="power iteration over several GPUs: 1 = every rank's block of x is its own slab(s); 2 = own block + one slab for "
                         "all blocks left of it + one for all blocks right of it (fewer passes over y)")
    ap.add_argument("--prefetch-x", type=int, default=None, help="1: bulk-prefetch each slab's x slice into L2 before its kernel")
    ap.add_argument("--own-slabs", action="store_true", help="with --shard: cut the slabs along the ranks' x blocks")
    ap.add_argument("--per-slab", action="store_true", help="add the kernel time of every slab to config")
    ap.add_argument("--strict", action="store_true", help="strict-order kernel (fp64 memcmp-equal to SoftwareSpMV)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run parity checks")
    ap.add_argument("--no-others", action="store_true", help="skip the other_configs (R-MAT, power iteration) of the default run")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
