#!/usr/bin/env python
"""One-screen summary of bench.py JSON lines: show_bench.py FILE..."""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    c = d["config"]
    print("%s: N=%d value %.1f ms %.4f frac %.3f (peak %.1f) kernel_ms %.4f e2e %.1f h2d %d parity %s slabs %s window %s kernels %s clocks %s" % (
        f, d["n_gpus"], d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["peak"], d["roofline"]["kernel_ms_per_step"],
        d["e2e"]["value"], d["e2e"]["h2d_bytes_per_step"], d["parity"]["ok"], c.get("slabs"), c.get("persist_window"), c.get("slab_kernels"),
        d["clocks"].get("sm_mhz")))
    print("   e2e timeline", d["e2e"].get("last_step_timeline_ms"), "cpu", d.get("cpu_baseline", {}).get("value"))
    for k, v in d.get("other_configs", {}).items():
        print("   %-20s %.1f GNNZ/s  %.3f ms/step slabs %s window %s kernels %s frac %.3f parity %s exch %s" % (
            k, v["value"], v["ms_per_step"], v.get("slabs"), v.get("persist_window"), v.get("slab_kernels"), v["roofline"]["frac"],
            v["parity"]["ok"], v.get("x_exchange_ms_last_step")))
