import sys, json
d = json.loads(sys.stdin.read())
ps = d["config"].get("per_slab") or []
print("%.4f ms/step  %.1f GNNZ/s  slabs %d" % (d["ms_per_step"], d["value"], d["config"]["slabs"]))
for i, p in enumerate(ps):
    print("   slab %2d  %.3f ms  nnz %10d  nonempty rows %9d  (%.2f per row, longest %d, %.1f G/s)" %
          (i, p["ms"], p["nnz"], p["nonempty_rows"], p["nnz"] / max(1, p["nonempty_rows"]), p.get("longest_row", 0), p["nnz"] / max(1e-9, p["ms"]) / 1e6))
