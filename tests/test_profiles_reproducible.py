"""The summaries under profiles/ are derived files: regenerate them from the committed raw captures with the
committed tools and compare, so that a number quoted in DESIGN.md can be traced to an ncu line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")


def test_launch_list_summary_matches_the_capture():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "launch_summary.py"), os.path.join(PROF, "r01_launches.csv")],
                         capture_output=True, text=True, check=True).stdout.strip().splitlines()
    committed = open(os.path.join(PROF, "r01_launch_list_summary.txt")).read().splitlines()
    for line in out:
        assert line in committed, line
    shares = [l for l in out if l.startswith("spmv_flag_kernel")]
    assert shares, "the SpMV kernel must appear in the launch list"


def test_round2_launch_list_summary_matches_the_capture():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "launch_summary.py"), os.path.join(PROF, "r02_launches.csv")],
                         capture_output=True, text=True, check=True).stdout.strip().splitlines()
    committed = open(os.path.join(PROF, "r02_launch_list_summary.txt")).read().splitlines()
    for line in out:
        assert line in committed, line
    flag = [l for l in out if l.startswith("within the SpMV steps")]
    assert flag and float(flag[0].split("spmv_flag_kernel")[1].split("%")[0]) > 90.0  # the kernel's share of the step


def test_recorded_traffic_is_the_sum_of_the_captured_launches():
    """every record of traffic.json (what bench.py prints as roofline.traffic) names the summary it was read from"""
    recs = json.load(open(os.path.join(PROF, "traffic.json")))
    assert recs[0]["workload"] == "uniform" and recs[0]["slabs"] == 3
    for rec in recs:
        summary = open(os.path.join(ROOT, rec["source"].split(":")[0])).read()
        line = [l for l in summary.splitlines() if "sum over the step" in l][0]
        assert int(line.split(":")[1].split()[0]) == rec["dram_bytes_per_step"]
        per_launch = [l for l in summary.splitlines() if l.strip().startswith("read ")]
        assert len(per_launch) == rec["slabs"]  # one launch per slab and step


def test_l2_ab_table_matches_the_raw_counters():
    """profiles/r02_l2_ab.txt: the x-gather hit rates are the evict_last (window on) / evict_normal (NO_PERSIST) lookup
    counters of the committed per-launch CSVs"""
    import csv
    text = open(os.path.join(PROF, "r02_l2_ab.txt")).read()
    for name, cls in (("r02_l2ab_3slabs_window.csv", "evict_last"), ("r02_l2ab_3slabs_nopersist.csv", "evict_normal"), ("r02_l2ab_1slab.csv", "evict_last")):
        rows = list(csv.DictReader(l for l in open(os.path.join(PROF, name)) if not l.startswith("==")))
        per = {}
        for r in rows:
            per.setdefault(r["ID"], {})[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
        for m in per.values():
            h = m["lts__t_sectors_srcunit_tex_op_read_%s_lookup_hit.sum" % cls]
            mi = m["lts__t_sectors_srcunit_tex_op_read_%s_lookup_miss.sum" % cls]
            assert "%11d / %10d   (%5.2f)" % (h, mi, 100 * h / (h + mi)) in text


def test_bench_lines_under_profiles_carry_the_contract_keys():
    need = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
            "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks"}
    for name in ("r01_bench_default.json", "r01_bench_8gpu.json", "r01_bench_8gpu_rmat.json", "r01_bench_8gpu_power_pull.json",
                 "r02_bench_1gpu.json", "r02_bench_2gpu.json", "r02_bench_4gpu.json", "r02_bench_8gpu.json",
                 "r02_bench_1gpu_band.json", "r02_bench_1gpu_banded.json"):
        d = json.loads(open(os.path.join(PROF, name)).read().strip().splitlines()[-1])
        assert need <= set(d), (name, need - set(d))
        assert set(d["roofline"]) >= {"bound", "achieved", "peak", "unit", "frac", "traffic"}
        assert d["gpu_launches"] > 0 and d["dtype"] == "f64" and "workload" in d["config"]
    ref = json.loads(open(os.path.join(PROF, "r01_bench_reference.json")).read().strip().splitlines()[-1])
    assert ref["impl"] == "reference" and ref["cpu_baseline"]["kind"] in ("reference", "port")
    assert ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["e2e"]["d2h_bytes_per_step"] == 0


def test_round2_bench_lines_carry_parity_and_the_other_configs():
    """correctness on the bench line, C4 and C5 next to C3 at every N"""
    vals = {}
    for n in (1, 2, 4, 8):
        d = json.loads(open(os.path.join(PROF, "r02_bench_%dgpu.json" % n)).read().strip().splitlines()[-1])
        assert d["n_gpus"] == n and d["parity"]["ok"] is True and d["parity"]["checked_rows"] == d["config"]["rows"]
        oc = d["other_configs"]
        power = ["banded_power"] if n == 1 else ["banded_power_nccl", "banded_power_pull"]
        for k in ["rmat"] + power:
            assert oc[k]["parity"]["ok"] is True and oc[k]["roofline"]["frac"] > 0
            if k.startswith("banded_power"):
                assert oc[k]["steps"] == 100
        vals[n] = (d["value"], oc["rmat"]["value"], oc[power[-1]]["value"], d["e2e"]["value"])
    assert vals[8][0] / vals[1][0] >= 6 and vals[8][1] / vals[1][1] >= 6 and vals[8][2] / vals[1][2] >= 6
    assert vals[8][3] > 2 * vals[1][3]  # end to end now scales (each rank uploads 1/N of x)


def test_band_workload_is_hbm_bound_and_one_slab():
    """pure band at >= 0.70 of the measured HBM peak; setup recognises the band and does not cut it"""
    d = json.loads(open(os.path.join(PROF, "r02_bench_1gpu_band.json")).read().strip().splitlines()[-1])
    assert d["config"]["slabs"] == 1 and d["roofline"]["frac"] >= 0.70 and d["parity"]["ok"] is True
    assert d["roofline"]["traffic"] is not None and d["roofline"]["traffic"] < 1.1 * d["roofline"]["algorithmic_bytes_per_step"]
