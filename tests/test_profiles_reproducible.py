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


def test_recorded_traffic_is_the_sum_of_the_captured_launches():
    rec = json.load(open(os.path.join(PROF, "traffic.json")))[0]
    summary = open(os.path.join(PROF, "r01_k3_final_ncu_summary.txt")).read()
    line = [l for l in summary.splitlines() if "sum over the step" in l][0]
    assert int(line.split(":")[1].split()[0]) == rec["dram_bytes_per_step"]
    # three launches (slabs) per step on C3, each about 2.4 GB read + 0.135 GB written
    per_launch = [l for l in summary.splitlines() if l.strip().startswith("read ")]
    assert len(per_launch) == rec["slabs"] == 3


def test_bench_lines_under_profiles_carry_the_contract_keys():
    need = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
            "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks"}
    for name in ("r01_bench_default.json", "r01_bench_8gpu.json", "r01_bench_8gpu_rmat.json", "r01_bench_8gpu_power_pull.json"):
        d = json.loads(open(os.path.join(PROF, name)).read().strip().splitlines()[-1])
        assert need <= set(d), (name, need - set(d))
        assert set(d["roofline"]) >= {"bound", "achieved", "peak", "unit", "frac", "traffic"}
        assert d["gpu_launches"] > 0 and d["dtype"] == "f64" and "workload" in d["config"]
    ref = json.loads(open(os.path.join(PROF, "r01_bench_reference.json")).read().strip().splitlines()[-1])
    assert ref["impl"] == "reference" and ref["cpu_baseline"]["kind"] in ("reference", "port")
    assert ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["e2e"]["d2h_bytes_per_step"] == 0
