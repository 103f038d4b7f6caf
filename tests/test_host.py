"""Host C++ side of the drop-in (spmv-vector-cache_b200/host): CudaSpMV : SpMV, SpMVFactory, the
matrix-directory loader and the port of the reference's benchmark loop (software/main.cpp:146-264).

CPU tests: it builds, refuses to run without a GPU (no CPU fallback), rejects the FPGA variant names.
GPU tests: the CSV transcript has diffFromGolden = 0 against golden.bin (fp64 directories), against
oracle-written golden vectors (uint64 directories), and -- oracle/_ref/spmvbench_ref, the same host code
compiled against the reference's own headers -- against SoftwareSpMV running in the same process."""
import os
import shutil
import subprocess

import numpy as np
import pytest

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "spmv-vector-cache_b200", "host")
REFBENCH = os.path.join(ROOT, "oracle", "_ref", "spmvbench_ref")


def have_gpu():
    return os.path.exists("/dev/nvidiactl") or os.path.exists("/dev/nvidia0")


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(os.path.join(HOST, "..", "libb200spmv.so")):
        subprocess.check_call(["make", "-C", os.path.join(HOST, ".."), "-j8"])
    subprocess.check_call(["make", "-C", HOST, "-j8"], stdout=subprocess.DEVNULL)
    return HOST


def run_bench(exe, root, confs, matrices, extra=()):
    p = subprocess.run([exe, "--root", root, "--confs", confs, "--matrices", matrices, *extra],
                       capture_output=True, text=True, timeout=300)
    return p.returncode, p.stdout, p.stderr


def parse_csv(out):
    """-> list of dict rows keyed by the header line (the line that starts with diffFromGolden)."""
    rows, keys = [], None
    for line in out.splitlines():
        cells = [c for c in line.strip().split(",")]
        if line.startswith("diffFromGolden,") or line.startswith("rows,"):
            keys = cells[:-1]
        elif keys and len(cells) == len(keys) + 1 and not line.startswith("#"):
            rows.append(dict(zip(keys, cells[:-1])))
    return rows


def test_host_builds_and_has_no_cpu_fallback(built):
    if have_gpu():
        pytest.skip("GPU present: covered by the gpu tests")
    rc, out, err = run_bench(os.path.join(built, "spmvbench"), H.MATRICES, "cuda", "circuit204")
    assert rc != 0
    assert "no CPU fallback" in err
    row = parse_csv(out)[0]
    assert row["diffFromGolden"] == "1" and row["cudaError"] != "0" and row["accType"] == "CUDA-B200"


def test_factory_rejects_fpga_and_unknown_names(built):
    rc, out, err = run_bench(os.path.join(built, "spmvbench"), H.MATRICES, "NewCache,bogus,sw", "i64")
    assert rc != 0
    assert "requires the Zynq programmable logic" in out
    assert "Accelerator signature unrecognized! bogus" in out
    assert "SoftwareSpMV is the reference's CPU code" in out


def test_loader_rejects_invalid_directory(built, tmp_path):
    d = tmp_path / "bad"
    d.mkdir()
    np.zeros(7, dtype="<u4").tofile(d / "bad-meta.bin")  # zero rows / cols / nz: SparseMatrix.cpp:33-38
    rc, out, err = run_bench(os.path.join(built, "spmvbench"), str(tmp_path), "cuda", "bad,missing")
    assert rc != 0 and "is invalid" in err and "cannot read" in err


@pytest.mark.gpu
def test_csv_transcript_fp64_against_golden_bin(built):
    names = ",".join(H.FP64_DIRS)
    rc, out, err = run_bench(os.path.join(built, "spmvbench"), H.MATRICES, "cuda-strict,cuda,cuda-nopersist,cuda-overwrite", names,
                             ["--iters", "3"])
    assert rc == 0, err
    rows = parse_csv(out)
    assert len(rows) == 4 * len(H.FP64_DIRS)
    for r in rows:
        m = H.load_matrix_dir(os.path.join(H.MATRICES, r["matrix"]))
        assert (int(r["rows"]), int(r["cols"]), int(r["nz"])) == (m.rows, m.cols, m.nnz)
        assert r["cudaError"] == "0" and r["accType"] == "CUDA-B200" and r["gpus"] == "1"
    strict = [r for r in rows if r["mode"] == "0"]
    assert len(strict) == len(H.FP64_DIRS) and all(r["diffFromGolden"] == "0" for r in strict)
    # re-associated sums: still bit-equal wherever rows have <= 1 entry or only exact sums (identity, row64k)
    for r in rows:
        if r["matrix"] in ("i64", "i1k", "i64k", "row64k"):
            assert r["diffFromGolden"] == "0"
    assert "GNNZ/s" in out


@pytest.mark.gpu
def test_csv_transcript_uint64_against_oracle_golden(built, oracle, tmp_path):
    names = [n for n in H.ALL_DIRS if n.endswith("-uint64")]
    for n in names:  # the uint64 directories ship no golden.bin: write the oracle's A*1 next to a copy
        shutil.copytree(os.path.join(H.MATRICES, n), tmp_path / n)
        m = H.load_matrix_dir(os.path.join(H.MATRICES, n))
        vals = m.data_raw.view(np.uint64)
        y = oracle.spmv_csc("u64", m.rows, m.cols, m.colptr, m.rowind, vals, np.ones(m.cols, np.uint64), np.zeros(m.rows, np.uint64))
        y.tofile(tmp_path / n / "golden.bin")
    rc, out, err = run_bench(os.path.join(built, "spmvbench_u64"), str(tmp_path), "cuda,cuda-strict,cuda-nopersist", ",".join(names))
    assert rc == 0, err
    rows = parse_csv(out)
    assert len(rows) == 3 * len(names) and all(r["diffFromGolden"] == "0" and r["cudaError"] == "0" for r in rows)


@pytest.mark.gpu
def test_same_host_code_inside_the_reference_interface():
    """oracle/_ref/spmvbench_ref = our CudaSpMV / factory compiled against the reference's SpMV.h and
    SparseMatrix.h and linked with its SoftwareSpMV.cpp: the golden vector is SoftwareSpMV's, in process
    (software/main.cpp:222-247), with and without markRowStarts (cold-miss-skip bits set in the indices)."""
    if not os.path.exists(REFBENCH):
        pytest.skip("oracle/_ref/spmvbench_ref not built (needs /root/reference at build time)")
    names = ",".join(H.FP64_DIRS)
    for cms in ("0", "1"):
        rc, out, err = run_bench(REFBENCH, H.MATRICES, "cuda-strict", names, ["--cms", cms])
        assert rc == 0, err
        rows = parse_csv(out)
        assert len(rows) == len(H.FP64_DIRS) and all(r["diffFromGolden"] == "0" for r in rows), out
    rc, out, err = run_bench(REFBENCH, H.MATRICES, "sw", "circuit204")
    assert rc == 0 and "spmvtime" in out


@pytest.mark.gpu
def test_preprocessing_and_vector_cache_vocabularies(built, oracle):
    """The reference's two other stat vocabularies through the host class and the CLI (missing 4-5):
    conf token `prep` = SoftwareSpMV::measurePreprocessingTimes (SoftwareSpMV.cpp:72-94) on the device -- maxAlive and
    maxColSpan must equal the reference's methods (oracle restatement, itself pinned against the compiled reference);
    `vcmodel<depth>` = the NewCache counters of HardwareSpMVNewCache.cpp:189-204 that a functional model can know,
    equal to the sequential oracle model for the same depth, with and without cold-miss skip."""
    names = ["circuit204", "i1k", "row64k"]
    confs = "cuda-prep-strict,cuda-vcmodel64,cuda-vcmodel1024-cms"
    rc, out, err = run_bench(os.path.join(built, "spmvbench"), H.MATRICES, confs, ",".join(names))
    assert rc == 0, err
    rows = parse_csv(out)
    assert len(rows) == 3 * len(names)
    for key in ("cmstime", "maxAliveTime", "maxColSpanTime", "maxAlive", "maxColSpan"):  # SoftwareSpMV.cpp:96-108
        assert key in rows[0]
    for r in rows[:len(names)]:
        m = H.load_matrix_dir(os.path.join(H.MATRICES, r["matrix"]))
        inds = m.rowind.copy()
        assert int(r["maxColSpan"]) == oracle.lib.oracle_max_col_span(H.C.c_uint32(m.cols), H.C.c_uint32(m.nnz), H.ptr(m.colptr), H.ptr(inds))
        assert int(r["maxAlive"]) == oracle.lib.oracle_max_alive(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(inds))
        assert r["diffFromGolden"] == "0"
    # the vcmodel confs print their own header (keys are built per conf list in the reference too: first run wins),
    # so re-run them alone to get their columns
    for conf, depth_log2, cms in (("cuda-vcmodel64", 6, 0), ("cuda-vcmodel1024-cms", 10, 1)):
        rc, out, err = run_bench(os.path.join(built, "spmvbench"), H.MATRICES, conf, ",".join(names))
        assert rc == 0, err
        for r in parse_csv(out):
            m = H.load_matrix_dir(os.path.join(H.MATRICES, r["matrix"]))
            inds = m.rowind.copy()
            if cms:
                oracle.lib.oracle_mark_row_starts(H.C.c_uint32(m.rows), H.C.c_uint32(m.nnz), H.ptr(inds), H.C.c_int(0), H.C.c_int(31))
            out7 = np.zeros(7, dtype=np.uint64)
            oracle.lib.oracle_vector_cache_model(H.C.c_uint32(inds.size), H.ptr(inds), H.C.c_int(depth_log2), H.C.c_int(cms), H.ptr(out7))
            want = dict(zip(("reads", "readMisses", "conflictEvictions", "coldMissSkips", "ddrReads", "flushWrites", "hits"), (int(v) for v in out7)))
            assert int(r["ocmDepth"]) == 1 << depth_log2 and int(r["cms"]) == cms
            for k in ("reads", "readMisses", "conflictEvictions", "coldMissSkips", "ddrReads", "flushWrites", "hits"):
                assert int(r[k]) == want[k], (conf, r["matrix"], k)
