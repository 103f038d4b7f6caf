"""MatrixMarket reader + matrix-directory writer (host/MatrixDirectory.cpp, host/mtx2dir.cpp; SURVEY.md
8f-2): regenerate the reference's circuit204 and circuit204-uint64 directories from its own
matrices/mtx/circuit204.mtx (tests/golden/circuit204.mtx) and compare BYTES with the fixtures the reference
ships (matrices/matrixutils.py:186-260 wrote them).  golden.bin needs the GPU (CudaSpMV strict order)."""
import os
import subprocess

import numpy as np
import pytest

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "spmv-vector-cache_b200", "host")
MTX = os.path.join(ROOT, "tests", "golden", "circuit204.mtx")


@pytest.fixture(scope="module")
def mtx2dir():
    subprocess.check_call(["make", "-C", HOST, "-j8"], stdout=subprocess.DEVNULL)
    return os.path.join(HOST, "mtx2dir")


def same_bytes(a, b):
    with open(a, "rb") as fa, open(b, "rb") as fb:
        return fa.read() == fb.read()


@pytest.mark.parametrize("name,extra", [("circuit204", ["--no-golden"]), ("circuit204-uint64", ["--uint64"])])
def test_directory_files_are_byte_identical(mtx2dir, tmp_path, name, extra):
    p = subprocess.run([mtx2dir, "--mtx", MTX, "--root", str(tmp_path), "--name", name, *extra], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert "Rows = 1020" in p.stdout and "NonZ = 5883" in p.stdout
    for part in ("meta", "indptr", "inds", "data"):
        assert same_bytes(tmp_path / name / f"{name}-{part}.bin", os.path.join(H.MATRICES, name, f"{name}-{part}.bin")), part


def test_symmetric_pattern_and_duplicates(mtx2dir, tmp_path):
    (tmp_path / "s.mtx").write_text("%%MatrixMarket matrix coordinate real symmetric\n% c\n3 3 4\n1 1 2.0\n3 1 5.0\n3 1 0.5\n2 2 -1.0\n")
    (tmp_path / "p.mtx").write_text("%%MatrixMarket matrix coordinate pattern general\n2 3 2\n2 3\n1 1\n")
    assert subprocess.run([mtx2dir, "--mtx", str(tmp_path / "s.mtx"), "--root", str(tmp_path), "--name", "s", "--no-golden"]).returncode == 0
    m = H.load_matrix_dir(str(tmp_path / "s"))
    assert (m.rows, m.cols, m.nnz) == (3, 3, 4)
    assert m.colptr.tolist() == [0, 2, 3, 4] and m.rowind.tolist() == [0, 2, 1, 0]
    assert m.data_raw.view(np.float64).tolist() == [2.0, 5.5, -1.0, 5.5]  # duplicates summed, mirrored entry added
    assert subprocess.run([mtx2dir, "--mtx", str(tmp_path / "p.mtx"), "--root", str(tmp_path), "--name", "p", "--no-golden"]).returncode == 0
    m = H.load_matrix_dir(str(tmp_path / "p"))
    assert (m.rows, m.cols, m.nnz) == (2, 3, 2) and m.colptr.tolist() == [0, 1, 1, 2] and m.data_raw.view(np.float64).tolist() == [1.0, 1.0]


def test_bad_input_is_refused(mtx2dir, tmp_path):
    (tmp_path / "bad.mtx").write_text("%%MatrixMarket matrix array real general\n2 2\n1\n2\n3\n4\n")
    p = subprocess.run([mtx2dir, "--mtx", str(tmp_path / "bad.mtx"), "--root", str(tmp_path), "--name", "bad"], capture_output=True, text=True)
    assert p.returncode == 1 and "coordinate" in p.stderr


@pytest.mark.gpu
def test_golden_bin_from_the_gpu_is_byte_identical(mtx2dir, tmp_path):
    p = subprocess.run([mtx2dir, "--mtx", MTX, "--root", str(tmp_path), "--name", "circuit204"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert same_bytes(tmp_path / "circuit204" / "golden.bin", os.path.join(H.MATRICES, "circuit204", "golden.bin"))
