import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "spmv-vector-cache_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu under gpurun)")
    # the checker library is cheap to (re)build and has no GPU dependency
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])


@pytest.fixture(scope="session")
def oracle():
    import helpers
    return helpers.OracleLib()


@pytest.fixture(scope="session")
def reflib():
    import helpers
    if not helpers.RefLib.available():
        pytest.skip("oracle/_ref not built (needs /root/reference; run `make -C oracle ref`)")
    return helpers.RefLib()


@pytest.fixture(scope="session")
def ref_outputs():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_outputs.npz"))
