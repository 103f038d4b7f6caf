"""CPU suite: the C-ABI library loads, exports every symbol include/b200spmv.h declares,
and refuses to run without a GPU (there is no CPU fallback in the product path)."""
import ctypes as C
import os
import re
import subprocess

import pytest

import helpers as H

PKG = os.path.join(H.ROOT, "spmv-vector-cache_b200")


@pytest.fixture(scope="module")
def libpath():
    p = os.path.join(PKG, "libb200spmv.so")
    if not os.path.exists(p):
        subprocess.check_call(["make", "-C", PKG, "-j8"])
    return p


def declared_symbols():
    with open(os.path.join(H.ROOT, "include", "b200spmv.h")) as f:
        return sorted(set(re.findall(r"\b(b200spmv_[a-z0-9_]+)\s*\(", f.read())))


def test_library_exports_every_declared_symbol(libpath):
    lib = C.CDLL(libpath)
    names = declared_symbols()
    assert len(names) >= 23
    for n in names:
        assert getattr(lib, n) is not None
    import b200spmv
    assert sorted(b200spmv.EXPORTS) == names


def test_no_extra_exports(libpath):
    out = subprocess.check_output(["nm", "-D", "--defined-only", libpath]).decode()
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert exported == declared_symbols()


def test_product_does_not_reference_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    for d, _, files in os.walk(PKG):
        if os.path.basename(d) == "build":
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".cpp", ".py", "Makefile")):
                with open(os.path.join(d, f), errors="ignore") as fh:
                    src = fh.read()
                assert "liboracle" not in src and "oracle/" not in src.replace("oracle/synth_host.c", ""), f


def test_create_fails_loudly_without_gpu(libpath):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = C.CDLL(libpath)
    lib.b200spmv_last_error.restype = C.c_char_p
    ctx = C.c_void_p()
    rc = lib.b200spmv_create(C.byref(ctx), 0)
    assert rc == -2 and not ctx.value
    assert b"no CPU fallback" in lib.b200spmv_last_error(None)
    import b200spmv
    with pytest.raises(b200spmv.B200Error):
        b200spmv.Engine(0)
