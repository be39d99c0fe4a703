"""The C-ABI library builds for sm_100a, loads, and exports every symbol include/mpcc_b200.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re
import subprocess

from conftest import ROOT
from pyscf_mpcc_b200 import _cabi


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mpcc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mpccb_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    _cabi.build()
    assert os.path.isfile(_cabi.LIB_PATH)
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"


def test_binding_covers_header():
    assert sorted(_cabi.SIGNATURES) == declared_symbols()
    lib = _cabi.load()
    assert b"sm_100a" in lib.mpccb_version()
    assert lib.mpccb_profile_stages() == 8


def test_library_is_sm100a_dmma_code():
    out = subprocess.run(["cuobjdump", "-sass", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert out.count("DMMA.8x8x4") > 500          # FP64 tensor-core atoms in the GEMM kernels


def test_no_product_import_of_oracle():
    """The product package must never route through the oracle (CPU fallback check)."""
    pkg = os.path.join(ROOT, "pyscf_mpcc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
