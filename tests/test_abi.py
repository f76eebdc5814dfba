"""The C-ABI library: it loads, and it exports every symbol include/peng_cuda.h declares (no compute calls:
this runs without a GPU)."""
import ctypes
import os
import re

import pytest


def test_header_symbols_exported(pkg):
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build()
    hdr = open(pkg.HEADER_PATH).read()
    declared = set(re.findall(r"\b(peng_cuda_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = ctypes.CDLL(pkg.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/peng_cuda.h but not exported"
    assert set(pkg.exported_symbols()) == declared, "Python prototypes out of sync with the header"
    pkg.load_library()


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the product must fail loudly instead of computing on the CPU."""
    lib = pkg.load_library()
    if lib.peng_cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.PengCudaError) as e:
        pkg.PengCuda(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_reference_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg_dir = os.path.join(root, "peng-motif_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "peng_oracle" not in text and "_oracle" not in text and "libpeng_sim" not in text, os.path.join(dirpath, f)
