"""CPU tests of the boundary: libsr_cuda.so loads, exports every symbol include/sr_cuda.h declares, and refuses to
run without a CUDA device (no CPU fallback)."""
import os
import re

import pytest

import software_renderer_b200 as srb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "sr_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sr_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    lib = srb.load_library()
    syms = declared_symbols()
    assert set(syms) == set(srb.ABI_SYMBOLS)
    for s in syms:
        assert getattr(lib, s) is not None


def test_version():
    assert srb.load_library().sr_version() == 1000


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(srb.SrError) as e:
        srb.CudaRenderer(0)
    assert e.value.code == -4


def test_product_package_does_not_import_oracle():
    pkg = os.path.join(ROOT, "software_renderer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "sr_oracle" not in text and "import oracle" not in text and "from oracle" not in text, f
