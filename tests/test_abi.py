"""CPU tests of the boundary: libbgpu.so loads, exports every symbol include/bgpu.h declares, and fails loudly
without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "bgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bgpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(bb):
    L = C.CDLL(bb.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/bgpu.h but not exported by libbgpu.so"


def test_python_binding_covers_header(bb):
    assert sorted(bb.SIGNATURES) == declared_symbols()


def test_header_compiles_as_c():
    import subprocess
    subprocess.check_call(["gcc", "-std=c99", "-fsyntax-only", "-x", "c", os.path.join(ROOT, "include", "bgpu.h")])


def test_no_device_fails_loudly(bb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(bb.BgpuError) as e:
        bb.Context(0)
    assert e.value.code == 4 and "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "arrow-ballista_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"import oracle|from oracle|#include\s*[<\"][^\n]*oracle|liboracle", txt), f
