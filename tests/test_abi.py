"""The C-ABI libraries load and export every symbol their headers declare (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

from montecarlopathtracing_b200 import device, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(%s\w+)\s*\(" % prefix, text)))


def test_cuda_library_exports_every_declared_symbol():
    names = declared("mcpt.h", "mcpt_")
    assert len(names) >= 35
    L = ctypes.CDLL(device.library_path())
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(device.EXPORTS) == names                  # the binding covers the whole header
    assert L.mcpt_abi_version() == 1


def test_host_library_exports_every_declared_symbol():
    names = declared("mcpt_host.h", "mcpt_host_")
    L = host.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path needs a machine without one")
    with pytest.raises(device.McptError) as e:
        device.Context(16, 16, 4, 1)
    assert "CUDA" in str(e.value) or "device" in str(e.value)


def test_null_context_is_an_error_not_a_crash():
    L = device.lib()
    L.mcpt_render.argtypes = [ctypes.c_void_p, ctypes.c_int]
    assert L.mcpt_render(None, 1) < 0
    assert b"null context" in L.mcpt_last_error()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "montecarlopathtracing_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) and f != "smoke.py":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_api" not in text and "liboracle" not in text and "mcpt_oracle" not in text, f
