"""The C-ABI library loads on a machine without a GPU and exports every symbol include/feelgpu.h declares;
without a device the entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import feelpp_b200 as fpp
from feelpp_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "feelgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(feelgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(capi.gpu_lib_path())
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n
    # and the Python binding table covers the header
    assert set(capi.GPU_SYMBOLS) | {"feelgpu_last_error", "feelgpu_eplan_host_error", "feelgpu_rplan_host_error"} == set(names)


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return
    try:
        fpp.Engine(0)
    except fpp.FeelGpuError as e:
        assert e.code == 2 and "CUDA" in str(e)
    else:
        raise AssertionError("context_create must fail without a CUDA device")


def test_product_does_not_reference_the_oracle():
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "feelpp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"feel_oracle|from oracle|import oracle|oracle/", txt):
                    bad.append(os.path.join(d, f))
    assert not bad, bad
