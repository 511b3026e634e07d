"""CPU-side checks of the drop-in boundary: the library loads without a GPU and exports
exactly the symbols include/pathnet_b200.h declares; no compute entry point is called."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "pathnet_b200.h")).read()
    return sorted(set(re.findall(r"PN_API\s+[A-Za-z0-9_\* ]+?\b(pn_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(native_lib):
    from pathnet_gym_b200 import _native
    decl = declared_symbols()
    assert len(decl) >= 20
    out = subprocess.run(["nm", "-D", "--defined-only", _native.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (pn_[a-z0-9_]+)", out)))
    assert exported == decl
    assert sorted(_native.SIGNATURES) == decl           # the ctypes binding covers every symbol
    for name in decl:
        assert hasattr(native_lib, name)


def test_abi_version_and_errors_without_gpu(native_lib):
    import ctypes as C
    from pathnet_gym_b200 import _native
    assert native_lib.pn_abi_version() == _native.PN_ABI_VERSION
    # invalid configuration is rejected before any CUDA call
    pc = _native.PnConfig()
    pc.L, pc.M, pc.F = 1, 10, 8
    h = C.c_void_p()
    assert native_lib.pn_create(C.byref(pc), C.byref(h)) == _native.PN_ERR_INVALID
    assert b"L = 1" in native_lib.pn_last_global_error()
    assert native_lib.pn_geopath_init(None, 4, 0, 3, 40, 3, 1, 0, None) == _native.PN_ERR_INVALID


def test_product_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from pathnet_gym_b200.population import PathNetPopulation, PathNetConfig
    with pytest.raises(RuntimeError):
        PathNetPopulation(PathNetConfig(), 1, 1)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pathnet_gym_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f
