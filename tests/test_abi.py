"""CPU tests: the C-ABI library loads, exports every symbol include/cavept.h declares, and
fails loudly (error code, no fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "cavept.h")).read()
    return re.findall(r"CAVEPT_API\s+[\w\s\*]+?\b(cavept_\w+)\s*\(", text)


@pytest.fixture(scope="module")
def lib():
    from caveengine_b200 import build

    build.build()
    from caveengine_b200 import load_library

    return load_library()


def test_header_and_binding_agree(lib):
    from caveengine_b200 import api

    declared = _declared_symbols()
    assert len(declared) >= 25
    bound = [name for name, _, _ in api.ABI]
    assert sorted(declared) == sorted(bound)


def test_every_declared_symbol_is_exported(lib):
    for name in _declared_symbols():
        assert hasattr(lib, name), name


def test_no_oracle_in_product():
    """The product library and package must not reference the oracle."""
    import subprocess

    from caveengine_b200 import library_path

    needed = subprocess.run(["ldd", library_path()], capture_output=True, text=True).stdout
    assert "cave_oracle" not in needed and "cave_ref" not in needed
    pkg = os.path.join(ROOT, "caveengine_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in src and "from oracle" not in src and "cave_oracle" not in src, f


def test_create_without_gpu_fails_loudly(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from caveengine_b200 import CaveptError, PathTracer

    with pytest.raises(CaveptError) as e:
        PathTracer(64, 36, 5)
    assert e.value.code == 3  # CAVEPT_ERR_CUDA


def test_invalid_arguments(lib):
    from caveengine_b200 import api

    ctx = C.c_void_p()
    cfg = api._Config(-1, 0, 0, 5, 0, 0)
    assert lib.cavept_create(C.byref(cfg), C.byref(ctx)) == api.ERR_INVALID_ARG
    assert lib.cavept_create(None, C.byref(ctx)) == api.ERR_INVALID_ARG
    assert lib.cavept_is_active(None) == 0
    assert lib.cavept_render(None, 0, 1) == api.ERR_INVALID_ARG
