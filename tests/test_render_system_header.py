"""include/cavept_render_system.hpp -- the header-only C++ facade with the shape of
cave/engine/renderer/path_tracer_render_system.h:10-21 -- compiles as C++20 with -Wall -Wextra -Werror against
include/cavept.h, links with libcavept.so and behaves: error codes without a device (CPU test), the reference's
call sequence on a device (GPU test)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "render_system_check.cpp")


def _build(tmp_path):
    from caveengine_b200 import build

    lib = build.build()
    exe = str(tmp_path / "render_system_check")
    libdir = os.path.dirname(lib)
    cmd = ["g++", "-std=c++20", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe, "-L", libdir, "-l:libcavept.so", f"-Wl,-rpath,{libdir}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-3000:]
    return exe


def test_header_compiles_and_fails_loudly_without_a_device(tmp_path):
    exe = _build(tmp_path)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "OK" in res.stdout


def test_header_is_plain_c_compatible_underneath(tmp_path):
    """include/cavept.h itself compiles as C11 (the boundary is a C ABI)."""
    c = tmp_path / "abi.c"
    c.write_text('#include "cavept.h"\nint main(void) { cavept_config c = {0}; (void)c; return sizeof(cavept_stats) > 0 ? 0 : 1; }\n')
    res = subprocess.run(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(c), "-o", str(tmp_path / "abi.o")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]


@pytest.mark.gpu
def test_header_drives_the_device_like_the_reference_facade(tmp_path):
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    exe = _build(tmp_path)
    res = subprocess.run([exe, "gpu"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "OK (device path)" in res.stdout
