"""Builds libcavept.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcavept.so")
SOURCES = ["kernels.cu", "cavept_api.cu", "cavept_group.cu", "cavept_static.cu", "scene_ingest.cpp"]
HEADERS = ["cv_common.h", "bvh_parity_body.h", "bvh_parity_driver.h", "wbvh_body.h", "trace_body.h", "trace_pool_body.h", "host_math.h", "kernels.h", "cavept_ctx.h", "../../include/cavept.h"]

NVCC_FLAGS = [f"-D{d}" for d in os.environ.get("CAVEPT_DEFINES", "").split()] + [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-fmad=false",  # float results follow the source's operation order (see cv_common.h)
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-fvisibility=hidden,-Wall,-Wno-unknown-pragmas",
    "-shared", "-cudart", "static", "-ldl",
    "-Xptxas", "-v",
]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS) or os.path.getmtime(__file__) > t


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = res.stdout + res.stderr
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if res.returncode != 0:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed building libcavept.so")
    if verbose:
        print(log)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
