#!/usr/bin/env python3
"""Generate build-dir-only shim headers for compiling the reference's hot-path
sources with plain g++ on Linux (test infrastructure; never shipped, never
committed: output goes to oracle/_ref/gen/, which is git-ignored).

Two generated files, each a mechanical rewrite of a reference header that is
read from /root/reference at build time:

 * engine/core/base/typedefs.h  -- the reference refuses Linux/GCC with three
   `#error` lines (cave/engine/core/base/typedefs.h:30-32, :74-76, :85-87).
   They are replaced by PLATFORM_*=NOT_IN_USE, empty WARNING_* macros, and a
   GENERATE_TRAP() that throws a C++ exception so the C wrapper can turn the
   reference's traps (bvh_accel.cpp:117-120,138-141) into an error code.
 * mb<K>/shared_path_tracer.h   -- `#define MAX_BOUNCE 5`
   (cave/shader/shared_path_tracer.h:25) rewritten to K, because
   BASELINE.json's configs ask for 8 bounces and the define is unconditional;
   and every `inout T name` parameter rewritten to `T& name`.  The header's own
   C++ prologue does `#define inout` (empty, :10), which makes the file PARSE as
   C++ but turns every GLSL inout parameter into pass-by-value: the RNG state
   never advances and HitTriangle's `p_ray.t = t` is lost, so the header compiled
   as-is does not compute what the shader computes.  GLSL inout is copy-in /
   copy-out, which (no aliasing here) is exactly a C++ reference.
"""
import os
import re
import sys


def gen_typedefs(ref_root: str, out_root: str) -> None:
    src = open(os.path.join(ref_root, "cave/engine/core/base/typedefs.h")).read()
    lines = src.split("\n")
    out = []
    n_err = 0
    for ln in lines:
        s = ln.strip()
        if s.startswith("#error"):
            n_err += 1
            if n_err == 1:
                out += ["#define PLATFORM_WINDOWS NOT_IN_USE",
                        "#define PLATFORM_APPLE   NOT_IN_USE",
                        "#define PLATFORM_WASM    NOT_IN_USE"]
            elif n_err == 2:
                out += ["#define WARNING_PUSH()",
                        "#define WARNING_POP()",
                        "#define WARNING_DISABLE(a, b)"]
            elif n_err == 3:
                out += ["struct CaveRefTrap {};",
                        "#define GENERATE_TRAP() throw ::CaveRefTrap {}"]
            else:
                raise SystemExit("unexpected extra #error in typedefs.h")
        else:
            out.append(ln)
    if n_err != 3:
        raise SystemExit(f"expected 3 #error lines in typedefs.h, found {n_err}")
    dst = os.path.join(out_root, "engine/core/base")
    os.makedirs(dst, exist_ok=True)
    open(os.path.join(dst, "typedefs.h"), "w").write("\n".join(out))


def gen_tracer_header(ref_root: str, out_root: str, max_bounce: int) -> None:
    src = open(os.path.join(ref_root, "cave/shader/shared_path_tracer.h")).read()
    new, n = re.subn(r"#define\s+MAX_BOUNCE\s+5\b", f"#define MAX_BOUNCE    {max_bounce}", src)
    if n != 1:
        raise SystemExit("MAX_BOUNCE define not found exactly once")
    new, n = re.subn(r"\binout\s+(\w+)\s+(\w+)", r"\1& \2", new)
    if n != 8:
        raise SystemExit(f"expected 8 inout parameters in shared_path_tracer.h, rewrote {n}")
    dst = os.path.join(out_root, f"mb{max_bounce}")
    os.makedirs(dst, exist_ok=True)
    open(os.path.join(dst, "shared_path_tracer.h"), "w").write(new)


if __name__ == "__main__":
    ref_root, out_root = sys.argv[1], sys.argv[2]
    gen_typedefs(ref_root, out_root)
    for k in sys.argv[3:]:
        gen_tracer_header(ref_root, out_root, int(k))
