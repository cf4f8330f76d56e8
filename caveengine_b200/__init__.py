"""caveengine_b200 -- B200-native drop-in for CaveEngine's progressive path tracer pass.

The product is the CUDA shared library `libcavept.so` (C ABI: include/cavept.h), built
in-tree by `caveengine_b200.build`.  This package is the thin host-side mirror of the
reference's renderer-facing surface
(cave/engine/renderer/path_tracer_render_system.h:10-21) over that ABI, plus the
procedural scene generators the tests and the benchmark share.

There is no CPU implementation here: importing works anywhere, but creating a
`PathTracer` raises unless the CUDA library is present and a GPU is visible.
"""
from .api import (  # noqa: F401
    BvhMode,
    CaveptError,
    KernelVariant,
    PathTracer,
    PathTracerMode,
    PinnedArray,
    SceneFile,
    TraceMode,
    library_path,
    load_library,
    write_mesh_file,
)
