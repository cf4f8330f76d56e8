#!/usr/bin/env python3
"""Generates tests/golden/isect_golden.npz from the reference's own CPU twins of the shader's
intersection routines: TestIntersection::RayTriangle / RayAabb,
cave/engine/math/intersection.cpp:18-70, compiled in place (oracle/_ref/libcave_ref_isect.so).
Needs /root/reference (build container).  Stores the inputs next to the reference's outputs."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libcave_ref_isect.so"))
    rng = np.random.default_rng(8020)
    n = 4096
    start = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    end = (start + rng.normal(size=(n, 3)) * rng.uniform(0.5, 6.0, (n, 1))).astype(np.float32)
    # triangles near the segment so a good share hits; every 8th ray axis-aligned (a zero direction
    # component: the slab test then runs on +-inf, as the shader relies on, shared_path_tracer.h:137)
    end[::8, 1] = start[::8, 1]
    end[::16, 2] = start[::16, 2]
    mid = start + (end - start) * rng.uniform(0.05, 1.3, (n, 1)).astype(np.float32)
    tri = (mid[:, None, :] + rng.normal(size=(n, 3, 3)) * rng.uniform(0.05, 1.5, (n, 1, 1))).astype(np.float32)
    tri[::5] = tri[::5, ::-1]  # flip some windings (back faces are culled)
    lo = (mid + rng.uniform(-1.5, 0.2, (n, 3))).astype(np.float32)
    hi = (lo + rng.uniform(0.0, 2.5, (n, 3))).astype(np.float32)
    hi[::7, 0] = lo[::7, 0]  # flat boxes
    rays6 = np.ascontiguousarray(np.concatenate([start, end], axis=1), dtype=np.float32)
    tris9 = np.ascontiguousarray(tri.reshape(n, 9))
    boxes6 = np.ascontiguousarray(np.concatenate([lo, hi], axis=1), dtype=np.float32)
    out_t = np.zeros((n, 2), dtype=np.float32)
    out_b = np.zeros((n, 2), dtype=np.float32)
    for f in (lib.cave_ref_ray_triangle, lib.cave_ref_ray_aabb):
        f.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.cave_ref_ray_triangle(n, rays6.ctypes.data, tris9.ctypes.data, out_t.ctypes.data)
    lib.cave_ref_ray_aabb(n, rays6.ctypes.data, boxes6.ctypes.data, out_b.ctypes.data)
    np.savez_compressed(os.path.join(OUT, "isect_golden.npz"), rays=rays6, tris=tris9, boxes=boxes6, ref_triangle=out_t, ref_aabb=out_b)
    print("triangle hits", int(out_t[:, 0].sum()), "box hits", int(out_b[:, 0].sum()), "of", n)


if __name__ == "__main__":
    main()
