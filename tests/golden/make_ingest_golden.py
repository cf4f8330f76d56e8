#!/usr/bin/env python3
"""Generates tests/golden/ingest_golden.npz (needs /root/reference; run in the build container):

  * `suzanne/*`: counts and SHA-256 digests of the arrays in the reference's own asset
    projects/sample3d/resources/models/monkey/Suzanne.mesh, read with an independent numpy
    parser of MeshAsset::SerializeBinary's layout (cave/engine/assets/mesh_asset.cpp:45-61),
    and the monkey.scene instance/material the engine would hand the path tracer;
  * `chains/*`: seeded TRS chains (entity, parent, grandparent, ...) and the world matrices the
    REFERENCE'S OWN TransformComponent::GetLocalMatrix + the UpdateHierarchy product give for
    them (oracle/_ref/libcave_ref_scene.so = cave/engine/scene/transform_component.cpp compiled
    in place over the vendored glm).
"""
import ctypes as C
import hashlib
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def parse_mesh(path):
    b = open(path, "rb").read()
    o = 0
    (subset_count,) = struct.unpack_from("<Q", b, o); o += 8
    (n,) = struct.unpack_from("<Q", b, o); o += 8
    assert n == subset_count
    o += 32 * n
    o += 4  # flags
    out = {}
    for name, dt, width in [("indices", "<u4", 1), ("positions", "<f4", 3), ("normals", "<f4", 3), ("tangents", "<f4", 3), ("uv0", "<f4", 2)]:
        (n,) = struct.unpack_from("<Q", b, o); o += 8
        out[name] = np.frombuffer(b, dt, n * width, o).copy()
        o += n * width * 4
    return out


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def main():
    g = {}
    m = parse_mesh(f"{REF}/projects/sample3d/resources/models/monkey/Suzanne.mesh")
    g["suzanne/counts"] = np.array([m["positions"].size // 3, m["indices"].size], dtype=np.int64)
    for k in ("indices", "positions", "normals"):
        g[f"suzanne/sha_{k}"] = sha(m[k])
    g["suzanne/material"] = np.array([0.600000024, 0.600000024, 0.600000024, 1.0, 0.0, 0.0], dtype=np.float32)  # base_color, roughness, metallic, emissive

    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libcave_ref_scene.so"))
    lib.cave_ref_world_matrix.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(20261104)
    chains, worlds, lens = [], [], []
    for case in range(48):
        depth = 1 + case % 4
        chain = np.zeros((depth, 10), dtype=np.float32)
        for k in range(depth):
            q = rng.normal(size=4)
            q /= np.linalg.norm(q)
            chain[k, 0:3] = rng.uniform(-5, 5, 3)
            chain[k, 3:7] = q
            chain[k, 7:10] = rng.uniform(0.2, 3.0, 3) * (1 if case % 5 else 0.008)
        if case == 0:
            chain[0] = [0, 0, 0, 0, 0, 0, 1, 1, 1, 1]
        if case == 1:
            chain[0] = [-0.0, 0, 0, 0, 0, 0, 1, -1, 1, 1]  # signed zeros and a mirror
        out = np.zeros(16, dtype=np.float32)
        lib.cave_ref_world_matrix(depth, chain.ctypes.data, out.ctypes.data)
        padded = np.zeros((4, 10), dtype=np.float32)
        padded[:depth] = chain
        chains.append(padded)
        worlds.append(out)
        lens.append(depth)
    g["chains/trs"] = np.stack(chains)
    g["chains/len"] = np.array(lens, dtype=np.int32)
    g["chains/world"] = np.stack(worlds)
    np.savez_compressed(os.path.join(OUT, "ingest_golden.npz"), **g)
    print("wrote ingest_golden.npz:", {k: v.shape for k, v in g.items()})


if __name__ == "__main__":
    main()
