#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE'S OWN CODE (oracle/_ref, i.e.
/root/reference's bvh_accel.cpp + box.cpp + shared_path_tracer.h compiled in place).
Run in the build container (needs /root/reference):  python tests/golden/make_golden.py

The reference ships no tests or golden vectors for this path, so
these files are the pins: each stores the inputs next to the reference's outputs, so the
checks do not depend on regenerating identical inputs on another host.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from caveengine_b200 import scenes  # noqa: E402
from oracle import oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def signed_zero_mesh():
    """Vertices at +0/-0 in every axis order: pins the a<b?a:b tie rule (common.h:50-58)."""
    pz, nz = np.float32(0.0), np.float32(-0.0)
    pos = np.array([[nz, pz, 1], [pz, nz, 2], [1, pz, nz], [nz, 1, pz], [2, nz, pz], [pz, 2, nz], [nz, nz, nz], [1, 1, pz], [-1, nz, pz], [pz, -1, nz],
                    [nz, pz, -1], [3, nz, 1]], dtype=np.float32)
    idx = np.array([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 1, 0, 6, 2, 6, 9, 5, 3, 8, 4, 7, 10, 11, 6, 0, 6, 1, 3, 9, 2, 11], dtype=np.uint32)
    return scenes.Mesh(pos, np.tile(np.array([0, 1, 0], dtype=np.float32), (pos.shape[0], 1)), idx)


def bvh_cases():
    cases = {
        "heightfield_8x8": scenes.heightfield_mesh(8, 8, 7, cell=0.5, height=0.6),
        "heightfield_13x7": scenes.heightfield_mesh(13, 7, 99),
        "sphere_12x8": scenes._sphere(12, 8, 0.5, 8, bump=0.1),
        "open_box": scenes._open_box(0.8, 0.8, 0.8),
        "quad": scenes._quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1)),
        "single_triangle": scenes.Mesh(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32), np.tile(np.array([0, 0, 1], dtype=np.float32), (3, 1)),
                                       np.array([0, 1, 2], dtype=np.uint32)),
        "signed_zero": signed_zero_mesh(),
        "cylinder_16x6": scenes._cylinder(16, 6, 0.4, 0.0, 2.0, 3),
        "drape_10x10": scenes._drape(10, 10, 4.5, 5.0, 95),
    }
    return cases


def main():
    ref = O.ref()
    if ref is None:
        raise SystemExit("oracle/_ref is not built; run `make -C oracle ref` first")
    # ---- BVH known answers ------------------------------------------------------------
    blob = {}
    for name, mesh in bvh_cases().items():
        nodes = ref.bvh_build(mesh.indices, mesh.positions)
        blob[name + "/positions"] = mesh.positions
        blob[name + "/indices"] = mesh.indices
        blob[name + "/nodes"] = nodes.view(np.uint8).reshape(-1, 48)
    np.savez_compressed(os.path.join(OUT, "bvh_golden.npz"), **blob)

    # digests of large builds (inputs are regenerated; the input digest guards that)
    big = {}
    for name, mesh in {"heightfield_10000": scenes.heightfield_mesh(), "sponza_mesh0_floor": scenes.sponza_scene().meshes[0], "sponza_mesh22_blob": scenes.sponza_scene().meshes[22]}.items():
        nodes = ref.bvh_build(mesh.indices, mesh.positions)
        big[name + "/input_sha256"] = np.frombuffer(hashlib.sha256(mesh.positions.tobytes() + mesh.indices.tobytes()).digest(), dtype=np.uint8)
        big[name + "/nodes_sha256"] = np.frombuffer(hashlib.sha256(nodes.tobytes()).digest(), dtype=np.uint8)
        big[name + "/n_nodes"] = np.array([nodes.shape[0]])
    np.savez_compressed(os.path.join(OUT, "bvh_digests.npz"), **big)

    # ---- matrix inverse (glm::inverse) -------------------------------------------------
    mats = np.stack([scenes.mat_identity(), scenes.mat_trs((1, 2, 3), 30.0, 0.7), scenes.mat_trs((-0.7, 0.0, 0.6), 18.0), scenes.mat_trs((0, 5, 0), -77.0, (1.0, 0.55, 2.0))])
    inv = np.stack([ref.mat4_inverse(m) for m in mats])
    np.savez_compressed(os.path.join(OUT, "inverse_golden.npz"), mats=mats, inverses=inv)

    # ---- images -----------------------------------------------------------------------
    blob = {}
    for name, scene, (w, h), mb, frames in [("tiny_mb5", scenes.tiny_scene(), (48, 27), 5, (0, 4)), ("cornell_mb5", scenes.cornell_scene(), (48, 27), 5, (3, 3)),
                                            ("tiny_mb8", scenes.tiny_scene(), (32, 18), 8, (10, 2)), ("tiny_mb1", scenes.tiny_scene(), (32, 18), 1, (0, 1))]:
        asm = ref.assemble(scene)
        img, st = ref.render(asm, scene.camera, w, h, frames[0], frames[1], mb, threads=1)
        for k in ("vertices", "indices", "bvhs", "meshes", "materials"):
            blob[f"{name}/{k}"] = getattr(asm, k).view(np.uint8)
        c = scene.camera
        blob[f"{name}/camera"] = np.array(list(c.position) + [c.fov_degree] + list(c.forward) + list(c.right) + list(c.up), dtype=np.float32)
        blob[f"{name}/params"] = np.array([w, h, mb, frames[0], frames[1], st["segments"]], dtype=np.int64)
        blob[f"{name}/image"] = img
    np.savez_compressed(os.path.join(OUT, "image_golden.npz"), **blob)
    for f in ("bvh_golden.npz", "bvh_digests.npz", "inverse_golden.npz", "image_golden.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)), "bytes")


if __name__ == "__main__":
    main()
