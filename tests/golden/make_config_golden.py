#!/usr/bin/env python3
"""Golden vectors for the scenes of BASELINE configs 3, 4 and 5, made by the REFERENCE'S OWN CODE
(oracle/_ref: bvh_accel.cpp + box.cpp + shared_path_tracer.h compiled in place from /root/reference).
Run in the build container:  python tests/golden/make_config_golden.py   -> tests/golden/config_golden.npz

  * per mesh of breakfast_1048576 (config 4) and of animated_500000 (config 5 at test size): SHA-256 of the
    inputs and of the reference builder's GpuPtBvh array;
  * small images of the three scenes rendered by the reference's shader code at equal seeds
    (stored next to the parameters; the inputs are regenerated and guarded by the input digests).
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
IMAGES = [("sponza", 8, 64, 36, 0, 2), ("breakfast", 8, 48, 27, 0, 2), ("animated", 5, 48, 27, 0, 2)]


def scene_of(name):
    return {"sponza": scenes.sponza_scene, "breakfast": scenes.breakfast_scene, "animated": lambda: scenes.animated_scene(500, 500)}[name]()


def sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return np.frombuffer(h.digest(), dtype=np.uint8)


def main():
    ref = O.ref()
    if ref is None:
        raise SystemExit("oracle/_ref is not built; run `make -C oracle ref` first")
    blob = {}
    for name, mb, w, h, f0, nf in IMAGES:
        sc = scene_of(name)
        used = sorted({i.mesh for i in sc.instances})
        for m in used:
            mesh = sc.meshes[m]
            nodes = ref.bvh_build(mesh.indices, mesh.positions)
            blob[f"{name}/mesh{m}/input_sha256"] = sha(mesh.positions, mesh.indices)
            blob[f"{name}/mesh{m}/nodes_sha256"] = sha(nodes)
            blob[f"{name}/mesh{m}/n_nodes"] = np.array([nodes.shape[0]])
        asm = ref.assemble(sc)
        img, st = ref.render(asm, sc.camera, w, h, f0, nf, mb)
        blob[f"{name}/params"] = np.array([w, h, mb, f0, nf, st["segments"]], dtype=np.int64)
        blob[f"{name}/image"] = img
        blob[f"{name}/scene_sha256"] = sha(asm.vertices, asm.indices, asm.meshes, asm.materials)
        print(name, "meshes", len(used), "segments", st["segments"], "mean", float(img[..., :3].mean()), flush=True)
    p = os.path.join(OUT, "config_golden.npz")
    np.savez_compressed(p, **blob)
    print(p, os.path.getsize(p), "bytes")


if __name__ == "__main__":
    main()
