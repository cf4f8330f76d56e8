#!/usr/bin/env python3
"""Small end-to-end pass through the C ABI for compute-sanitizer (memcheck / racecheck): upload (parity + SAH +
LBVH builders), all three kernel variants, a tile launch, an instance refit, the tone passes."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import caveengine_b200 as cb
from caveengine_b200 import scenes

sc = scenes.tiny_scene()
for bvh in (cb.BvhMode.PARITY, cb.BvhMode.LBVH):
    with cb.PathTracer(64, 36, 5, bvh_mode=bvh) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(sc)
        pt.set_camera(sc.camera)
        for v in (cb.KernelVariant.QUEUE, cb.KernelVariant.PIXEL, cb.KernelVariant.POOL):
            pt.set_kernel_variant(v)
            pt.reset()
            pt.render(0, 2, sync=True)
        pt.render_tile(5, 3, 40, 20, 2, 1, sync=True)
        moved = copy.deepcopy(sc)
        moved.instances[1].transform = scenes.mat_trs((1.3, 1.4, 1.2), yaw_deg=12.0)
        pt.update_instances(moved.instances)
        pt.render(0, 1, sync=True)
        img = pt.read_accum()
        pt.read_tonemapped_rgba8()
        pt.read_tonemapped_rgba16f()
        print(bvh.name, "mean", float(img[..., :3].mean()), "refit", pt.stats()["last_update_was_refit"])
with cb.PathTracer(64, 36, 5, trace_mode=cb.TraceMode.REFERENCE_WALK) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(sc)
    pt.set_camera(sc.camera)
    pt.render(0, 1, sync=True)
    print("walk mean", float(pt.read_accum()[..., :3].mean()))
# parity builder with both phases (level loop + one block per item of <= 1024 triangles) on a 10,000-triangle mesh
hf = scenes.heightfield_scene()
with cb.PathTracer(32, 18, 2, bvh_mode=cb.BvhMode.PARITY) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(hf)
    st = pt.stats()
    print("heightfield parity build: levels", st["build_levels"], "launches", st["build_launches"], "nodes", st["node_count"])
    # the radix sort, allocating and static-scratch entry points
    import torch

    rng = np.random.default_rng(1)
    keys = rng.integers(0, 1 << 30, size=20_000, dtype=np.uint64).astype(np.uint32)
    vals = np.arange(keys.size, dtype=np.int32)
    k1, v1 = keys.copy(), vals.copy()
    pt.sort_pairs(k1, v1, 30)
    dk, dv = torch.from_numpy(keys.view(np.int32).copy()).cuda(), torch.from_numpy(vals.copy()).cuda()
    torch.cuda.synchronize()
    pt.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), keys.size, 30)
    pt.sync()
    print("static sort equal", bool(np.array_equal(dv.cpu().numpy(), v1)))
print("SANITIZE_PROBE_DONE")
