import sys, os, json, numpy as np
sys.path.insert(0, os.getcwd())
import caveengine_b200 as cb
from caveengine_b200 import scenes
scene = scenes.animated_scene(1600, 1600)
poses = [scenes.animated_sheet(0.05 * (k + 1), 1600, 1600, with_normals=False).positions for k in range(5)]
with cb.PathTracer(1920, 1080, 8, bvh_mode=cb.BvhMode.LBVH) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    if hasattr(pt, "set_kernel_variant"):
        pt.set_kernel_variant(cb.KernelVariant.PIXEL)
    pt.upload_scene(scene); pt.set_camera(scene.camera); pt.render(0, 1, sync=True)
    ms = []
    for k in range(5):
        pt.rebuild_mesh(0, poses[k]); pt.reset(); pt.reset_stats(); pt.render(k, 1, sync=True)
        ms.append(pt.stats()["last_render_ms"])
    # the same launch again without a rebuild in between (scene data as the previous render left it in L2)
    again = []
    for k in range(3):
        pt.reset(); pt.render(k, 1, sync=True); again.append(pt.stats()["last_render_ms"])
    print(json.dumps({"after_rebuild_ms": ms, "repeat_ms": again}))
