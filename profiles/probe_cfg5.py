import sys, os
sys.path.insert(0, "/root/repo")
import caveengine_b200 as cb
from caveengine_b200 import scenes
scene = scenes.animated_scene(1600, 1600)
pose = scenes.animated_sheet(0.05, 1600, 1600, with_normals=False).positions
with cb.PathTracer(1920, 1080, 8, bvh_mode=cb.BvhMode.LBVH) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(scene)
    pt.set_camera(scene.camera)
    pt.rebuild_mesh(0, pose)
    print(pt.stats()["last_fast_build_ms"])
