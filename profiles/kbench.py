#!/usr/bin/env python3
"""Quick kernel benchmark (measurement helper, not the contract bench): config 3's scene and image, `--spp` frames
per launch, device time from the library's CUDA events.  One JSON line per (library, kernel variant, spp).

    python profiles/kbench.py [--libs variants/lib_0.so ...] [--variants pool pixel] [--spp 32 256] [--reps 3]
Each library is measured in its own process (CAVEPT_LIBRARY)."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(args):
    import caveengine_b200 as cb
    from caveengine_b200 import scenes

    sc = {"sponza": scenes.sponza_scene, "breakfast": scenes.breakfast_scene, "cornell": scenes.cornell_scene}[args.scene]()
    w, h = (3840, 2160) if args.scene == "breakfast" and not args.size else (args.size or (1920, 1080))
    mb = 5 if args.scene == "cornell" else 8
    tag = os.environ.get("KBENCH_TAG", "")
    with cb.PathTracer(w, h, mb, bvh_mode=cb.BvhMode.FAST) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(sc)
        pt.set_camera(sc.camera)
        for variant in args.variants:
            pt.set_kernel_variant(getattr(cb.KernelVariant, variant.upper()))
            pt.set_count_visits(True)
            pt.reset_stats()
            pt.render(1 << 20, 1, sync=True)
            c = pt.stats()
            pt.set_count_visits(False)
            for spp in args.spp:
                pt.reset()
                pt.render(0, spp, sync=True)  # warm
                ms = []
                segs = 0
                for r in range(args.reps):
                    pt.reset()
                    pt.reset_stats()
                    pt.render(r * spp, spp, sync=True)
                    st = pt.stats()
                    ms.append(st["last_render_ms"])
                    segs = st["segments"]
                best = min(ms)
                print(json.dumps({"lib": tag, "scene": args.scene, "variant": variant, "spp": spp, "M_seg_per_s": segs / best / 1e3, "ms_best": best, "ms_all": ms,
                                  "nodes_per_seg": c["node_visits"] / c["segments"], "tris_per_seg": c["triangle_tests"] / c["segments"]}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--libs", nargs="*", default=[])
    ap.add_argument("--variants", nargs="*", default=["queue", "pixel"])
    ap.add_argument("--spp", nargs="*", type=int, default=[32])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--scene", default="sponza")
    ap.add_argument("--size", nargs=2, type=int, default=None)
    ap.add_argument("--child", action="store_true")
    args = ap.parse_args()
    if args.size:
        args.size = tuple(args.size)
    if args.child or not args.libs:
        return one(args)
    for lib in args.libs:
        tag = lib
        txt = os.path.splitext(lib)[0] + ".txt"
        if os.path.exists(txt):
            tag = open(txt).read().strip() or lib
        env = dict(os.environ, CAVEPT_LIBRARY=os.path.abspath(lib), KBENCH_TAG=tag)
        cmd = [sys.executable, os.path.abspath(__file__), "--child", "--variants", *args.variants, "--spp", *[str(s) for s in args.spp], "--reps", str(args.reps), "--scene", args.scene]
        if args.size:
            cmd += ["--size", str(args.size[0]), str(args.size[1])]
        try:
            subprocess.run(cmd, env=env, timeout=240, check=False)
        except subprocess.TimeoutExpired:
            print(json.dumps({"lib": tag, "error": "timeout"}), flush=True)


if __name__ == "__main__":
    main()
