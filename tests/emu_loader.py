"""TEST CODE -- builds and wraps tests/emu/libcavept_emu.so, the serial host emulation
of the cavept kernel bodies (see tests/emu/emu.cpp)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "caveengine_b200", "csrc")
LIB = os.path.join(HERE, "emu", "libcavept_emu.so")
SRC = os.path.join(HERE, "emu", "emu.cpp")

FASTNODE_DT = np.dtype([("mn", "<f4", 3), ("a", "<i4"), ("mx", "<f4", 3), ("b", "<i4")])
FASTTRI_DT = np.dtype([("v0", "<f4", 4), ("v1", "<f4", 4), ("v2", "<f4", 4), ("pad", "<f4", 4)])


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")]
    return any(os.path.getmtime(d) > t for d in deps)


def load() -> C.CDLL:
    if _stale():
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-fvisibility=hidden", "-I", CSRC, "-o", LIB, SRC])
    lib = C.CDLL(LIB)
    lib.emu_render.restype = C.c_longlong
    return lib


WARP_LIB = os.path.join(HERE, "emu", "libcavept_emu_warp.so")


def load_warp() -> C.CDLL:
    """The same bodies with warps of 32 host threads (-DCV_WARP_EMU): votes and the job draw go through a barrier."""
    deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")]
    if not os.path.exists(WARP_LIB) or any(os.path.getmtime(d) > os.path.getmtime(WARP_LIB) for d in deps):
        subprocess.check_call(["g++", "-std=c++20", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-fvisibility=hidden", "-pthread", "-DCV_WARP_EMU", "-I", CSRC,
                               "-o", WARP_LIB, SRC])
    lib = C.CDLL(WARP_LIB)
    lib.emu_render_warps.restype = C.c_longlong
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class EmuScene:
    """Runs the product's build pipeline (parity BVH + production world-space BVH) in emulation
    on the scene-global arrays of an oracle-assembled scene."""

    def __init__(self, asm, scene, parity_sub_max=None):
        """parity_sub_max: largest item the parity builder hands to its subtree phase (None = the product's constant,
        0 = level loop only)."""
        from oracle import oracle as O

        self.lib = load()
        self.lib.emu_set_parity_sub_max(C.c_int(-1 if parity_sub_max is None else int(parity_sub_max)))
        self.asm = asm
        built = []  # input mesh ids in first-appearance order
        for inst in scene.instances:
            if inst.mesh not in built:
                built.append(inst.mesh)
        tri_counts = [scene.meshes[m].n_tris for m in built]
        self.tri_off = np.concatenate([[0], np.cumsum(tri_counts)]).astype(np.int32)
        n_tris = int(self.tri_off[-1])
        n_meshes = len(built)
        self.n_nodes = 2 * n_tris - n_meshes
        self.nodes = np.zeros(self.n_nodes, dtype=O.BVH_DT)
        self.parent = np.zeros(self.n_nodes, dtype=np.int32)
        self.count = np.zeros(self.n_nodes, dtype=np.int32)
        levels, err = C.c_int(), C.c_int()
        self.rc = self.lib.emu_bvh_build(_p(asm.vertices), _p(asm.indices), C.c_int(n_tris), _p(self.tri_off), C.c_int(n_meshes), _p(self.nodes),
                                         _p(self.parent), _p(self.count), C.byref(levels), C.byref(err))
        self.levels, self.error_bits = levels.value, err.value
        if self.rc != 0:
            return
        self.build_world(asm, scene, built)

    def build_world(self, asm, scene, built, force_median=False, lbvh=False, hybrid=False):
        """Production (world-space) BVH, as cavept_api.cu's build_world() drives it."""
        n_inst = len(scene.instances)
        tri_counts = [scene.meshes[i.mesh].n_tris for i in scene.instances]
        self.inst_prim_off = np.concatenate([[0], np.cumsum(tri_counts)]).astype(np.int32)
        self.inst_tri_off = np.array([int(self.tri_off[built.index(i.mesh)]) for i in scene.instances], dtype=np.int32)
        P = int(self.inst_prim_off[-1])
        self.n_prims = P
        self.fnodes = np.zeros(2 * (P + 8), dtype=FASTNODE_DT)
        self.qpairs = np.zeros((P + 8, 8), dtype=np.uint32)  # QPair: q[6], r0, r1
        self.q4nodes = np.zeros((P + 8, 16), dtype=np.uint32)  # Q4Node: q[12], r[4]
        self.lib.emu_bind_q4(_p(self.q4nodes))
        self.pair_depth = np.zeros(P + 8, dtype=np.uint8)    # depth of the interior node owning pair k (the refit's schedule)
        self.lib.emu_bind_pair_depth(_p(self.pair_depth))
        self.grid = np.zeros(6, dtype=np.float32)            # origin xyz, scale xyz
        self.ftris = np.zeros(P, dtype=FASTTRI_DT)
        self.fnormals = np.zeros(P, dtype=FASTTRI_DT)
        report = np.zeros(4 + 124, dtype=np.int32)
        pad = C.c_float()
        self.wrc = self.lib.emu_wbvh_build(_p(asm.vertices), _p(asm.indices), _p(asm.meshes), C.c_int(n_inst), _p(self.inst_prim_off), _p(self.inst_tri_off),
                                           C.c_int(4 if hybrid else ((3 if force_median else 2) if lbvh else (1 if force_median else 0))), _p(self.fnodes), _p(self.qpairs), _p(self.ftris), _p(self.fnormals), _p(report), C.byref(pad),
                                           _p(self.grid))
        self.fast_levels, self.n_pairs, self.root_ref, self.wb_error_bits = [int(v) for v in report[:4]]
        self.level_items = report[4:4 + self.fast_levels].copy()
        self.pad = pad.value

    def refit(self, meshes, moved):
        """wbvh_body.h refit_wbvh on the tree of build_world(): `meshes` = the GpuPtMesh array with the new transforms,
        `moved` = one flag per instance.  Returns (scene box min/max, surface-area sum)."""
        moved = np.ascontiguousarray(moved, dtype=np.uint8)
        out = np.zeros(7, dtype=np.float32)
        self.lib.emu_bind_q4(_p(self.q4nodes))
        self.lib.emu_bind_pair_depth(_p(self.pair_depth))
        self.lib.emu_wbvh_refit(_p(self.asm.vertices), _p(self.asm.indices), _p(meshes), C.c_int(meshes.shape[0]), C.c_int(self.n_prims), C.c_float(self.pad), _p(self.grid),
                                C.c_int(self.n_pairs), C.c_int(self.fast_levels), _p(moved), _p(self.fnodes), _p(self.qpairs), _p(self.ftris), _p(self.fnormals), _p(out))
        return out[:6].copy(), float(out[6])

    def render(self, camera, width, height, first_frame, n_frames, max_bounce, fast, accum=None, reset_first=True):
        from oracle.oracle import _cam

        if accum is None:
            accum = np.zeros((height, width, 4), dtype=np.float32)
        cam = _cam(camera)
        counters = (C.c_ulonglong * 3)()
        a = self.asm
        self.lib.emu_bind_q4(_p(self.q4nodes))
        segs = self.lib.emu_render(_p(a.vertices), _p(a.indices), _p(self.nodes), _p(a.meshes), C.c_int(a.meshes.shape[0]), _p(a.materials), _p(self.fnodes),
                                   _p(self.qpairs), _p(self.grid), _p(self.ftris), _p(self.fnormals), C.c_int(self.root_ref), C.byref(cam), C.c_int(max_bounce), C.c_int(width),
                                   C.c_int(height), C.c_int(first_frame), C.c_int(n_frames), C.c_int(1 if reset_first else 0), _p(accum), C.c_int(int(fast)),
                                   counters)
        return accum, {"segments": int(segs), "node_visits": int(counters[1]), "tri_tests": int(counters[2])}

    def bind_textures(self, uvs=None, textures=(), base_maps=None, mr_maps=None, env=None):
        """Inputs of the textured instantiation (render(..., fast=6)): per scene-global-vertex uvs (V, 2), RGBA float
        textures [(H, W, 4)], per-material map indices, equirect env (H, W, 3).  The per-slot uv records are emitted here
        the way kernels.cu FnEmitUvs does."""
        a = self.asm
        self._tex_keep = []
        fuvs = None
        if uvs is not None:
            uvs = np.ascontiguousarray(uvs, dtype=np.float32)
            tri = self.ftris["v1"][:, 3].view(np.uint32).astype(np.int64)
            inst = self.ftris["v0"][:, 3].view(np.uint32).astype(np.int64)
            m = a.meshes["transform"].reshape(-1, 16).astype(np.float32)
            det = (m[:, 0] * (m[:, 5] * m[:, 10] - m[:, 9] * m[:, 6]) - m[:, 4] * (m[:, 1] * m[:, 10] - m[:, 9] * m[:, 2]) + m[:, 8] * (m[:, 1] * m[:, 6] - m[:, 5] * m[:, 2]))
            idx = a.indices["tri"][tri].astype(np.int64)
            swap = det[inst] < 0
            i1 = np.where(swap, idx[:, 2], idx[:, 1])
            i2 = np.where(swap, idx[:, 1], idx[:, 2])
            fuvs = np.zeros((self.n_prims, 8), dtype=np.float32)
            fuvs[:, 0:2] = uvs[idx[:, 0]]
            fuvs[:, 2:4] = uvs[i1]
            fuvs[:, 4:6] = uvs[i2]
        table = np.zeros((max(len(textures), 1), 4), dtype=np.int32)
        texels = []
        off = 0
        for k, t in enumerate(textures):
            t = np.ascontiguousarray(t, dtype=np.float32)
            table[k] = (off, t.shape[1], t.shape[0], 0)
            texels.append(t.reshape(-1, 4))
            off += t.shape[0] * t.shape[1]
        texels = np.ascontiguousarray(np.concatenate(texels)) if texels else np.zeros((1, 4), dtype=np.float32)
        maps = None
        if base_maps is not None:
            maps = np.ascontiguousarray(np.stack([np.asarray(base_maps, dtype=np.int32), np.asarray(mr_maps, dtype=np.int32)], axis=1))
        envp, ew, eh = None, 0, 0
        if env is not None:
            env = np.asarray(env, dtype=np.float32)
            eh, ew = env.shape[:2]
            envp = np.ascontiguousarray(np.concatenate([env, np.ones((eh, ew, 1), dtype=np.float32)], axis=2))
        self._tex_keep = [fuvs, table, texels, maps, envp]
        self.lib.emu_bind_textures(None if fuvs is None or maps is None else _p(fuvs), _p(table), _p(texels), None if maps is None or fuvs is None else _p(maps), None if envp is None else _p(envp), C.c_int(ew), C.c_int(eh))

    def primary_hits(self, camera, width, height, frame=0, mode=1):
        """(FastTri slot or -1, ray.t) of every pixel's camera ray: mode 0 = all triangles tested (no tree), 1 = the
        production walk."""
        from oracle.oracle import _cam

        cam = _cam(camera)
        slot = np.zeros((height, width), dtype=np.int32)
        t = np.zeros((height, width), dtype=np.float32)
        self.lib.emu_primary_hits(_p(self.fnodes), _p(self.qpairs), _p(self.grid), _p(self.ftris), C.c_int(self.n_prims), C.c_int(self.root_ref), C.byref(cam),
                                  C.c_int(width), C.c_int(height), C.c_int(frame), C.c_int(mode), _p(slot), _p(t))
        return slot, t

    def render_warps(self, camera, width, height, first_frame, n_frames, max_bounce, n_warps=2, tile=None, accum=None, reset_first=True):
        """k_trace_queue on n_warps concurrent warps of 32 host threads each (real divergence between lanes)."""
        from oracle.oracle import _cam

        lib = load_warp()
        if accum is None:
            accum = np.zeros((height, width, 4), dtype=np.float32)
        cam = _cam(camera)
        counters = (C.c_ulonglong * 3)()
        t = (C.c_int * 4)(*tile) if tile is not None else None
        a = self.asm
        segs = lib.emu_render_warps(_p(a.vertices), _p(a.indices), _p(self.nodes), _p(a.meshes), C.c_int(a.meshes.shape[0]), _p(a.materials), _p(self.fnodes),
                                    _p(self.qpairs), _p(self.grid), _p(self.ftris), _p(self.fnormals), C.c_int(self.root_ref), C.byref(cam), C.c_int(max_bounce),
                                    C.c_int(width), C.c_int(height), C.c_int(first_frame), C.c_int(n_frames), C.c_int(1 if reset_first else 0), _p(accum), C.c_int(n_warps), t,
                                    counters)
        return accum, {"segments": int(segs), "node_visits": int(counters[1]), "tri_tests": int(counters[2])}
