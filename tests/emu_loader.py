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
FASTTRI_DT = np.dtype([("v0", "<f4", 4), ("v1", "<f4", 4), ("v2", "<f4", 4)])


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


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class EmuScene:
    """Runs the product's build pipeline (parity BVH -> compact layout -> TLAS) in emulation
    on the scene-global arrays of an oracle-assembled scene."""

    def __init__(self, asm, scene):
        from oracle import oracle as O

        self.lib = load()
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
        n_inst = len(scene.instances)
        self.fnodes = np.zeros(self.n_nodes + n_meshes + 2 * n_inst + 8, dtype=FASTNODE_DT)
        self.ftris = np.zeros(n_tris, dtype=FASTTRI_DT)
        self.fnormals = np.zeros(n_tris, dtype=FASTTRI_DT)
        self.mesh_root_ref = np.zeros(n_meshes, dtype=np.int32)
        self.n_pairs = self.lib.emu_compact(_p(self.nodes), _p(self.parent), _p(self.count), C.c_int(self.n_nodes), _p(asm.vertices), _p(asm.indices),
                                            _p(self.fnodes), _p(self.ftris), _p(self.fnormals), _p(self.mesh_root_ref))
        mesh_of_inst = np.array([built.index(i.mesh) for i in scene.instances], dtype=np.int32)
        root_node = (2 * self.tri_off[:-1] - np.arange(n_meshes)).astype(np.int32)
        self.tlas_base = 2 * self.n_pairs
        root_ref = C.c_int()
        self.n_tlas = self.lib.emu_tlas(_p(self.nodes), _p(asm.meshes), C.c_int(n_inst), _p(mesh_of_inst), _p(root_node), C.c_int(self.tlas_base), _p(self.fnodes),
                                        C.byref(root_ref))
        self.tlas_root = root_ref.value
        self.inst_fast_root = self.mesh_root_ref[mesh_of_inst].astype(np.int32)

    def render(self, camera, width, height, first_frame, n_frames, max_bounce, fast, accum=None, reset_first=True):
        from oracle.oracle import _cam

        if accum is None:
            accum = np.zeros((height, width, 4), dtype=np.float32)
        cam = _cam(camera)
        counters = (C.c_ulonglong * 3)()
        a = self.asm
        segs = self.lib.emu_render(_p(a.vertices), _p(a.indices), _p(self.nodes), _p(a.meshes), C.c_int(a.meshes.shape[0]), _p(a.materials), _p(self.fnodes),
                                   _p(self.ftris), _p(self.fnormals), _p(self.inst_fast_root), C.c_int(self.tlas_root), C.byref(cam), C.c_int(max_bounce), C.c_int(width),
                                   C.c_int(height), C.c_int(first_frame), C.c_int(n_frames), C.c_int(1 if reset_first else 0), _p(accum), C.c_int(1 if fast else 0),
                                   counters)
        return accum, {"segments": int(segs), "node_visits": int(counters[1]), "tri_tests": int(counters[2])}
