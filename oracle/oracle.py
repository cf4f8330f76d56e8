"""TEST INFRASTRUCTURE ONLY -- Python access to the CPU oracle.

Two back ends with the same interface:
  * "port": oracle/libcave_oracle.so, our fp32 restatement (cave_oracle.cpp);
  * "ref":  oracle/_ref/*.so, the reference's own sources compiled in place from
            /root/reference (oracle/Makefile `make ref`).  Present only where it was
            built (this container; it travels to the GPU box as a prebuilt file).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product package never does.

The scene assembly below restates, in numpy integer/copy operations,
cave/engine/renderer/path_tracer/path_tracer.cpp:27-47 (ConstructMesh), :72-118
(AppendVertices/AppendIndices/AppendBvhs), :120-201 (UpdateAccelStructure) and
:203-310 (CreateAccelStructure).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

# reference GPU layouts, cave/shader/structured_buffer.hlsl.h:24-61
BVH_DT = np.dtype([("min", "<f4", 3), ("miss", "<i4"), ("max", "<f4", 3), ("hit", "<i4"), ("pad", "<f4", 2), ("leaf", "<i4"), ("tri", "<i4")])
VERTEX_DT = np.dtype([("position", "<f4", 3), ("pad1", "<i4"), ("normal", "<f4", 3), ("pad2", "<i4")])
INDEX_DT = np.dtype([("tri", "<i4", 3), ("pad1", "<i4")])
MESH_DT = np.dtype([("transform", "<f4", 16), ("transform_inv", "<f4", 16), ("pad", "<f4", 2), ("root", "<i4"), ("material", "<i4")])
MATERIAL_DT = np.dtype([("base_color", "<f4", 3), ("roughness", "<f4"), ("emissive", "<f4", 3), ("metallic", "<f4")])
assert BVH_DT.itemsize == 48 and VERTEX_DT.itemsize == 32 and INDEX_DT.itemsize == 16
assert MESH_DT.itemsize == 144 and MATERIAL_DT.itemsize == 32


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("fov_degree", C.c_float), ("forward", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3)]


def _cam(cam) -> Camera:
    return Camera((C.c_float * 3)(*cam.position), cam.fov_degree, (C.c_float * 3)(*cam.forward), (C.c_float * 3)(*cam.right), (C.c_float * 3)(*cam.up))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def build_port(force: bool = False) -> str:
    path = os.path.join(HERE, "libcave_oracle.so")
    src = os.path.join(HERE, "cave_oracle.cpp")
    if force or not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)
    return path


def build_ref() -> bool:
    """Compile the reference's own sources (needs /root/reference). Returns availability."""
    if os.path.isdir("/root/reference/cave"):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)
    return os.path.exists(os.path.join(REF_DIR, "libcave_ref_bvh.so"))


class AssembledScene:
    """Scene-global arrays in the reference's GPU layouts."""

    def __init__(self, vertices, indices, bvhs, meshes, materials, mesh_node_ranges):
        self.vertices, self.indices, self.bvhs, self.meshes, self.materials = vertices, indices, bvhs, meshes, materials
        self.mesh_node_ranges = mesh_node_ranges  # per input mesh id: (first node, node count) or None


class Oracle:
    kind = "?"

    # ---- BVH ---------------------------------------------------------------
    def bvh_build(self, indices: np.ndarray, positions: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def mat4_inverse(self, m: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    # ---- scene assembly (a10-a12) ------------------------------------------
    def assemble(self, scene) -> AssembledScene:
        verts, idxs, nodes = [], [], []
        n_v = n_t = n_b = 0
        mesh_meta = {}  # mesh id -> (rootBvhId, material of the FIRST renderer using it) path_tracer.cpp:227-231
        ranges = [None] * len(scene.meshes)
        for inst in scene.instances:  # CreateAccelStructure: iterate renderers, build each unique mesh once
            if inst.mesh in mesh_meta:
                continue
            mesh = scene.meshes[inst.mesh]
            mesh_meta[inst.mesh] = (n_b, inst.material)
            b = self.bvh_build(mesh.indices, mesh.positions).copy()
            ranges[inst.mesh] = (n_b, b.shape[0])
            # AppendBvhs: links shifted by the node offset only when >= 0; triangleIndex by the triangle offset
            b["hit"] = np.where(b["hit"] >= 0, b["hit"] + n_b, b["hit"])
            b["miss"] = np.where(b["miss"] >= 0, b["miss"] + n_b, b["miss"])
            b["tri"] = np.where(b["tri"] >= 0, b["tri"] + n_t, b["tri"])
            v = np.zeros(mesh.positions.shape[0], dtype=VERTEX_DT)
            v["position"] = mesh.positions
            v["normal"] = mesh.normals
            i = np.zeros(mesh.n_tris, dtype=INDEX_DT)
            i["tri"] = mesh.indices.reshape(-1, 3).astype(np.int32) + n_v  # AppendIndices
            verts.append(v)
            idxs.append(i)
            nodes.append(b)
            n_v += v.shape[0]
            n_t += i.shape[0]
            n_b += b.shape[0]
        # UpdateAccelStructure: materials in std::map<Guid,...> order == mesh id order here, de-duplicated
        lookup, mats = {}, []
        for mesh_id in sorted(mesh_meta):
            mat_id = mesh_meta[mesh_id][1]
            if mat_id in lookup:
                continue
            lookup[mat_id] = len(mats)
            m = scene.materials[mat_id]
            g = np.zeros((), dtype=MATERIAL_DT)
            g["base_color"] = np.asarray(m.base_color, dtype=np.float32)
            g["emissive"] = g["base_color"] * np.float32(m.emissive)
            g["roughness"] = m.roughness
            g["metallic"] = m.metallic
            mats.append(g)
        meshes = np.zeros(len(scene.instances), dtype=MESH_DT)
        for k, inst in enumerate(scene.instances):
            root, mat_id = mesh_meta[inst.mesh]
            meshes[k]["transform"] = inst.transform
            meshes[k]["transform_inv"] = self.mat4_inverse(np.asarray(inst.transform, dtype=np.float32))
            meshes[k]["root"] = root
            meshes[k]["material"] = lookup[mat_id]
        return AssembledScene(np.concatenate(verts), np.concatenate(idxs), np.concatenate(nodes), meshes, np.array(mats, dtype=MATERIAL_DT), ranges)

    # ---- tracer ------------------------------------------------------------
    def render(self, asm: AssembledScene, camera, width, height, first_frame, n_frames, max_bounce, accum: Optional[np.ndarray] = None,
               reset_first=True, rect=None, threads=0):
        raise NotImplementedError


class PortOracle(Oracle):
    kind = "port"

    def __init__(self):
        self.lib = C.CDLL(build_port())
        self.lib.caveoracle_bvh_build.restype = C.c_int
        self.lib.caveoracle_trace_render.restype = C.c_longlong
        self.lib.caveoracle_mat4_inverse.restype = None
        self.lib.caveoracle_tonemap.restype = None

    def bvh_build(self, indices, positions):
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        positions = np.ascontiguousarray(positions, dtype=np.float32)
        n = indices.size // 3
        out = np.zeros(max(2 * n - 1, 1), dtype=BVH_DT)
        rc = self.lib.caveoracle_bvh_build(_ptr(indices), C.c_int(indices.size), _ptr(positions), C.c_int(positions.shape[0]), _ptr(out), C.c_int(out.shape[0]))
        if rc < 0:
            raise ValueError(f"oracle bvh_build failed rc={rc}")
        return out[:rc]

    def mat4_inverse(self, m):
        m = np.ascontiguousarray(m, dtype=np.float32)
        out = np.zeros(16, dtype=np.float32)
        self.lib.caveoracle_mat4_inverse(_ptr(m), _ptr(out))
        return out

    def render(self, asm, camera, width, height, first_frame, n_frames, max_bounce, accum=None, reset_first=True, rect=None, threads=0):
        if accum is None:
            accum = np.zeros((height, width, 4), dtype=np.float32)
        x0, y0, x1, y1 = rect if rect else (0, 0, width, height)
        counters = (C.c_longlong * 4)()
        ms = C.c_double()
        cam = _cam(camera)
        segs = self.lib.caveoracle_trace_render(
            _ptr(asm.vertices), _ptr(asm.indices), _ptr(asm.bvhs), _ptr(asm.meshes), C.c_int(asm.meshes.shape[0]), _ptr(asm.materials),
            C.byref(cam), C.c_int(max_bounce), C.c_int(width), C.c_int(height), C.c_int(x0), C.c_int(y0), C.c_int(x1), C.c_int(y1),
            C.c_int(first_frame), C.c_int(n_frames), C.c_int(1 if reset_first else 0), _ptr(accum), C.c_int(threads or os.cpu_count() or 1),
            counters, C.byref(ms))
        stats = {"segments": int(segs), "node_visits": int(counters[1]), "tri_tests": int(counters[2]), "mesh_iters": int(counters[3]), "ms": ms.value}
        return accum, stats

    def tonemap(self, accum):
        accum = np.ascontiguousarray(accum, dtype=np.float32)
        out = np.zeros_like(accum)
        self.lib.caveoracle_tonemap(_ptr(accum), _ptr(out), C.c_longlong(accum.size // 4))
        return out


class RefOracle(Oracle):
    """The reference's own code (oracle/_ref)."""

    kind = "reference"

    def __init__(self):
        self.bvh = C.CDLL(os.path.join(REF_DIR, "libcave_ref_bvh.so"))
        self.bvh.caveref_bvh_build.restype = C.c_int
        self.bvh.caveref_last_error.restype = C.c_char_p
        self._trace = {}
        self._loaded_scene = {}
        self._port = None
        self.last_build_ms = (0.0, 0.0)

    @staticmethod
    def available() -> bool:
        return os.path.exists(os.path.join(REF_DIR, "libcave_ref_bvh.so"))

    def bvh_build(self, indices, positions):
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        positions = np.ascontiguousarray(positions, dtype=np.float32)
        n = indices.size // 3
        out = np.zeros(max(2 * n - 1, 1), dtype=BVH_DT)
        ms = (C.c_double * 2)()
        rc = self.bvh.caveref_bvh_build(_ptr(indices), C.c_int(indices.size), _ptr(positions), C.c_int(positions.shape[0]), _ptr(out), C.c_int(out.shape[0]), ms)
        if rc < 0:
            raise ValueError(f"reference bvh build trapped rc={rc}: {self.bvh.caveref_last_error().decode()}")
        self.last_build_ms = (ms[0], ms[1])
        return out[:rc]

    def mat4_inverse(self, m):
        m = np.ascontiguousarray(m, dtype=np.float32)
        out = np.zeros(16, dtype=np.float32)
        self.bvh.caveref_mat4_inverse(_ptr(m), _ptr(out))
        return out

    def trace_lib(self, max_bounce):
        if max_bounce not in self._trace:
            p = os.path.join(REF_DIR, f"libcave_ref_trace_mb{max_bounce}.so")
            if not os.path.exists(p):
                raise FileNotFoundError(p)
            lib = C.CDLL(p)
            lib.caveref_trace_render.restype = C.c_longlong
            lib.caveref_trace_load.restype = C.c_int
            self._trace[max_bounce] = lib
        return self._trace[max_bounce]

    def render(self, asm, camera, width, height, first_frame, n_frames, max_bounce, accum=None, reset_first=True, rect=None, threads=0):
        lib = self.trace_lib(max_bounce)
        if self._loaded_scene.get(max_bounce) is not asm:
            rc = lib.caveref_trace_load(_ptr(asm.vertices), C.c_int(asm.vertices.shape[0]), _ptr(asm.indices), C.c_int(asm.indices.shape[0]),
                                        _ptr(asm.bvhs), C.c_int(asm.bvhs.shape[0]), _ptr(asm.meshes), C.c_int(asm.meshes.shape[0]),
                                        _ptr(asm.materials), C.c_int(asm.materials.shape[0]))
            if rc != 0:
                raise ValueError("scene exceeds the reference wrapper's static capacities")
            self._loaded_scene[max_bounce] = asm
        if accum is None:
            accum = np.zeros((height, width, 4), dtype=np.float32)
        x0, y0, x1, y1 = rect if rect else (0, 0, width, height)
        ms = C.c_double()
        cam = _cam(camera)
        segs = lib.caveref_trace_render(C.byref(cam), C.c_int(width), C.c_int(height), C.c_int(x0), C.c_int(y0), C.c_int(x1), C.c_int(y1),
                                        C.c_int(first_frame), C.c_int(n_frames), C.c_int(1 if reset_first else 0), _ptr(accum),
                                        C.c_int(threads or os.cpu_count() or 1), C.byref(ms))
        return accum, {"segments": int(segs), "ms": ms.value}


_port_singleton = None
_ref_singleton = None


def port() -> PortOracle:
    global _port_singleton
    if _port_singleton is None:
        _port_singleton = PortOracle()
    return _port_singleton


def ref() -> Optional[RefOracle]:
    global _ref_singleton
    if _ref_singleton is None and RefOracle.available():
        _ref_singleton = RefOracle()
    return _ref_singleton


def best() -> Oracle:
    """The reference's own code when it was built here, else the port."""
    return ref() or port()
