"""ctypes binding of include/cavept.h and a Python mirror of the reference facade.

Reference surface mirrored (cave/engine/renderer/path_tracer_render_system.h:10-21):
    enum class PathTracerMode { NONE, INTERACTIVE, TILED }
    SetPathTracerMode(mode)            -> PathTracer.set_mode
    IsPathTracerActive()               -> PathTracer.is_active
    RequestPathTracerUpdate(cam,scene) -> PathTracer.request_update(scene)
    Bind/UnbindPathTracerData + PathTracerPassFunc's Dispatch
        (cave/engine/render_graph/common_passes.cpp:977-999) -> PathTracer.render
"""
from __future__ import annotations

import ctypes as C
import enum
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("CAVEPT_LIBRARY") or os.path.join(_HERE, "libcavept.so")  # CAVEPT_LIBRARY: another build of the same library (kernel A/B runs)
_lib: Optional[C.CDLL] = None


class CaveptError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"cavept error {code}: {message}")
        self.code = code


class PathTracerMode(enum.IntEnum):  # path_tracer_render_system.h:10-14
    NONE = 0
    INTERACTIVE = 1
    TILED = 2


class TraceMode(enum.IntEnum):
    FAST = 0
    REFERENCE_WALK = 1


class KernelVariant(enum.IntEnum):
    POOL = 0  # persistent warps + job queue + two pixel streams per lane (experiment; slower)
    PIXEL = 1  # one thread per pixel for the whole launch (default: measured fastest of the pair-walk variants)
    QUEUE = 2  # persistent warps drawing pixel jobs from a global queue
    WIDE4 = 3  # PIXEL's launch shape over four-wide node records


class BvhMode(enum.IntEnum):
    PARITY = 0  # reference-layout per-mesh BVHs (bit-exact) + the production tree
    FAST = 1  # production tree only
    LBVH = 2  # production tree only, Morton-sort builder (per-frame rebuilds)
    HYBRID = 3  # production tree only: LBVH's radix tree above, binned-SAH treelets of <= 768 triangles below


ERR_INVALID_ARG, ERR_DEGENERATE_GEOMETRY, ERR_CUDA, ERR_NOT_READY, ERR_CAPACITY, ERR_NCCL = 1, 2, 3, 4, 5, 6
GROUP_ID_BYTES = 128


class _Config(C.Structure):
    _fields_ = [("device", C.c_int), ("width", C.c_int), ("height", C.c_int), ("max_bounce", C.c_int), ("trace_mode", C.c_int), ("bvh_mode", C.c_int)]


class _Mesh(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("normals", C.c_void_p), ("indices", C.c_void_p), ("vertex_count", C.c_int), ("index_count", C.c_int)]


class _Instance(C.Structure):
    _fields_ = [("mesh", C.c_int), ("material", C.c_int), ("transform", C.c_float * 16)]


class _Material(C.Structure):
    _fields_ = [("base_color", C.c_float * 3), ("roughness", C.c_float), ("metallic", C.c_float), ("emissive", C.c_float)]


class _Texture(C.Structure):
    _fields_ = [("pixels", C.c_void_p), ("width", C.c_int), ("height", C.c_int), ("format", C.c_int)]


class Stats(C.Structure):
    _fields_ = [
        ("segments", C.c_uint64), ("node_visits", C.c_uint64), ("triangle_tests", C.c_uint64), ("kernel_launches", C.c_uint64),
        ("last_build_ms", C.c_float), ("last_fast_build_ms", C.c_float), ("last_render_ms", C.c_float),
        ("triangle_count", C.c_int), ("node_count", C.c_int), ("fast_node_count", C.c_int), ("instance_count", C.c_int),
        ("build_levels", C.c_int), ("build_launches", C.c_int), ("fast_build_levels", C.c_int), ("fast_build_launches", C.c_int),
        ("last_update_instances_ms", C.c_float), ("last_update_was_refit", C.c_int), ("collectives", C.c_uint64),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# every symbol include/cavept.h declares: (name, restype, argtypes)
_P = C.c_void_p
ABI = [
    ("cavept_create", C.c_int, [C.POINTER(_Config), C.POINTER(_P)]),
    ("cavept_destroy", None, [_P]),
    ("cavept_last_error", C.c_char_p, [_P]),
    ("cavept_get_image_size", C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("cavept_set_mode", C.c_int, [_P, C.c_int]),
    ("cavept_is_active", C.c_int, [_P]),
    ("cavept_upload_scene", C.c_int, [_P, C.POINTER(_Mesh), C.c_int, C.POINTER(_Instance), C.c_int, C.POINTER(_Material), C.c_int]),
    ("cavept_update_instances", C.c_int, [_P, C.POINTER(_Instance), C.c_int]),
    ("cavept_update_materials", C.c_int, [_P, C.POINTER(_Material), C.c_int]),
    ("cavept_set_refit", C.c_int, [_P, C.c_int]),
    ("cavept_rebuild_mesh", C.c_int, [_P, C.c_int, _P, _P]),
    ("cavept_get_mesh_bvh", C.c_int, [_P, C.c_int, _P, C.c_int]),
    ("cavept_get_fast_tree", C.c_int, [_P, _P, C.c_int]),
    ("cavept_set_mesh_uvs", C.c_int, [_P, C.c_int, _P]),
    ("cavept_upload_textures", C.c_int, [_P, _P, C.c_int]),
    ("cavept_set_material_maps", C.c_int, [_P, _P, _P, C.c_int]),
    ("cavept_set_env_map", C.c_int, [_P, _P, C.c_int, C.c_int]),
    ("cavept_get_scene_counts", C.c_int, [_P] + [C.POINTER(C.c_int)] * 5),
    ("cavept_get_scene_arrays", C.c_int, [_P] * 6),
    ("cavept_set_camera", C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float]),
    ("cavept_reset", C.c_int, [_P]),
    ("cavept_render", C.c_int, [_P, C.c_int, C.c_int]),
    ("cavept_sync", C.c_int, [_P]),
    ("cavept_render_tile", C.c_int, [_P] + [C.c_int] * 6),
    ("cavept_read_accum", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_read_tonemapped_rgba8", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_read_tonemapped_rgba32f", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_read_tonemapped_rgba16f", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_write_ppm", C.c_int, [_P, C.c_char_p]),
    ("cavept_last_render_ms", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("cavept_write_accum", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_get_stats", C.c_int, [_P, C.POINTER(Stats)]),
    ("cavept_reset_stats", C.c_int, [_P]),
    ("cavept_set_count_visits", C.c_int, [_P, C.c_int]),
    ("cavept_set_kernel_variant", C.c_int, [_P, C.c_int]),
    ("cavept_stream", _P, [_P]),
    ("cavept_accum_device_ptr", _P, [_P]),
    ("cavept_accum_to_sums", C.c_int, [_P]),
    ("cavept_accum_from_sums", C.c_int, [_P]),
    ("cavept_group_unique_id", C.c_int, [_P, C.c_size_t]),
    ("cavept_group_create", C.c_int, [_P, C.c_int, C.c_int, _P, C.c_size_t]),
    ("cavept_group_destroy", C.c_int, [_P]),
    ("cavept_group_info", C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("cavept_group_frame_block", C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("cavept_group_composite", C.c_int, [_P, C.c_int]),
    ("cavept_read_composite", C.c_int, [_P, _P, C.c_size_t]),
    ("cavept_composite_device_ptr", _P, [_P]),
    ("cavept_group_row_band", C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("cavept_group_gather_bands", C.c_int, [_P, C.c_int]),
    ("cavept_measure_gather", C.c_int, [_P, C.c_size_t, C.c_int, C.POINTER(C.c_double)]),
    ("cavept_sort_pairs", C.c_int, [_P, _P, _P, C.c_int, C.c_int]),
    ("cavept_sort_pairs_static", C.c_int, [_P, _P, _P, C.c_int, C.c_int]),
    ("cavept_host_alloc", _P, [C.c_size_t]),
    ("cavept_host_free", None, [_P]),
    ("cavept_probe_intersections", C.c_int, [_P, C.c_int, _P, _P, _P, _P]),
    ("cavept_scene_file_open", C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(_P), C.c_char_p, C.c_size_t]),
    ("cavept_mesh_file_open", C.c_int, [C.c_char_p, C.POINTER(_P), C.c_char_p, C.c_size_t]),
    ("cavept_scene_file_close", None, [_P]),
    ("cavept_scene_file_counts", C.c_int, [_P] + [C.POINTER(C.c_int)] * 4),
    ("cavept_scene_file_meshes", C.POINTER(_Mesh), [_P]),
    ("cavept_scene_file_instances", C.POINTER(_Instance), [_P]),
    ("cavept_scene_file_materials", C.POINTER(_Material), [_P]),
    ("cavept_upload_scene_file", C.c_int, [_P, _P]),
    ("cavept_mesh_file_write", C.c_int, [C.c_char_p, C.POINTER(_Mesh)]),
]


def library_path() -> str:
    return _LIB_PATH


def load_library() -> C.CDLL:
    """Loads libcavept.so and binds every ABI symbol.  Fails loudly if it is missing:
    there is no fallback implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise FileNotFoundError(f"{_LIB_PATH} is missing -- build it with `python -m caveengine_b200.build` (needs nvcc); cavept has no CPU fallback")
    lib = C.CDLL(_LIB_PATH)
    for name, restype, argtypes in ABI:
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


# numpy dtypes of the reference GPU layouts (cave/shader/structured_buffer.hlsl.h:24-61)
BVH_DT = np.dtype([("min", "<f4", 3), ("miss", "<i4"), ("max", "<f4", 3), ("hit", "<i4"), ("pad", "<f4", 2), ("leaf", "<i4"), ("tri", "<i4")])
FASTNODE_DT = np.dtype([("mn", "<f4", 3), ("a", "<i4"), ("mx", "<f4", 3), ("b", "<i4")])  # production-tree record (csrc/trace_body.h FastNode)
VERTEX_DT = np.dtype([("position", "<f4", 3), ("pad1", "<i4"), ("normal", "<f4", 3), ("pad2", "<i4")])
INDEX_DT = np.dtype([("tri", "<i4", 3), ("pad1", "<i4")])
MESH_DT = np.dtype([("transform", "<f4", 16), ("transform_inv", "<f4", 16), ("pad", "<f4", 2), ("root", "<i4"), ("material", "<i4")])
MATERIAL_DT = np.dtype([("base_color", "<f4", 3), ("roughness", "<f4"), ("emissive", "<f4", 3), ("metallic", "<f4")])


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class SceneFile:
    """A `.scene` (or a bare `.mesh`) of the engine parsed by the library's ingest
    (include/cavept.h, section "asset ingest").  Host-side only:
    usable without a GPU.  `to_scene(camera)` copies the arrays into a `scenes.Scene` so the
    same inputs can be handed to the oracle."""

    def __init__(self, path: str, resource_root: Optional[str] = None):
        self._lib = load_library()
        self._h = _P()
        err = C.create_string_buffer(512)
        if path.endswith(".mesh"):
            rc = self._lib.cavept_mesh_file_open(path.encode(), C.byref(self._h), err, len(err))
        else:
            rc = self._lib.cavept_scene_file_open(path.encode(), None if resource_root is None else resource_root.encode(), C.byref(self._h), err, len(err))
        if rc != 0:
            self._h = _P()
            raise CaveptError(rc, err.value.decode(errors="replace"))
        n = [C.c_int() for _ in range(4)]
        self._lib.cavept_scene_file_counts(self._h, *[C.byref(x) for x in n])
        self.mesh_count, self.instance_count, self.material_count, self.skipped_renderers = [x.value for x in n]

    def close(self):
        if getattr(self, "_h", None):
            self._lib.cavept_scene_file_close(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    def to_scene(self, camera=None, name: str = "file"):
        from . import scenes

        def arr(ptr, n, dt):
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n * np.dtype(dt).itemsize,)).view(dt).copy() if n else np.zeros(0, dt)

        meshes, cm = [], self._lib.cavept_scene_file_meshes(self._h)
        for k in range(self.mesh_count):
            m = cm[k]
            meshes.append(scenes.Mesh(arr(m.positions, m.vertex_count * 3, np.float32).reshape(-1, 3), arr(m.normals, m.vertex_count * 3, np.float32).reshape(-1, 3),
                                      arr(m.indices, m.index_count, np.uint32)))
        ci = self._lib.cavept_scene_file_instances(self._h)
        instances = [scenes.Instance(ci[k].mesh, ci[k].material, np.array(list(ci[k].transform), dtype=np.float32)) for k in range(self.instance_count)]
        cmat = self._lib.cavept_scene_file_materials(self._h)
        f = np.float32
        materials = [scenes.Material(tuple(f(x) for x in cmat[k].base_color), f(cmat[k].roughness), f(cmat[k].metallic), f(cmat[k].emissive)) for k in range(self.material_count)]
        return scenes.Scene(name, meshes, instances, materials, camera)


class PinnedArray:
    """numpy view over page-locked host memory (cavept_host_alloc): fill `.array` in place and pass it
    to `PathTracer.rebuild_mesh` for a DMA upload.  Freed on close()/garbage collection."""

    def __init__(self, shape, dtype=np.float32):
        self._lib = load_library()
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = self._lib.cavept_host_alloc(n)
        if not self._ptr:
            raise CaveptError(ERR_CUDA, f"cavept_host_alloc({n}) failed")
        self.array = np.ctypeslib.as_array(C.cast(self._ptr, C.POINTER(C.c_uint8)), shape=(n,)).view(dtype).reshape(shape)

    def close(self):
        if getattr(self, "_ptr", None):
            self.array = None
            self._lib.cavept_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_mesh_file(path: str, mesh) -> None:
    """Writes a `scenes.Mesh` in the engine's `.mesh` layout (cavept_mesh_file_write)."""
    lib = load_library()
    pos = np.ascontiguousarray(mesh.positions, dtype=np.float32)
    nrm = np.ascontiguousarray(mesh.normals, dtype=np.float32)
    idx = np.ascontiguousarray(mesh.indices, dtype=np.uint32)
    cm = _Mesh(pos.ctypes.data, nrm.ctypes.data, idx.ctypes.data, pos.shape[0], idx.size)
    rc = lib.cavept_mesh_file_write(path.encode(), C.byref(cm))
    if rc != 0:
        raise CaveptError(rc, f"cannot write {path}")


def group_unique_id() -> bytes:
    """cavept_group_unique_id: 128 bytes made on rank 0, to be handed to every rank of the group."""
    lib = load_library()
    buf = C.create_string_buffer(GROUP_ID_BYTES)
    rc = lib.cavept_group_unique_id(buf, GROUP_ID_BYTES)
    if rc != 0:
        raise CaveptError(rc, "cavept_group_unique_id failed (is NCCL loadable? set CAVEPT_NCCL_LIBRARY)")
    return buf.raw


def group_frame_block(first_frame: int, n_frames: int, rank: int, nranks: int):
    lib = load_library()
    f, c = C.c_int(), C.c_int()
    rc = lib.cavept_group_frame_block(first_frame, n_frames, rank, nranks, C.byref(f), C.byref(c))
    if rc != 0:
        raise CaveptError(rc, "cavept_group_frame_block: invalid arguments")
    return f.value, c.value


def group_row_band(height: int, rank: int, nranks: int, align: int = 8):
    lib = load_library()
    y, h = C.c_int(), C.c_int()
    rc = lib.cavept_group_row_band(height, rank, nranks, align, C.byref(y), C.byref(h))
    if rc != 0:
        raise CaveptError(rc, "cavept_group_row_band: invalid arguments")
    return y.value, h.value


class PathTracer:
    """One context == the reference's file-static PathTracer
    (cave/engine/renderer/path_tracer_render_system.cpp:8-10), on one GPU."""

    def __init__(self, width: int = 1920, height: int = 1080, max_bounce: int = 5, trace_mode: TraceMode = TraceMode.FAST, device: int = -1,
                 bvh_mode: BvhMode = BvhMode.PARITY):
        self._lib = load_library()
        self._ctx = _P()
        self.width, self.height, self.max_bounce = width, height, max_bounce
        cfg = _Config(device, width, height, max_bounce, int(trace_mode), int(bvh_mode))
        rc = self._lib.cavept_create(C.byref(cfg), C.byref(self._ctx))
        if rc != 0:
            self._ctx = _P()
            raise CaveptError(rc, "cavept_create failed (is a CUDA device visible? cavept has no CPU path)")
        self._keep = None
        self._resident = False

    # -- lifetime ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.cavept_destroy(self._ctx)
            self._ctx = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise CaveptError(rc, self._lib.cavept_last_error(self._ctx).decode(errors="replace"))

    # -- SetPathTracerMode / IsPathTracerActive -----------------------------------------
    def set_mode(self, mode: PathTracerMode):
        self._check(self._lib.cavept_set_mode(self._ctx, int(mode)))

    def is_active(self) -> bool:
        return bool(self._lib.cavept_is_active(self._ctx))

    # -- RequestPathTracerUpdate ---------------------------------------------------------
    def request_update(self, scene):
        """PathTracer::Update (path_tracer.cpp:49-70): no-op unless the mode is
        INTERACTIVE/TILED; first call uploads the scene and builds the BVHs."""
        if not self._resident:
            self.upload_scene(scene)
            self._resident = True
        self.set_camera(scene.camera)

    def upload_scene(self, scene):
        meshes = (_Mesh * len(scene.meshes))()
        keep = []
        for k, m in enumerate(scene.meshes):
            pos = np.ascontiguousarray(m.positions, dtype=np.float32)
            nrm = np.ascontiguousarray(m.normals, dtype=np.float32)
            idx = np.ascontiguousarray(m.indices, dtype=np.uint32)
            keep += [pos, nrm, idx]
            meshes[k] = _Mesh(pos.ctypes.data, nrm.ctypes.data, idx.ctypes.data, pos.shape[0], idx.size)
        inst = self._instances(scene.instances)
        mats = self._materials(scene.materials)
        self._check(self._lib.cavept_upload_scene(self._ctx, meshes, len(scene.meshes), inst, len(scene.instances), mats, len(scene.materials)))

    def upload_scene_file(self, scene_file: "SceneFile"):
        """cavept_upload_scene_file: the parsed arrays go to the device without a Python copy."""
        self._check(self._lib.cavept_upload_scene_file(self._ctx, scene_file.handle))
        self._resident = True

    @staticmethod
    def _instances(instances):
        arr = (_Instance * len(instances))()
        for k, i in enumerate(instances):
            arr[k] = _Instance(i.mesh, i.material, (C.c_float * 16)(*[float(x) for x in np.asarray(i.transform, dtype=np.float32)]))
        return arr

    @staticmethod
    def _materials(materials):
        arr = (_Material * len(materials))()
        for k, m in enumerate(materials):
            arr[k] = _Material(_f3(m.base_color), m.roughness, m.metallic, m.emissive)
        return arr

    def update_instances(self, instances):
        self._check(self._lib.cavept_update_instances(self._ctx, self._instances(instances), len(instances)))

    def set_refit(self, enable: bool):
        self._check(self._lib.cavept_set_refit(self._ctx, 1 if enable else 0))

    # -- textured materials / environment map (the reference's switched-off RayColor) ---------
    def set_mesh_uvs(self, mesh: int, uvs: np.ndarray):
        uv = np.ascontiguousarray(uvs, dtype=np.float32)
        self._check(self._lib.cavept_set_mesh_uvs(self._ctx, mesh, uv.ctypes.data))

    def upload_textures(self, textures):
        """textures: list of (H, W, 4) arrays, uint8 (unorm) or float32."""
        arr = (_Texture * max(len(textures), 1))()
        keep = []
        for k, t in enumerate(textures):
            a = np.ascontiguousarray(t, dtype=np.uint8 if np.asarray(t).dtype == np.uint8 else np.float32)
            assert a.ndim == 3 and a.shape[2] == 4
            keep.append(a)
            arr[k] = _Texture(a.ctypes.data, a.shape[1], a.shape[0], 0 if a.dtype == np.uint8 else 1)
        self._check(self._lib.cavept_upload_textures(self._ctx, arr, len(textures)))

    def set_material_maps(self, base_color_maps, material_maps):
        a = np.ascontiguousarray(base_color_maps, dtype=np.int32)
        b = np.ascontiguousarray(material_maps, dtype=np.int32)
        assert a.shape == b.shape
        self._check(self._lib.cavept_set_material_maps(self._ctx, a.ctypes.data, b.ctypes.data, a.shape[0]))

    def set_env_map(self, rgb: Optional[np.ndarray]):
        if rgb is None:
            self._check(self._lib.cavept_set_env_map(self._ctx, None, 0, 0))
            return
        a = np.ascontiguousarray(rgb, dtype=np.float32)
        assert a.ndim == 3 and a.shape[2] == 3
        self._check(self._lib.cavept_set_env_map(self._ctx, a.ctypes.data, a.shape[1], a.shape[0]))

    def update_materials(self, materials):
        self._check(self._lib.cavept_update_materials(self._ctx, self._materials(materials), len(materials)))

    def rebuild_mesh(self, mesh: int, positions: np.ndarray, normals: Optional[np.ndarray] = None):
        pos = np.ascontiguousarray(positions, dtype=np.float32)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        self._check(self._lib.cavept_rebuild_mesh(self._ctx, mesh, pos.ctypes.data, None if nrm is None else nrm.ctypes.data))

    # -- parity read-backs ----------------------------------------------------------------
    def get_mesh_bvh(self, mesh: int, n_tris: int) -> np.ndarray:
        out = np.zeros(max(2 * n_tris - 1, 1), dtype=BVH_DT)
        rc = self._lib.cavept_get_mesh_bvh(self._ctx, mesh, out.ctypes.data, out.shape[0])
        if rc < 0:
            self._check(-rc)
        return out[:rc]

    def get_fast_tree(self) -> np.ndarray:
        """The production tree's fp32 records (sibling pairs adjacent), as cavept_get_fast_tree returns them."""
        cap = max(2 * self.stats()["fast_node_count"], 2)
        out = np.zeros(cap, dtype=FASTNODE_DT)
        rc = self._lib.cavept_get_fast_tree(self._ctx, out.ctypes.data, cap)
        if rc < 0:
            self._check(-rc)
        return out[:rc]

    def get_scene_arrays(self):
        n = [C.c_int() for _ in range(5)]
        self._check(self._lib.cavept_get_scene_counts(self._ctx, *[C.byref(x) for x in n]))
        nv, nt, nb, ni, nm = [x.value for x in n]
        v, i, b = np.zeros(nv, dtype=VERTEX_DT), np.zeros(nt, dtype=INDEX_DT), np.zeros(nb, dtype=BVH_DT)
        m, mat = np.zeros(ni, dtype=MESH_DT), np.zeros(nm, dtype=MATERIAL_DT)
        self._check(self._lib.cavept_get_scene_arrays(self._ctx, v.ctypes.data, i.ctypes.data, b.ctypes.data, m.ctypes.data, mat.ctypes.data))
        return {"vertices": v, "indices": i, "bvhs": b, "meshes": m, "materials": mat}

    # -- per-frame constants ---------------------------------------------------------------
    def set_camera(self, cam):
        self._check(self._lib.cavept_set_camera(self._ctx, _f3(cam.position), _f3(cam.right), _f3(cam.up), _f3(cam.forward), float(cam.fov_degree)))

    def reset(self):
        self._check(self._lib.cavept_reset(self._ctx))

    # -- the pass ----------------------------------------------------------------------------
    def render(self, first_frame_index: int, n_frames: int = 1, sync: bool = False):
        self._check(self._lib.cavept_render(self._ctx, first_frame_index, n_frames))
        if sync:
            self.sync()

    def render_tile(self, x0: int, y0: int, w: int, h: int, first_frame_index: int, n_frames: int = 1, sync: bool = False):
        """PathTracerMode::TILED: the pass over one pixel rectangle (c_tileOffset, cbuffer.hlsl.h:183)."""
        self._check(self._lib.cavept_render_tile(self._ctx, x0, y0, w, h, first_frame_index, n_frames))
        if sync:
            self.sync()

    def sync(self):
        self._check(self._lib.cavept_sync(self._ctx))

    # -- read-back ---------------------------------------------------------------------------
    def read_accum(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._check(self._lib.cavept_read_accum(self._ctx, out.ctypes.data, out.nbytes))
        return out

    def write_accum(self, accum: np.ndarray):
        a = np.ascontiguousarray(accum, dtype=np.float32)
        self._check(self._lib.cavept_write_accum(self._ctx, a.ctypes.data, a.nbytes))

    def read_tonemapped_rgba8(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        self._check(self._lib.cavept_read_tonemapped_rgba8(self._ctx, out.ctypes.data, out.nbytes))
        return out

    def read_tonemapped_rgba16f(self) -> np.ndarray:
        """RT_FMT_TONE (R16G16B16A16_FLOAT): (H, W, 4) float16."""
        out = np.empty((self.height, self.width, 4), dtype=np.float16)
        self._check(self._lib.cavept_read_tonemapped_rgba16f(self._ctx, out.ctypes.data, out.nbytes))
        return out

    def write_ppm(self, path: str):
        self._check(self._lib.cavept_write_ppm(self._ctx, path.encode()))

    def last_render_ms(self) -> Optional[float]:
        """Non-blocking: None while the last launch is still running."""
        ms = C.c_float()
        rc = self._lib.cavept_last_render_ms(self._ctx, C.byref(ms))
        if rc == ERR_NOT_READY:
            return None
        self._check(rc)
        return ms.value

    def read_tonemapped(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._check(self._lib.cavept_read_tonemapped_rgba32f(self._ctx, out.ctypes.data, out.nbytes))
        return out

    # -- measurement / multi-GPU plumbing ---------------------------------------------------
    def stats(self) -> dict:
        s = Stats()
        self._check(self._lib.cavept_get_stats(self._ctx, C.byref(s)))
        return s.as_dict()

    def reset_stats(self):
        self._check(self._lib.cavept_reset_stats(self._ctx))

    def set_kernel_variant(self, variant: "KernelVariant"):
        self._check(self._lib.cavept_set_kernel_variant(self._ctx, int(variant)))

    def set_count_visits(self, enable: bool):
        self._check(self._lib.cavept_set_count_visits(self._ctx, 1 if enable else 0))

    @property
    def stream_handle(self) -> int:
        return int(self._lib.cavept_stream(self._ctx) or 0)

    @property
    def accum_device_ptr(self) -> int:
        return int(self._lib.cavept_accum_device_ptr(self._ctx) or 0)

    def accum_to_sums(self):
        self._check(self._lib.cavept_accum_to_sums(self._ctx))

    def accum_from_sums(self):
        self._check(self._lib.cavept_accum_from_sums(self._ctx))

    # -- multi-GPU group (include/cavept.h "multi-GPU group") -------------------------------------
    def group_create(self, rank: int, nranks: int, unique_id: Optional[bytes] = None):
        buf = C.create_string_buffer(unique_id, GROUP_ID_BYTES) if unique_id is not None else None
        self._check(self._lib.cavept_group_create(self._ctx, rank, nranks, buf, GROUP_ID_BYTES if buf is not None else 0))

    def group_destroy(self):
        self._check(self._lib.cavept_group_destroy(self._ctx))

    def group_info(self):
        r, n, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self._lib.cavept_group_info(self._ctx, C.byref(r), C.byref(n), C.byref(v)))
        return {"rank": r.value, "nranks": n.value, "nccl_version": v.value}

    def group_composite(self, dst_rank: int = 0):
        self._check(self._lib.cavept_group_composite(self._ctx, dst_rank))

    def group_gather_bands(self, align: int = 8):
        self._check(self._lib.cavept_group_gather_bands(self._ctx, align))

    def read_composite(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._check(self._lib.cavept_read_composite(self._ctx, out.ctypes.data, out.nbytes))
        return out

    def measure_gather(self, working_set_bytes: int, iters: int = 2000) -> float:
        """cavept_measure_gather: random 32-byte LDG.E.256 lane-gathers per second from a working set of this size."""
        out = C.c_double()
        self._check(self._lib.cavept_measure_gather(self._ctx, int(working_set_bytes), int(iters), C.byref(out)))
        return out.value

    def sort_pairs(self, keys: np.ndarray, values: np.ndarray, key_bits: int = 32):
        """Device radix sort of (uint32 key, int32 value) pairs, in place (stable)."""
        assert keys.dtype == np.uint32 and values.dtype == np.int32 and keys.flags.c_contiguous and values.flags.c_contiguous and keys.size == values.size
        self._check(self._lib.cavept_sort_pairs(self._ctx, keys.ctypes.data, values.ctypes.data, int(keys.size), int(key_bits)))

    STATIC_SORT_MAX_PAIRS = 1 << 20  # CAVEPT_STATIC_SORT_MAX_PAIRS

    def sort_pairs_static(self, d_keys: int, d_values: int, count: int, key_bits: int = 32):
        """cavept_sort_pairs_static: the same sort on DEVICE arrays (addresses), in place, asynchronous on the context's
        stream, with no allocation -- its scratch is a static device array in the library.  count is limited to
        STATIC_SORT_MAX_PAIRS; calls that overlap on one GPU race on the shared scratch."""
        self._check(self._lib.cavept_sort_pairs_static(self._ctx, C.c_void_p(int(d_keys)), C.c_void_p(int(d_values)), int(count), int(key_bits)))

    def probe_intersections(self, rays: np.ndarray, triangles: np.ndarray, boxes: np.ndarray) -> np.ndarray:
        """cavept_probe_intersections: (n,7) rays, (n,9) triangles, (n,6) boxes -> (n,8) results."""
        r = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 7)
        t = np.ascontiguousarray(triangles, dtype=np.float32).reshape(-1, 9)
        b = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
        assert r.shape[0] == t.shape[0] == b.shape[0]
        out = np.zeros((r.shape[0], 8), dtype=np.float32)
        self._check(self._lib.cavept_probe_intersections(self._ctx, r.shape[0], r.ctypes.data, t.ctypes.data, b.ctypes.data, out.ctypes.data))
        return out

    def accum_as_torch(self):
        """Zero-copy torch view of the device accumulator (H, W, 4) float32, for
        torch.distributed collectives.  torch is plumbing only."""
        import torch

        class _Holder:
            pass

        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (self.height, self.width, 4), "typestr": "<f4", "data": (self.accum_device_ptr, False), "version": 3, "strides": None}
        t = torch.as_tensor(h, device="cuda")
        self._keep = h
        return t
