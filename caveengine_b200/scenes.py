"""Deterministic procedural scenes of the shapes BASELINE.json names (no assets, no
network).  Host-side numpy only; feeds both the product (through the C ABI) and the
oracle, so both see byte-identical inputs.

A scene mirrors what the reference's uploader reads from the ECS
(cave/engine/renderer/path_tracer/path_tracer.cpp:203-310):
  meshes    -- MeshAsset{indices u32, positions vec3, normals vec3}
               (cave/engine/assets/mesh_asset.h:44-46)
  instances -- MeshRendererComponent + TransformComponent world matrix (column-major
               glm mat4) + material (first material of the renderer)
  materials -- MaterialComponent{base_color, metallic, roughness, emissive}
               (cave/engine/scene/material_component.h:12-21)
  camera    -- c_cameraPosition/Forward/Right/Up/FovDegree
               (cave/engine/runtime/render_system.cpp:89-93, up = cross(front, right) :273)

Winding: the reference culls back faces (shared_path_tracer.h:106), so every
triangle is wound so that cross(B-A, C-A) points toward the side it is seen from.
"""
from __future__ import annotations

import dataclasses
from typing import List

import numpy as np

f32 = np.float32


@dataclasses.dataclass
class Mesh:
    positions: np.ndarray  # (V,3) f32
    normals: np.ndarray  # (V,3) f32
    indices: np.ndarray  # (T*3,) u32

    @property
    def n_tris(self) -> int:
        return self.indices.size // 3


@dataclasses.dataclass
class Instance:
    mesh: int
    material: int
    transform: np.ndarray  # (16,) f32 column-major


@dataclasses.dataclass
class Material:
    base_color: tuple
    roughness: float
    metallic: float
    emissive: float  # scalar, GpuPtMaterial.emissive = base_color * emissive (path_tracer.cpp:141)


@dataclasses.dataclass
class Camera:
    position: tuple
    forward: tuple
    right: tuple
    up: tuple
    fov_degree: float = 50.0  # camera_component.h:68


@dataclasses.dataclass
class Scene:
    name: str
    meshes: List[Mesh]
    instances: List[Instance]
    materials: List[Material]
    camera: Camera

    @property
    def n_tris(self) -> int:
        return sum(self.meshes[i.mesh].n_tris for i in self.instances)


# ----------------------------------------------------------------------------
# integer-hash noise: identical on every machine (no libm, no numpy RNG streams)
# ----------------------------------------------------------------------------
def _hash_u32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64)
    x = (x ^ 61) ^ (x >> 16)
    x = (x * 9) & 0xFFFFFFFF
    x = x ^ (x >> 4)
    x = (x * 0x27D4EB2D) & 0xFFFFFFFF
    x = x ^ (x >> 15)
    return x.astype(np.uint32)


def hash01(seed: int, *idx: np.ndarray) -> np.ndarray:
    """Uniform [0,1) floats from integer lattice coordinates."""
    h = np.uint32(seed * 2654435761 & 0xFFFFFFFF)
    acc = np.zeros(np.broadcast(*idx).shape, dtype=np.uint32) + h
    for k, i in enumerate(idx):
        acc = _hash_u32(acc ^ (np.asarray(i).astype(np.uint32) * np.uint32(0x9E3779B1 + 2 * k)))
    return (acc >> np.uint32(8)).astype(np.float64) / float(1 << 24)


def _round6(a: np.ndarray) -> np.ndarray:
    """Quantise to 2^-16 so trig ulp differences between hosts cannot change the f32 scene."""
    return (np.round(np.asarray(a, dtype=np.float64) * 65536.0) / 65536.0).astype(f32)


def _smooth_normals(pos: np.ndarray, idx: np.ndarray) -> np.ndarray:
    tri = idx.reshape(-1, 3)
    p = pos.astype(np.float64)
    fn = np.cross(p[tri[:, 1]] - p[tri[:, 0]], p[tri[:, 2]] - p[tri[:, 0]])
    n = np.zeros_like(p)
    for k in range(3):
        np.add.at(n, tri[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return _round6(n / ln)


def grid_mesh(nu: int, nv: int, fn, flip: bool = False, wrap_u: bool = False) -> Mesh:
    """Parametric surface over an nu x nv quad lattice; fn(u, v) -> (x, y, z) with
    u, v in [0,1].  Front face = cross(dP/du, dP/dv) unless flip."""
    cu = nu if wrap_u else nu + 1
    iu, iv = np.meshgrid(np.arange(cu), np.arange(nv + 1), indexing="ij")
    u = iu.astype(np.float64) / nu
    v = iv.astype(np.float64) / nv
    x, y, z = fn(u, v, iu, iv)
    pos = _round6(np.stack([x, y, z], axis=-1).reshape(-1, 3))
    qi, qj = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    qi = qi.reshape(-1)
    qj = qj.reshape(-1)
    qi1 = (qi + 1) % cu if wrap_u else qi + 1
    a = qi * (nv + 1) + qj
    b = qi1 * (nv + 1) + qj
    c = qi1 * (nv + 1) + qj + 1
    d = qi * (nv + 1) + qj + 1
    if flip:
        tris = np.stack([a, c, b, a, d, c], axis=-1)
    else:
        tris = np.stack([a, b, c, a, c, d], axis=-1)
    idx = tris.reshape(-1).astype(np.uint32)
    return Mesh(pos, _smooth_normals(pos, idx), idx)


def merge(meshes: List[Mesh]) -> Mesh:
    pos, nrm, idx = [], [], []
    off = 0
    for m in meshes:
        pos.append(m.positions)
        nrm.append(m.normals)
        idx.append(m.indices + np.uint32(off))
        off += m.positions.shape[0]
    return Mesh(np.concatenate(pos), np.concatenate(nrm), np.concatenate(idx).astype(np.uint32))


def translate(m: Mesh, t) -> Mesh:
    return Mesh((m.positions + np.asarray(t, dtype=f32)).astype(f32), m.normals, m.indices)


def mat_identity() -> np.ndarray:
    return np.eye(4, dtype=f32).T.reshape(-1).copy()


def mat_trs(t=(0, 0, 0), yaw_deg: float = 0.0, s=1.0) -> np.ndarray:
    """Column-major T * R_y * S as float32 (values rounded so hosts agree)."""
    a = np.deg2rad(yaw_deg)
    c, sn = float(_round6(np.cos(a))), float(_round6(np.sin(a)))
    sx, sy, sz = (s, s, s) if np.isscalar(s) else s
    m = np.array([[c * sx, 0, sn * sz, t[0]], [0, sy, 0, t[1]], [-sn * sx, 0, c * sz, t[2]], [0, 0, 0, 1]], dtype=np.float64)
    return m.T.reshape(-1).astype(f32)


# ----------------------------------------------------------------------------
# config 1: 10k-triangle jittered heightfield
# ----------------------------------------------------------------------------
def heightfield_mesh(nx: int = 50, nz: int = 100, seed: int = 12345, cell: float = 0.2, jitter: float = 0.3,
                     height: float = 2.0) -> Mesh:
    def fn(u, v, iu, iv):
        jx = (hash01(seed, iu, iv, 0) - 0.5) * jitter * cell
        jz = (hash01(seed, iu, iv, 1) - 0.5) * jitter * cell
        y = hash01(seed, iu, iv, 2) * height
        return iu * cell + jx, y, iv * cell + jz

    # cross(dP/du, dP/dv) = cross(+x, +z) = -y, so flip to face +y
    return grid_mesh(nx, nz, fn, flip=True)


def heightfield_scene(nx: int = 50, nz: int = 100, seed: int = 12345) -> Scene:
    mesh = heightfield_mesh(nx, nz, seed)
    ex, ez = nx * 0.2, nz * 0.2
    cam = look_at((ex * 0.5, 6.0, -ez * 0.35), (ex * 0.5, 0.5, ez * 0.5))
    mats = [Material((0.7, 0.6, 0.5), 0.8, 0.0, 0.0)]
    return Scene(f"heightfield_{mesh.n_tris}", [mesh], [Instance(0, 0, mat_identity())], mats, cam)


def look_at(eye, target, fov: float = 50.0) -> Camera:
    eye = np.asarray(eye, dtype=np.float64)
    fwd = np.asarray(target, dtype=np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    r6 = lambda a: tuple(float(x) for x in _round6(a))
    return Camera(r6(eye), r6(fwd), r6(right), r6(up), fov)


# ----------------------------------------------------------------------------
# config 2: Cornell box, 8 meshes / 32 triangles
# ----------------------------------------------------------------------------
def _quad(p0, p1, p2, p3) -> Mesh:
    """Quad p0..p3 counter-clockwise seen from its front side."""
    pos = np.array([p0, p1, p2, p3], dtype=f32)
    idx = np.array([0, 1, 2, 0, 2, 3], dtype=np.uint32)
    n = np.cross(pos[1].astype(np.float64) - pos[0], pos[2].astype(np.float64) - pos[0])
    n = n / np.linalg.norm(n)
    return Mesh(pos, np.tile(_round6(n), (4, 1)), idx)


def _open_box(w: float, h: float, d: float) -> Mesh:
    """Axis-aligned box [-w/2,w/2]x[0,h]x[-d/2,d/2], outward faces, no bottom: 10 triangles."""
    x0, x1, z0, z1 = -w / 2, w / 2, -d / 2, d / 2
    faces = [
        _quad((x0, h, z1), (x1, h, z1), (x1, h, z0), (x0, h, z0)),  # top (+y)
        _quad((x0, 0, z1), (x1, 0, z1), (x1, h, z1), (x0, h, z1)),  # +z
        _quad((x1, 0, z0), (x0, 0, z0), (x0, h, z0), (x1, h, z0)),  # -z
        _quad((x1, 0, z1), (x1, 0, z0), (x1, h, z0), (x1, h, z1)),  # +x
        _quad((x0, 0, z0), (x0, 0, z1), (x0, h, z1), (x0, h, z0)),  # -x
    ]
    return merge(faces)


def cornell_scene() -> Scene:
    s = 2.0  # half extent; room is [-2,2] x [0,4] x [-2,2], open toward -z where the camera sits
    meshes = [
        _quad((-s, 0, -s), (-s, 0, s), (s, 0, s), (s, 0, -s)),  # floor, faces +y
        _quad((-s, 2 * s, -s), (s, 2 * s, -s), (s, 2 * s, s), (-s, 2 * s, s)),  # ceiling, faces -y
        _quad((-s, 0, s), (-s, 2 * s, s), (s, 2 * s, s), (s, 0, s)),  # back wall, faces -z
        _quad((-s, 0, -s), (-s, 2 * s, -s), (-s, 2 * s, s), (-s, 0, s)),  # left wall, faces +x
        _quad((s, 0, -s), (s, 0, s), (s, 2 * s, s), (s, 2 * s, -s)),  # right wall, faces -x
        _quad((-0.6, 3.99, -0.6), (0.6, 3.99, -0.6), (0.6, 3.99, 0.6), (-0.6, 3.99, 0.6)),  # light, faces -y
        _open_box(1.2, 2.4, 1.2),  # tall box
        _open_box(1.2, 1.2, 1.2),  # short box
    ]
    mats = [
        Material((0.73, 0.73, 0.73), 0.9, 0.0, 0.0),  # white
        Material((0.65, 0.05, 0.05), 0.9, 0.0, 0.0),  # red
        Material((0.12, 0.45, 0.15), 0.9, 0.0, 0.0),  # green
        Material((1.0, 0.9, 0.7), 1.0, 0.0, 12.0),  # light
        Material((0.8, 0.8, 0.85), 0.15, 0.9, 0.0),  # glossy metal box
    ]
    inst = [
        Instance(0, 0, mat_identity()),
        Instance(1, 0, mat_identity()),
        Instance(2, 0, mat_identity()),
        Instance(3, 1, mat_identity()),
        Instance(4, 2, mat_identity()),
        Instance(5, 3, mat_identity()),
        Instance(6, 4, mat_trs((-0.7, 0.0, 0.6), yaw_deg=18.0)),
        Instance(7, 0, mat_trs((0.7, 0.0, -0.5), yaw_deg=-17.0)),
    ]
    cam = look_at((0.0, 2.0, -5.2), (0.0, 2.0, 0.0), fov=50.0)
    return Scene("cornell_32", meshes, inst, mats, cam)


# ----------------------------------------------------------------------------
# config 3: Sponza-shaped atrium, 262,144 triangles in 25 meshes / 25 materials
# ----------------------------------------------------------------------------
def _relief(seed, iu, iv, amp):
    return (hash01(seed, iu, iv) - 0.5) * amp


def _plane_xz(nu, nv, x0, x1, z0, z1, y, seed, amp, up=True) -> Mesh:
    def fn(u, v, iu, iv):
        return x0 + (x1 - x0) * u, y + _relief(seed, iu, iv, amp), z0 + (z1 - z0) * v

    return grid_mesh(nu, nv, fn, flip=up)  # cross(+x,+z) = -y


def _wall_xy(nu, nv, x0, x1, y0, y1, z, seed, amp, face_neg_z=True) -> Mesh:
    def fn(u, v, iu, iv):
        return x0 + (x1 - x0) * u, y0 + (y1 - y0) * v, z + _relief(seed, iu, iv, amp)

    return grid_mesh(nu, nv, fn, flip=face_neg_z)  # cross(+x,+y) = +z


def _wall_zy(nu, nv, z0, z1, y0, y1, x, seed, amp, face_neg_x=True) -> Mesh:
    def fn(u, v, iu, iv):
        return x + _relief(seed, iu, iv, amp), y0 + (y1 - y0) * v, z0 + (z1 - z0) * u

    return grid_mesh(nu, nv, fn, flip=not face_neg_x)  # cross(+z,+y) = -x


def _cylinder(nseg, nring, radius, y0, y1, seed, flute=0.04) -> Mesh:
    def fn(u, v, iu, iv):
        a = 2.0 * np.pi * u
        r = radius * (1.0 + flute * np.cos(8 * a)) + _relief(seed, iu % nseg, iv, 0.004)
        return r * np.cos(a), y0 + (y1 - y0) * v, r * np.sin(a)

    # cross(dP/da, dP/dy) points inward for this parametrisation -> flip for outward
    return grid_mesh(nseg, nring, fn, flip=True, wrap_u=True)


def _sphere(nseg, nring, radius, seed, bump=0.0, squash=1.0) -> Mesh:
    def fn(u, v, iu, iv):
        a = 2.0 * np.pi * u
        # keep away from the poles so no zero-area triangles / coincident centroids
        th = np.pi * (0.02 + 0.96 * v)
        r = radius * (1.0 + bump * (hash01(seed, iu % nseg, iv) - 0.5))
        return r * np.sin(th) * np.cos(a), squash * r * np.cos(th), r * np.sin(th) * np.sin(a)

    return grid_mesh(nseg, nring, fn, flip=False, wrap_u=True)


def _arch(nu, nv, span, tube, seed) -> Mesh:
    """Half torus standing in the xy plane, feet on y=0, spanning x in [-span/2, span/2]."""
    R = span / 2.0

    def fn(u, v, iu, iv):
        a = np.pi * u  # along the arch
        b = 2.0 * np.pi * v  # around the tube
        rr = R + tube * np.cos(b) + _relief(seed, iu, iv % nv, 0.003)
        return -rr * np.cos(a), rr * np.sin(a), tube * np.sin(b)

    m = grid_mesh(nu, nv, fn, flip=False)
    return m


def _drape(nu, nv, w, h, seed) -> Mesh:
    def fn(u, v, iu, iv):
        wave = 0.18 * np.sin(7.0 * np.pi * u + 1.3 * (seed % 5)) * (0.3 + 0.7 * (1.0 - v))
        return w * (u - 0.5), h * v, wave + _relief(seed, iu, iv, 0.006)

    return grid_mesh(nu, nv, fn, flip=False)


def sponza_scene(seed: int = 2024) -> Scene:
    L, W, H = 30.0, 12.0, 14.0  # atrium: x in [-15,15], z in [-6,6], y in [0,14]
    meshes: List[Mesh] = []
    inst_tf: List[np.ndarray] = []

    def add(mesh: Mesh, tf=None):
        meshes.append(mesh)
        inst_tf.append(mat_identity() if tf is None else tf)

    add(_plane_xz(128, 64, -L / 2, L / 2, -W / 2, W / 2, 0.0, seed + 1, 0.02, up=True))  # 0 floor 16384
    add(_wall_xy(128, 48, -L / 2, L / 2, 0.0, H, W / 2, seed + 2, 0.03, face_neg_z=True))  # 1 wall +z 12288
    add(_wall_xy(128, 48, -L / 2, L / 2, 0.0, H, -W / 2, seed + 3, 0.03, face_neg_z=False))  # 2 wall -z 12288
    add(_wall_zy(64, 48, -W / 2, W / 2, 0.0, H, L / 2, seed + 4, 0.03, face_neg_x=True))  # 3 end +x 6144
    add(_wall_zy(64, 48, -W / 2, W / 2, 0.0, H, -L / 2, seed + 5, 0.03, face_neg_x=False))  # 4 end -x 6144
    roof = merge([
        _plane_xz(128, 16, -L / 2, L / 2, -W / 2, -W / 6, H, seed + 6, 0.02, up=False),
        _plane_xz(128, 16, -L / 2, L / 2, W / 6, W / 2, H, seed + 7, 0.02, up=False),
    ])
    add(roof)  # 5 roof strips 8192 (open slot in the middle lets the sky in)
    # 6-9 column rows: 8 columns of 32x32 quads each (2048 tris) -> 16384 per mesh
    col_x = [-13.0 + 26.0 * i / 7.0 for i in range(8)]
    for r, (z, y0, y1, rad) in enumerate([(-3.6, 0.0, 6.0, 0.38), (3.6, 0.0, 6.0, 0.38), (-3.6, 7.0, 12.5, 0.28), (3.6, 7.0, 12.5, 0.28)]):
        cols = [translate(_cylinder(32, 32, rad, y0, y1, seed + 10 + r * 8 + i), (0.0, 0.0, 0.0)) for i in range(8)]
        cols = [translate(c, (col_x[i], 0.0, 0.0)) for i, c in enumerate(cols)]
        add(merge(cols), mat_trs((0.0, 0.0, z)))
    # 10-13 arches: 7 per mesh of 32x16 quads (1024 tris) -> 7168 per mesh
    for r, (z, y) in enumerate([(-3.6, 6.0), (3.6, 6.0), (-3.6, 12.5), (3.6, 12.5)]):
        span = col_x[1] - col_x[0]
        arcs = [translate(_arch(32, 16, span * 0.9, 0.16, seed + 50 + r * 7 + i), (0.5 * (col_x[i] + col_x[i + 1]), 0.0, 0.0)) for i in range(7)]
        add(merge(arcs), mat_trs((0.0, y - 1.2 if r >= 2 else y - 1.6, z), s=(1.0, 0.55, 1.0)))
    # 14-15 gallery floors 128x8 -> 2048 each (seen from below and above: two-sided by doubling is avoided; face down)
    add(_plane_xz(128, 8, -L / 2, L / 2, -W / 2, -3.6, 6.6, seed + 80, 0.01, up=False))
    add(_plane_xz(128, 8, -L / 2, L / 2, 3.6, W / 2, 6.6, seed + 81, 0.01, up=False))
    # 16-21 drapes 64x64 -> 8192 each
    drape_pos = [(-9.0, 7.2, -3.2, 0.0), (0.0, 7.2, -3.2, 0.0), (9.0, 7.2, -3.2, 0.0), (-9.0, 7.2, 3.2, 180.0), (0.0, 7.2, 3.2, 180.0), (9.0, 7.2, 3.2, 180.0)]
    for i, (x, y, z, yaw) in enumerate(drape_pos):
        add(_drape(64, 64, 4.5, 5.0, seed + 90 + i), mat_trs((x, y, z), yaw_deg=yaw + 180.0))
    # 22 lion-head-ish displaced blob 128x64 -> 16384
    add(_sphere(128, 64, 1.1, seed + 100, bump=0.25, squash=1.2), mat_trs((12.5, 2.4, 0.0), yaw_deg=30.0))
    # 23 four vases 64x32 -> 4096 each
    vases = [translate(_sphere(64, 32, 0.55, seed + 110 + i, bump=0.05, squash=1.6), (x, 0.95, z)) for i, (x, z) in enumerate([(-10.0, -1.5), (-4.0, 1.5), (3.0, -1.5), (8.0, 1.5)])]
    add(merge(vases))
    # 24 ten emissive lamps 32x32 -> 2048 each = 20480
    lamps = [translate(_sphere(32, 32, 0.22, seed + 120 + i), (-13.5 + 3.0 * i, 5.2, (-2.9 if i % 2 == 0 else 2.9))) for i in range(10)]
    add(merge(lamps))

    assert len(meshes) == 25
    total = sum(m.n_tris for m in meshes)
    assert total == 262144, total

    mats: List[Material] = []
    for i in range(25):
        c = tuple(float(f32(0.2 + 0.7 * hash01(seed + 200, np.array(i), np.array(k)))) for k in range(3))
        rough = float(f32(0.1 + 0.9 * hash01(seed + 201, np.array(i))))
        metal = (0.0, 0.0, 0.5, 0.9)[int(hash01(seed + 202, np.array(i)) * 4) % 4]
        mats.append(Material(c, rough, metal, 0.0))
    mats[24] = Material((1.0, 0.85, 0.6), 1.0, 0.0, 20.0)  # lamps
    mats[22] = Material(mats[22].base_color, 0.35, 0.9, 0.0)  # glossy blob
    inst = [Instance(i, i, inst_tf[i]) for i in range(25)]
    cam = look_at((-13.5, 3.2, 0.4), (6.0, 4.6, -0.3), fov=50.0)
    return Scene("sponza_262144", meshes, inst, mats, cam)


def tiny_scene(seed: int = 7) -> Scene:
    """A few hundred triangles, several instances sharing meshes, for fast tests."""
    hf = heightfield_mesh(8, 8, seed, cell=0.5, height=0.6)
    ball = _sphere(12, 8, 0.5, seed + 1, bump=0.1)
    box = _open_box(0.8, 0.8, 0.8)
    light = _quad((-1.0, 3.0, -1.0), (1.0, 3.0, -1.0), (1.0, 3.0, 1.0), (-1.0, 3.0, 1.0))
    mats = [
        Material((0.6, 0.6, 0.55), 0.9, 0.0, 0.0),
        Material((0.8, 0.3, 0.2), 0.4, 0.5, 0.0),
        Material((0.9, 0.9, 0.9), 0.1, 0.9, 0.0),
        Material((1.0, 1.0, 0.9), 1.0, 0.0, 8.0),
    ]
    inst = [
        Instance(0, 0, mat_identity()),
        Instance(1, 1, mat_trs((1.0, 1.2, 1.5))),
        Instance(1, 2, mat_trs((2.6, 1.0, 2.4), yaw_deg=40.0, s=0.7)),  # same mesh: takes material 1 (reference quirk)
        Instance(2, 2, mat_trs((3.0, 0.55, 1.0), yaw_deg=25.0)),
        Instance(3, 3, mat_trs((2.0, 0.0, 2.0))),
    ]
    cam = look_at((2.0, 2.4, -2.2), (2.0, 0.8, 2.0), fov=55.0)
    return Scene("tiny", [hf, ball, box, light], inst, mats, cam)


# ----------------------------------------------------------------------------
# config 4: breakfast-room-shaped scene, 1,048,576 triangles, emissive panels + glossy tableware
# ----------------------------------------------------------------------------
def breakfast_scene(seed: int = 4242) -> Scene:
    """Closed room (10 x 4 x 8 m) with two ceiling light panels, a long table with glossy
    (metallic 0.9, roughness 0.1) tableware, eight chairs and a blinds wall.  17 meshes,
    12 materials, 1,048,576 triangles."""
    X, Y, Z = 10.0, 4.0, 8.0
    meshes: List[Mesh] = []
    inst: List[Instance] = []

    def add(mesh: Mesh, material: int, tf=None):
        meshes.append(mesh)
        inst.append(Instance(len(meshes) - 1, material, mat_identity() if tf is None else tf))

    add(_plane_xz(256, 256, -X / 2, X / 2, -Z / 2, Z / 2, 0.0, seed + 1, 0.004, up=True), 0)  # floor 131072
    add(_plane_xz(128, 128, -X / 2, X / 2, -Z / 2, Z / 2, Y, seed + 2, 0.004, up=False), 1)  # ceiling 32768
    add(_wall_xy(256, 128, -X / 2, X / 2, 0.0, Y, Z / 2, seed + 3, 0.01, face_neg_z=True), 1)  # 65536
    add(_wall_xy(256, 128, -X / 2, X / 2, 0.0, Y, -Z / 2, seed + 4, 0.01, face_neg_z=False), 1)  # 65536
    add(_wall_zy(256, 128, -Z / 2, Z / 2, 0.0, Y, X / 2, seed + 5, 0.01, face_neg_x=True), 2)  # 65536
    add(_wall_zy(256, 128, -Z / 2, Z / 2, 0.0, Y, -X / 2, seed + 6, 0.01, face_neg_x=False), 2)  # 65536
    # two emissive ceiling panels, 64 x 64 quads each -> 16384 in one mesh
    panels = merge([_plane_xz(64, 64, x0, x0 + 2.0, -1.5, 1.5, Y - 0.05, seed + 7 + i, 0.0, up=False) for i, x0 in enumerate((-3.5, 1.5))])
    add(panels, 3)
    # table top (both faces) 2 x 256 x 128 quads -> 131072, four legs 64 x 64 -> 32768
    top = merge([_plane_xz(256, 128, -3.0, 3.0, -1.0, 1.0, 0.95, seed + 10, 0.002, up=True), _plane_xz(256, 128, -3.0, 3.0, -1.0, 1.0, 0.90, seed + 11, 0.002, up=False)])
    add(top, 4)
    legs = merge([translate(_cylinder(64, 64, 0.06, 0.0, 0.9, seed + 12 + i, flute=0.0), (x, 0.0, z)) for i, (x, z) in enumerate([(-2.8, -0.85), (2.8, -0.85), (-2.8, 0.85), (2.8, 0.85)])])
    add(legs, 5)
    # glossy tableware: teapot-ish blob 256 x 128 -> 65536; 8 cups 64 x 32 -> 32768; 8 plates (squashed spheres) 64 x 32 -> 32768
    add(_sphere(256, 128, 0.28, seed + 20, bump=0.06, squash=0.85), 6, mat_trs((0.0, 1.22, 0.0), yaw_deg=20.0))
    seats = [(-2.25 + 1.5 * i, s) for s in (-1, 1) for i in range(4)]
    add(merge([translate(_sphere(64, 32, 0.07, seed + 30 + i, bump=0.0, squash=1.3), (x, 1.05, 0.55 * s)) for i, (x, s) in enumerate(seats)]), 7)
    add(merge([translate(_sphere(64, 32, 0.16, seed + 40 + i, bump=0.0, squash=0.08), (x, 0.97, 0.62 * s)) for i, (x, s) in enumerate(seats)]), 8)
    # eight chairs: seat + back as displaced sheets, 2 x 64 x 64 quads each chair -> 131072, instanced from one mesh per side
    def chair(k):
        seat = _plane_xz(64, 64, -0.25, 0.25, -0.25, 0.25, 0.5, seed + 50 + k, 0.004, up=True)
        back = _wall_xy(64, 64, -0.25, 0.25, 0.5, 1.1, 0.25, seed + 60 + k, 0.004, face_neg_z=True)
        return merge([seat, back])

    add(merge([translate(chair(i), (x, 0.0, 1.45)) for i, (x, s) in enumerate(seats[:4])]), 9)  # 65536
    add(merge([translate(chair(4 + i), (x, 0.0, 1.45)) for i, (x, s) in enumerate(seats[:4])]), 9, mat_trs((0.0, 0.0, 0.0), yaw_deg=180.0))  # 65536
    # blinds: 32 slats of 64 x 32 quads in front of the +x wall -> 131072
    slats = []
    for i in range(32):
        y = 0.6 + 0.09 * i

        def fn(u, v, iu, iv, y=y, i=i):
            return X / 2 - 0.15 - 0.06 * v, y + 0.04 * v + _relief(seed + 70 + i, iu, iv, 0.001), -3.0 + 6.0 * u

        slats.append(grid_mesh(64, 32, fn, flip=False))
    add(merge(slats), 10)
    # sideboard: a bumpy box-like blob 128 x 64 -> 16384, and a vase 128 x 128 -> 32768, lamp globe 128 x 64 -> 16384
    add(_sphere(128, 64, 0.5, seed + 80, bump=0.03, squash=0.7), 11, mat_trs((-4.0, 0.45, -3.0), s=(2.0, 1.0, 0.8)))
    add(_sphere(128, 128, 0.22, seed + 81, bump=0.04, squash=1.8), 6, mat_trs((-4.0, 1.25, -3.0)))
    assert sum(meshes[i.mesh].n_tris for i in inst) == 1 << 20
    mats = [
        Material((0.55, 0.42, 0.30), 0.6, 0.0, 0.0),  # 0 floor wood
        Material((0.85, 0.85, 0.82), 0.95, 0.0, 0.0),  # 1 plaster
        Material((0.70, 0.74, 0.80), 0.9, 0.0, 0.0),  # 2 painted walls
        Material((1.0, 0.96, 0.88), 1.0, 0.0, 12.0),  # 3 emissive panels
        Material((0.80, 0.80, 0.78), 0.25, 0.0, 0.0),  # 4 table top
        Material((0.30, 0.30, 0.32), 0.3, 0.9, 0.0),  # 5 legs, brushed metal
        Material((0.95, 0.93, 0.88), 0.1, 0.9, 0.0),  # 6 glossy teapot / vase
        Material((0.90, 0.85, 0.70), 0.1, 0.9, 0.0),  # 7 cups
        Material((0.92, 0.92, 0.95), 0.1, 0.9, 0.0),  # 8 plates
        Material((0.45, 0.30, 0.20), 0.7, 0.0, 0.0),  # 9 chairs
        Material((0.88, 0.88, 0.85), 0.5, 0.5, 0.0),  # 10 blinds
        Material((0.35, 0.25, 0.20), 0.4, 0.0, 0.0),  # 11 sideboard
    ]
    cam = look_at((-4.3, 1.9, 3.3), (0.3, 0.9, -0.3), fov=50.0)
    return Scene("breakfast_1048576", meshes, inst, mats, cam)


# ----------------------------------------------------------------------------
# config 5: animated 5M-triangle sheet; vertex positions are a function of time
# ----------------------------------------------------------------------------
def animated_sheet(t: float, nu: int = 1600, nv: int = 1600, seed: int = 555, with_normals: bool = True) -> Mesh:
    """nu x nv quads (default 5,120,000 triangles) over a 40 x 40 m square: travelling waves
    plus fixed per-vertex relief (so no two centroids coincide).  Seen from above."""

    def fn(u, v, iu, iv):
        y = 0.6 * np.sin(6.0 * np.pi * u + 1.7 * t) * np.cos(4.0 * np.pi * v - 1.1 * t) + 0.25 * np.sin(18.0 * np.pi * (u + v) + 2.3 * t)
        return 40.0 * (u - 0.5), y + _relief(seed, iu, iv, 0.01), 40.0 * (v - 0.5)

    if with_normals:
        return grid_mesh(nu, nv, fn, flip=True)
    # positions only (the per-frame path of a caller that animates vertices): analytic normals are not needed by the BVH
    iu, iv = np.meshgrid(np.arange(nu + 1), np.arange(nv + 1), indexing="ij")
    x, y, z = fn(iu.astype(np.float64) / nu, iv.astype(np.float64) / nv, iu, iv)
    return Mesh(_round6(np.stack([x, y, z], axis=-1).reshape(-1, 3)), None, None)


def animated_scene(nu: int = 1600, nv: int = 1600, seed: int = 555) -> Scene:
    sheet = animated_sheet(0.0, nu, nv, seed)
    lamp = _plane_xz(8, 8, -6.0, 6.0, -6.0, 6.0, 9.0, seed + 1, 0.0, up=False)
    mats = [Material((0.55, 0.65, 0.75), 0.5, 0.5, 0.0), Material((1.0, 0.95, 0.9), 1.0, 0.0, 10.0)]
    inst = [Instance(0, 0, mat_identity()), Instance(1, 1, mat_identity())]
    cam = look_at((-16.0, 7.5, -16.0), (2.0, 0.0, 2.0), fov=50.0)
    return Scene(f"animated_{sheet.n_tris}", [sheet, lamp], inst, mats, cam)
