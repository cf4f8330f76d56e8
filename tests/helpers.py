"""Shared helpers for the parity tests (TEST CODE)."""
from __future__ import annotations

import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def golden_cases(npz):
    return sorted({k.split("/")[0] for k in npz.files})


def sha(*arrays) -> bytes:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.digest()


class GoldenCamera:
    def __init__(self, v):
        self.position, self.fov_degree = tuple(v[0:3]), float(v[3])
        self.forward, self.right, self.up = tuple(v[4:7]), tuple(v[7:10]), tuple(v[10:13])


def golden_scene(npz, name):
    """AssembledScene + camera + params from image_golden.npz."""
    from oracle import oracle as O

    asm = O.AssembledScene(
        npz[f"{name}/vertices"].view(O.VERTEX_DT), npz[f"{name}/indices"].view(O.INDEX_DT), npz[f"{name}/bvhs"].view(O.BVH_DT),
        npz[f"{name}/meshes"].view(O.MESH_DT), npz[f"{name}/materials"].view(O.MATERIAL_DT), None)
    w, h, mb, f0, nf, segs = [int(x) for x in npz[f"{name}/params"]]
    return asm, GoldenCamera(npz[f"{name}/camera"]), dict(width=w, height=h, max_bounce=mb, first_frame=f0, n_frames=nf, segments=segs), npz[f"{name}/image"]


def image_mismatch(a, b, rel=1e-5, floor=1e-3):
    """Fraction of pixels whose rgb differs by more than rel * (|b| + floor); alpha must match exactly."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape
    assert np.array_equal(a[..., 3], b[..., 3]), "sample counts (alpha) differ"
    bad = (np.abs(a[..., :3] - b[..., :3]) > rel * (np.abs(b[..., :3]) + floor)).any(axis=-1)
    return float(bad.mean())


def rmse(a, b):
    a = np.asarray(a, dtype=np.float64)[..., :3]
    b = np.asarray(b, dtype=np.float64)[..., :3]
    return float(np.sqrt(np.mean((a - b) ** 2)))


def coincident_centroid_mesh(n=6):
    """> 4 triangles with identical centroids: the reference's builder traps on this
    (SURVEY.md section 8a note i; bvh_accel.cpp:138-141)."""
    from caveengine_b200 import scenes

    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32)
    pos = np.tile(tri, (n, 1))
    idx = np.arange(3 * n, dtype=np.uint32)
    return scenes.Mesh(pos, np.tile(np.array([0, 0, 1], dtype=np.float32), (3 * n, 1)), idx)


def isect_golden_inputs():
    """Inputs of isect_golden.npz in the probe's layout: rays (n,7) = origin, direction = end - start
    (fp32, as TestIntersection computes it), t = 1 (Ray::m_dist's initial value)."""
    g = load_golden("isect_golden.npz")
    start, end = g["rays"][:, 0:3], g["rays"][:, 3:6]
    rays = np.concatenate([start, (end - start).astype(np.float32), np.ones((start.shape[0], 1), dtype=np.float32)], axis=1).astype(np.float32)
    return g, rays


def check_probe_against_reference_twins(out, g):
    """SURVEY.md section 8 row a20.  `out` = (n,8) probe results of the product's HitTriangle / HitBvh.
    Triangle: same Moller-Trumbore in the same fp32 order -> hit flags and t bit-identical to
    TestIntersection::RayTriangle.  Box: same slab arithmetic, different accept rule -- the engine's CPU twin
    also rejects tmin <= 0 (origin inside or box behind), the shader's HitBvh does not -- so
    reference hit <=> (HitBvh hit and tmin > 0), and on a hit m_dist == tmin bit for bit."""
    rt, rb = g["ref_triangle"], g["ref_aabb"]
    assert np.array_equal(out[:, 0], rt[:, 0]), f"{int((out[:, 0] != rt[:, 0]).sum())} triangle hit flags differ"
    assert out[:, 1].tobytes() == rt[:, 1].tobytes(), "triangle t differs from the reference twin"
    assert rt[:, 0].sum() > 100 and (rt[:, 0] == 0).sum() > 100
    ours_box = (out[:, 4] == 1.0) & (out[:, 5] > 0.0)
    assert np.array_equal(ours_box, rb[:, 0] == 1.0), f"{int((ours_box != (rb[:, 0] == 1.0)).sum())} box decisions differ"
    hit = rb[:, 0] == 1.0
    assert out[hit, 5].tobytes() == rb[hit, 1].tobytes(), "slab tmin differs from the reference twin"
    assert hit.sum() > 100
