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


def grid_with_coincident_cluster(n_grid=40, n_cluster=6):
    """A regular grid of 2 * n_grid^2 well-separated triangles plus, far away, a cluster of > 4 triangles with identical
    centroids: the first splits are fine, the degenerate set surfaces only several levels down -- inside the parity
    builder's subtree phase (one block per item of <= 1024 triangles), which has to report it like the level loop does."""
    from caveengine_b200 import scenes

    pos, idx = [], []
    for j in range(n_grid):
        for i in range(n_grid):
            x, z = float(i), float(j)
            base = len(pos)
            pos += [[x, 0, z], [x + 0.9, 0, z], [x, 0, z + 0.9], [x + 0.9, 0.1, z + 0.9]]
            idx += [base, base + 1, base + 2, base + 1, base + 3, base + 2]
    tri = np.array([[1000, 0, 0], [1001, 0, 0], [1000, 1, 0]], dtype=np.float32)
    for _ in range(n_cluster):
        base = len(pos)
        pos += tri.tolist()
        idx += [base, base + 1, base + 2]
    pos = np.array(pos, dtype=np.float32)
    return scenes.Mesh(pos, np.tile(np.array([0, 1, 0], dtype=np.float32), (pos.shape[0], 1)), np.array(idx, dtype=np.uint32))


def random_mesh(seed, n_tris):
    """Seeded triangle soup for builder fuzzing: a few clusters of small triangles of random size, a tenth of them
    axis-aligned (zero-extent boxes: the MakeValid padding), coordinates snapped to a coarse lattice so that equal keys,
    equal centroids components and +0 / -0 turn up; never > 4 identical centroids (the reference traps on those)."""
    from caveengine_b200 import scenes

    rng = np.random.default_rng(seed)
    n_clusters = int(rng.integers(1, 6))
    centres = rng.uniform(-20, 20, size=(n_clusters, 3))
    c = centres[rng.integers(0, n_clusters, size=n_tris)] + rng.normal(0, 3.0, size=(n_tris, 3))
    size = rng.uniform(0.05, 1.5, size=(n_tris, 1, 1))
    tri = c[:, None, :] + rng.normal(0, 1.0, size=(n_tris, 3, 3)) * size
    flat = rng.random(n_tris) < 0.1
    axis = rng.integers(0, 3, size=n_tris)
    for k in np.nonzero(flat)[0]:
        tri[k, :, axis[k]] = tri[k, 0, axis[k]]
    tri = np.round(tri * 8.0) / 8.0  # lattice: ties and exact zeros
    neg = rng.random(tri.shape) < 0.05
    tri = np.where((tri == 0.0) & neg, -0.0, tri).astype(np.float32)
    # de-duplicate centroids so no set of > 4 triangles shares one (nudge repeats along x by distinct lattice steps)
    cent = np.round(tri.mean(axis=1) * 24.0).astype(np.int64)
    _, inv, counts = np.unique(cent, axis=0, return_inverse=True, return_counts=True)
    seen = {}
    for k in range(n_tris):
        j = int(inv[k])
        r = seen.get(j, 0)
        seen[j] = r + 1
        if r >= 3:
            tri[k, :, 0] += np.float32(0.375 * (r - 2) + 50.0 * (1 + j % 7))
    pos = tri.reshape(-1, 3).astype(np.float32)
    return scenes.Mesh(pos, np.tile(np.array([0, 1, 0], dtype=np.float32), (pos.shape[0], 1)), np.arange(pos.shape[0], dtype=np.uint32))


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


def lbvh_chain_mesh(cluster=2048, seed=11):
    """Geometry whose binary radix tree over 30-bit Morton codes is deeper than the traversal stack:
    one tiny triangle in each of the cells (2^j,0,0), (0,2^j,0), (0,0,2^j), j = 0..9, of the 1024^3 Morton
    grid (a chain of 30 one-sided splits) and `cluster` triangles inside cell (0,0,0) (equal codes, split by
    index bits: +log2(cluster) levels).  Forces the LBVH builder's balanced fallback (wbvh_body.h build_lbvh)."""
    from caveengine_b200 import scenes

    rng = np.random.default_rng(seed)
    centres = [rng.uniform(0.05, 0.45, size=(cluster, 3))]
    for j in range(10):
        for axis in range(3):
            c = np.full(3, 0.25)
            c[axis] = float(1 << j) + 0.25
            centres.append(c[None, :])
    centres.append(np.full((1, 3), 1024.0))  # pins the centroid bounds: one Morton cell == one unit
    c = np.concatenate(centres).astype(np.float32)
    n = c.shape[0]
    tri = np.array([[-0.01, 0.0, -0.01], [0.0, 0.0, 0.012], [0.012, 0.0, -0.008]], dtype=np.float32)  # faces +y
    pos = (c[:, None, :] + tri[None, :, :]).reshape(-1, 3).astype(np.float32)
    nrm = np.tile(np.array([0, 1, 0], dtype=np.float32), (3 * n, 1))
    return scenes.Mesh(pos, nrm, np.arange(3 * n, dtype=np.uint32))


def lbvh_chain_scene(cluster=2048):
    from caveengine_b200 import scenes

    mesh = lbvh_chain_mesh(cluster)
    floor = scenes._quad((-8.0, -0.5, -8.0), (-8.0, -0.5, 40.0), (40.0, -0.5, 40.0), (40.0, -0.5, -8.0))
    mats = [scenes.Material((0.8, 0.7, 0.6), 0.8, 0.0, 2.0), scenes.Material((0.6, 0.6, 0.6), 0.9, 0.0, 0.0)]
    inst = [scenes.Instance(0, 0, scenes.mat_identity()), scenes.Instance(1, 1, scenes.mat_identity())]
    return scenes.Scene("lbvh_chain", [mesh, floor], inst, mats, scenes.look_at((0.25, 1.2, 0.25), (0.25, 0.0, 0.26), fov=25.0))


def image_metrics(got, want):
    """The figures the image tolerances are stated in (all relative to the oracle image `want`):
    rel_rmse   per-pixel RMSE of rgb / mean radiance
    bias       |mean(got) - mean(want)| / mean(want)
    bad_1e4    fraction of pixels with a channel beyond 1e-4 * (|want| + 1e-3)
    far_5pct   fraction of pixels with a channel beyond 5 % * (|want| + 1e-3)
    exact      fraction of pixels whose rgb is bit-identical"""
    g = np.asarray(got)[..., :3].astype(np.float64)
    r = np.asarray(want)[..., :3].astype(np.float64)
    mean = float(r.mean())
    d = np.abs(g - r)
    return {
        "rel_rmse": float(np.sqrt((d ** 2).mean())) / mean,
        "bias": abs(float(g.mean()) - mean) / mean,
        "bad_1e4": float((d > 1e-4 * (np.abs(r) + 1e-3)).any(axis=-1).mean()),
        "far_5pct": float((d > 0.05 * (np.abs(r) + 1e-3)).any(axis=-1).mean()),
        "exact": float((g == r).all(axis=-1).mean()),
    }


def check_metrics(m, limits, what):
    """Every limit is an upper bound except `exact` (a lower bound).  The assert message carries the measured values.
    With $CAVEPT_PARITY_LOG set, the measured figures are appended to that file (one JSON line per check): this is
    how the stated bounds were derived (profiles/r02_parity_measured.jsonl)."""
    log = os.environ.get("CAVEPT_PARITY_LOG")
    if log:
        import json

        with open(log, "a") as f:
            f.write(json.dumps({"case": what, **m, "limits": limits}) + "\n")
    for k, lim in limits.items():
        ok = m[k] >= lim if k == "exact" else m[k] <= lim
        assert ok, f"{what}: {k} = {m[k]:.6g} violates the stated bound {lim:.6g}; measured {m}"


# ---- textured materials / environment map (cavept.h "textured materials ...") ----------------------------------
def texture_case(scene, seed=3):
    """Deterministic inputs of the textured path for `scene`: per-mesh planar uvs, two RGBA32F textures (a checker with
    transparent holes; a metal-rough map), per-material map indices and a small equirectangular environment map.
    Returns a dict; "uvs_global" / "maps_gpu" are the same data in the orders the emulator wants (scene-global vertex
    order = meshes in order of first use; GpuPtMaterial-slot order = first material of each used mesh, by mesh id)."""
    rng = np.random.default_rng(seed)
    uvs = [np.ascontiguousarray(np.stack([m.positions[:, 0] * 0.9 + m.positions[:, 1] * 0.31, m.positions[:, 2] * 0.8 - m.positions[:, 1] * 0.17], axis=1), dtype=np.float32)
           for m in scene.meshes]
    n = 16
    yy, xx = np.mgrid[0:n, 0:n]
    base = np.zeros((n, n, 4), dtype=np.float32)
    base[..., :3] = np.where(((xx // 2 + yy // 2) % 2 == 0)[..., None], np.array([0.85, 0.3, 0.25], dtype=np.float32), np.array([0.2, 0.5, 0.8], dtype=np.float32))
    base[..., 3] = np.where((xx % 8 < 2) & (yy % 8 < 2), 0.0, 1.0)  # holes: alpha cut-out
    mr = np.zeros((8, 8, 4), dtype=np.float32)
    mr[..., 1] = rng.uniform(0.1, 1.0, (8, 8))  # roughness in .g
    mr[..., 2] = rng.uniform(0.0, 0.9, (8, 8))  # metallic in .b
    mr[..., 3] = 1.0
    eh, ew = 8, 16
    env = np.zeros((eh, ew, 3), dtype=np.float32)
    env[..., 0] = np.linspace(0.1, 0.9, ew)[None, :]
    env[..., 1] = np.linspace(0.2, 0.6, eh)[:, None]
    env[..., 2] = 0.5
    nm = len(scene.materials)
    base_maps = np.array([0 if k % 2 == 0 else -1 for k in range(nm)], dtype=np.int32)
    mr_maps = np.array([1 if k % 3 != 2 else -1 for k in range(nm)], dtype=np.int32)
    first, used = {}, []
    for inst in scene.instances:
        if inst.mesh not in first:
            first[inst.mesh] = inst.material
            used.append(inst.mesh)
    slots = []
    for mesh_id in sorted(first):
        if first[mesh_id] not in slots:
            slots.append(first[mesh_id])
    return {"uvs": uvs, "textures": [base, mr], "base_maps": base_maps, "mr_maps": mr_maps, "env": env,
            "uvs_global": np.concatenate([uvs[m] for m in used]), "maps_gpu": (base_maps[slots], mr_maps[slots]), "slots": slots}


def apply_texture_case(pt, scene, case, maps=True, env=True):
    """Feeds a texture_case to a PathTracer through the C ABI."""
    if maps:
        for k in range(len(scene.meshes)):
            pt.set_mesh_uvs(k, case["uvs"][k])
        pt.upload_textures(case["textures"])
        pt.set_material_maps(case["base_maps"], case["mr_maps"])
    if env:
        pt.set_env_map(case["env"])
