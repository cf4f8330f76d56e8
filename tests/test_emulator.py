"""CPU tests of the PRODUCT's algorithms: the kernel bodies under caveengine_b200/csrc are
stepped serially by tests/emu (no GPU here) and compared with the reference-generated golden
vectors and the oracle.  The -m gpu tests repeat these through the real kernels."""
import numpy as np
import pytest

import emu_loader as E
import helpers as H
from caveengine_b200 import scenes


def _single_mesh_scene(mesh):
    return scenes.Scene("case", [mesh], [scenes.Instance(0, 0, scenes.mat_identity())], [scenes.Material((0.5, 0.5, 0.5), 0.5, 0.0, 0.0)], scenes.look_at((0, 1, -3), (0, 0, 0)))


def test_level_synchronous_builder_is_bit_exact_vs_reference_golden(port_oracle):
    g = H.load_golden("bvh_golden.npz")
    for name in H.golden_cases(g):
        mesh = scenes.Mesh(g[f"{name}/positions"], np.zeros_like(g[f"{name}/positions"]), g[f"{name}/indices"])
        sc = _single_mesh_scene(mesh)
        asm = port_oracle.assemble(sc)
        emu = E.EmuScene(asm, sc)
        assert emu.rc == 0, name
        assert emu.nodes.tobytes() == g[f"{name}/nodes"].tobytes(), f"{name}: differs from the reference builder"


@pytest.mark.parametrize("sub_max", [0, 1, 7, 64, 100_000])
def test_subtree_phase_of_the_parity_builder_emits_the_same_bytes(port_oracle, scene_cache, sub_max):
    """The parity builder's second phase (one block per queued item, bvh_parity_body.h parity_subtree_*) at several
    hand-over sizes -- level loop only, hand-over at 1 / 7 / 64 triangles, every root's children at once -- against the
    reference builder's bytes: golden meshes (signed zeros, ties, <= 4-triangle nodes) and a multi-mesh scene."""
    g = H.load_golden("bvh_golden.npz")
    for name in H.golden_cases(g):
        mesh = scenes.Mesh(g[f"{name}/positions"], np.zeros_like(g[f"{name}/positions"]), g[f"{name}/indices"])
        sc = _single_mesh_scene(mesh)
        emu = E.EmuScene(port_oracle.assemble(sc), sc, parity_sub_max=sub_max)
        assert emu.rc == 0, name
        assert emu.nodes.tobytes() == g[f"{name}/nodes"].tobytes(), f"{name}: differs from the reference builder at sub_max={sub_max}"
    sc = scene_cache("heightfield")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc, parity_sub_max=sub_max)
    assert emu.rc == 0 and emu.nodes.tobytes() == asm.bvhs.tobytes()
    ref = E.EmuScene(asm, sc, parity_sub_max=0)
    assert np.array_equal(emu.parent, ref.parent) and np.array_equal(emu.count, ref.count)


@pytest.mark.parametrize("name", ["tiny", "cornell", "heightfield"])
def test_multi_mesh_build_matches_oracle_assembly(port_oracle, scene_cache, name):
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    assert emu.rc == 0
    assert emu.nodes.tobytes() == asm.bvhs.tobytes()  # includes AppendBvhs offsets (path_tracer.cpp:90-118)
    # subtree sizes: every subtree over n triangles has 2n-1 nodes
    assert np.all(emu.count >= 1)
    leaf = asm.bvhs["leaf"] == 1
    assert np.array_equal(emu.count == 1, leaf)


def test_degenerate_geometry_reports_error(port_oracle):
    sc = _single_mesh_scene(H.coincident_centroid_mesh(6))
    asm_v = np.zeros(18, dtype=port_oracle.assemble.__globals__["VERTEX_DT"])
    asm_v["position"] = sc.meshes[0].positions
    asm_i = np.zeros(6, dtype=port_oracle.assemble.__globals__["INDEX_DT"])
    asm_i["tri"] = sc.meshes[0].indices.reshape(-1, 3)

    class Asm:
        vertices, indices, meshes, materials = asm_v, asm_i, None, None

    emu = E.EmuScene(Asm, sc)
    assert emu.rc != 0 and emu.error_bits != 0


@pytest.mark.parametrize("sub_max", [0, 16, 1024])
def test_random_meshes_build_to_the_reference_bytes_in_both_phases(best_oracle, sub_max):
    """Seeded triangle soups (clusters, zero-extent boxes, lattice ties, signed zeros; 1 .. 3000 triangles) through the
    level loop only, a hand-over at 16 triangles and the product's hand-over: the reference builder's bytes each time."""
    for seed in range(80):
        n = [1, 2, 5, 17, 100, 700, 1500, 3000, 2049, 1025][seed % 10]
        mesh = H.random_mesh(seed, n)
        sc = _single_mesh_scene(mesh)
        want = best_oracle.bvh_build(mesh.indices, mesh.positions)
        emu = E.EmuScene(best_oracle.assemble(sc), sc, parity_sub_max=sub_max)
        assert emu.rc == 0, (seed, n)
        assert emu.nodes.tobytes() == want.tobytes(), f"seed {seed} ({n} triangles), sub_max {sub_max}"


@pytest.mark.parametrize("sub_max", [0, 64, 1024])
def test_degenerate_cluster_deep_in_a_mesh_reports_error_from_either_phase(port_oracle, sub_max):
    """The coincident-centroid set sits several levels down: with sub_max >= 64 it is met inside a subtree build."""
    sc = _single_mesh_scene(H.grid_with_coincident_cluster())
    mesh = sc.meshes[0]
    asm_v = np.zeros(mesh.positions.shape[0], dtype=port_oracle.assemble.__globals__["VERTEX_DT"])
    asm_v["position"] = mesh.positions
    asm_i = np.zeros(mesh.indices.size // 3, dtype=port_oracle.assemble.__globals__["INDEX_DT"])
    asm_i["tri"] = mesh.indices.reshape(-1, 3)

    class Asm:  # the oracle refuses to assemble this mesh (its builder reports the same degenerate set)
        vertices, indices, meshes, materials = asm_v, asm_i, None, None

    with pytest.raises(ValueError):
        port_oracle.assemble(sc)
    emu = E.EmuScene(Asm, sc, parity_sub_max=sub_max)
    assert emu.rc != 0 and (emu.error_bits & 2), (emu.rc, emu.error_bits)  # BUILD_ERR_EMPTY_SIDE
    ok = _single_mesh_scene(H.grid_with_coincident_cluster(n_cluster=4))  # <= 4 triangles: SplitByAxis, no buckets, fine
    emu = E.EmuScene(port_oracle.assemble(ok), ok, parity_sub_max=sub_max)
    assert emu.rc == 0


@pytest.mark.parametrize("name", ["tiny_mb5", "cornell_mb5", "tiny_mb8", "tiny_mb1"])
@pytest.mark.parametrize("fast", [False, True])
def test_traversals_match_reference_golden_images(name, fast):
    g = H.load_golden("image_golden.npz")
    asm, cam, p, want = H.golden_scene(g, name)
    # rebuild the scene topology the product way from the golden arrays
    n_inst = asm.meshes.shape[0]
    roots = sorted(set(int(r) for r in asm.meshes["root"]))
    tri_of_root = {}
    for r in roots:  # first triangle of each mesh = smallest triangleIndex reachable is its offset
        tri_of_root[r] = None

    class FakeMesh:
        def __init__(self, n):
            self.n_tris = n

    # mesh triangle counts from the node ranges: mesh k owns nodes [root_k, root_{k+1})
    bounds = roots + [asm.bvhs.shape[0]]
    meshes = [FakeMesh((bounds[k + 1] - bounds[k] + 1) // 2) for k in range(len(roots))]
    instances = [scenes.Instance(roots.index(int(asm.meshes["root"][i])), 0, asm.meshes["transform"][i]) for i in range(n_inst)]
    # first-appearance order must equal root order for the offsets to line up
    order = []
    for i in instances:
        if i.mesh not in order:
            order.append(i.mesh)
    assert order == sorted(order)
    sc = scenes.Scene("golden", meshes, instances, [], cam)
    emu = E.EmuScene(asm, sc)
    assert emu.rc == 0
    assert emu.nodes.tobytes() == asm.bvhs.tobytes()
    img, st = emu.render(cam, p["width"], p["height"], p["first_frame"], p["n_frames"], p["max_bounce"], fast=fast)
    # fp32-pure arithmetic vs the header-as-C++ double promotions can flip a rare path
    assert abs(st["segments"] - p["segments"]) <= max(2, p["segments"] // 1000)
    if fast:  # world-space triangles: transformed instances differ by fp32 rounding of the transform
        assert H.image_mismatch(img, want, rel=1e-4) <= 0.01
    else:
        assert H.image_mismatch(img, want, rel=1e-5) <= 0.002


def test_fast_traversal_equals_reference_walk_bitwise_for_identity_instances(port_oracle, scene_cache):
    """World-space triangles of an identity instance are the mesh-space triangles, so the
    production traversal returns HitScene's (t, u, v) bit for bit."""
    sc = scene_cache("heightfield")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=False)
    b, sb = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=True)
    assert sa["segments"] == sb["segments"]
    assert H.image_mismatch(b, a, rel=0.0, floor=0.0) <= 0.001  # exact except exact-tie pixels
    assert sb["node_visits"] < sa["node_visits"]
    oracle_img, so = port_oracle.render(asm, sc.camera, 64, 36, 0, 3, 5)
    assert so["segments"] == sa["segments"] and so["node_visits"] == sa["node_visits"] and so["tri_tests"] == sa["tri_tests"]
    assert a.tobytes() == oracle_img.tobytes()


def test_fast_traversal_matches_reference_walk_with_transformed_instances(port_oracle, scene_cache):
    """Transformed instances: the reference intersects in mesh space with the inverse-transformed
    ray, the production tree in world space -- equal up to fp32 rounding of the transform."""
    sc = scene_cache("tiny")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=False)
    b, sb = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=True)
    assert abs(sa["segments"] - sb["segments"]) <= max(2, sa["segments"] // 500)
    assert H.image_mismatch(b, a, rel=1e-4) <= 0.01
    assert sb["node_visits"] < sa["node_visits"]


def _world_tree_check(emu, asm, sc, holes_ok=False):
    """Walks the production tree: every (instance, triangle) pair sits in exactly one leaf, every
    pair of sibling records is referenced once, every box contains what hangs below it."""
    P = emu.n_prims
    assert emu.wrc == 0 and emu.wb_error_bits == 0
    fn = emu.fnodes
    seen_slots = np.zeros(P, dtype=np.int32)
    seen_pairs = np.zeros(max(emu.n_pairs, 1), dtype=np.int32)
    a0 = emu.ftris["v0"][:, :3]  # records hold A, AB = B - A, AC = C - A (world space)
    verts = np.stack([a0, a0 + emu.ftris["v1"][:, :3], a0 + emu.ftris["v2"][:, :3]], axis=1)  # (P, 3, 3)
    max_depth = 0

    def visit(ref, depth, box):
        nonlocal max_depth
        max_depth = max(max_depth, depth)
        if ref < 0:
            first, count = ref & 0x1FFFFFFF, ((ref >> 29) & 3) + 1
            seen_slots[first:first + count] += 1
            if box is not None:
                v = verts[first:first + count].reshape(-1, 3)
                slack = 0.5 * emu.pad  # A + AB re-rounds B by <= 1 ulp; the boxes are padded by emu.pad
                assert np.all(v >= box[0] - slack) and np.all(v <= box[1] + slack)
            return count
        assert ref % 2 == 0
        seen_pairs[ref // 2] += 1
        total = 0
        for k in (0, 1):
            n = fn[ref + k]
            if box is not None:
                assert np.all(n["mn"] >= box[0]) and np.all(n["mx"] <= box[1])
            got = visit(int(n["a"]), depth + 1, (n["mn"], n["mx"]))
            assert got == int(n["b"])
            total += got
        return total

    assert visit(emu.root_ref, 0, None) == P
    # the quantised pairs the tracer fetches: each 15-bit grid box contains its fp32 box, and
    # the references are the fp32 records' (interior: index of the child's own pair)
    org, scl = emu.grid[:3].astype(np.float64), emu.grid[3:].astype(np.float64)
    for k in range(emu.n_pairs):
        if seen_pairs[k] == 0 and holes_ok:
            continue  # heap slot of a leaf position in the implicit LBVH tree: never referenced
        q = emu.qpairs[k]
        for s_ in (0, 1):
            w = [int(v) for v in q[3 * s_: 3 * s_ + 3]]
            lo = np.array([v & 0xFFFF for v in w], dtype=np.float64)  # per axis: min | max << 16
            hi = np.array([v >> 16 for v in w], dtype=np.float64)
            assert np.all(hi <= 32767) and np.all(lo >= 0)
            n = fn[2 * k + s_]
            assert np.all(org + lo / scl <= n["mn"].astype(np.float64)) and np.all(org + hi / scl >= n["mx"].astype(np.float64))
            a, r = int(n["a"]), int(np.int32(q[6 + s_]))
            assert r == (a >> 1 if a >= 0 else a)
    assert np.all(seen_slots == 1)
    assert np.all(seen_pairs[: emu.n_pairs] <= 1) if holes_ok else np.all(seen_pairs[: emu.n_pairs] == 1)
    assert max_depth <= emu.fast_levels
    # slots carry (instance, triangle): together they enumerate every pair once
    inst = emu.ftris["v0"][:, 3].view(np.int32)
    tri = emu.ftris["v1"][:, 3].view(np.int32)
    want = []
    for i, ins in enumerate(sc.instances):
        t0 = int(emu.inst_tri_off[i])
        want += [(i, t0 + k) for k in range(sc.meshes[ins.mesh].n_tris)]
    assert sorted(zip(inst.tolist(), tri.tolist())) == sorted(want)
    # world-space vertices = instance transform * mesh-space vertex; normals stay in mesh space
    k = np.arange(P)
    m = asm.meshes["transform"][inst].reshape(P, 4, 4)  # column-major: m[p, c, r]
    local = asm.vertices["position"][asm.indices["tri"][tri]]  # (P, 3, 3)
    world = np.einsum("pcr,pvc->pvr", m[:, :3, :3], local) + m[:, None, 3, :3]
    flipped = np.linalg.det(m[:, :3, :3].astype(np.float64)) < 0
    world[flipped] = world[flipped][:, [0, 2, 1]]
    assert np.allclose(verts, world, rtol=1e-6, atol=1e-6)
    nrm = asm.vertices["normal"][asm.indices["tri"][tri]]
    nrm[flipped] = nrm[flipped][:, [0, 2, 1]]
    nrm_world = np.einsum("pcr,pvc->pvr", m[:, :3, :3], nrm)  # normals times the world matrix (w = 0), not normalised
    assert np.allclose(emu.fnormals["v1"][:, :3], nrm_world[:, 1], rtol=1e-6, atol=1e-7) and np.allclose(emu.fnormals["v2"][:, :3], nrm_world[:, 2], rtol=1e-6, atol=1e-7)
    identity = np.all(m == np.eye(4, dtype=np.float32)[None], axis=(1, 2))
    assert np.array_equal(emu.fnormals["v0"][identity][:, :3] + 0.0, nrm[identity][:, 0] + 0.0)  # bit-exact for identity instances (up to the sign of zero)
    assert np.array_equal(emu.fnormals["v0"][:, 3].view(np.int32), inst)
    assert np.array_equal(emu.fnormals["v1"][:, 3].view(np.int32), asm.meshes["material"][inst])
    return max_depth


@pytest.mark.parametrize("name", ["tiny", "cornell", "heightfield"])
def test_world_tree_invariants(port_oracle, scene_cache, name):
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    depth = _world_tree_check(emu, asm, sc)
    # the balanced (median-split) fallback used when a SAH tree would overflow the traversal stack
    built = []
    for inst in sc.instances:
        if inst.mesh not in built:
            built.append(inst.mesh)
    emu.build_world(asm, sc, built, force_median=True)
    depth_median = _world_tree_check(emu, asm, sc)
    assert depth_median <= int(np.ceil(np.log2(max(emu.n_prims, 2)))) + 1
    img_a, _ = emu.render(sc.camera, 32, 18, 0, 1, 3, fast=True)
    emu.build_world(asm, sc, built, force_median=False)
    img_b, _ = emu.render(sc.camera, 32, 18, 0, 1, 3, fast=True)
    assert H.image_mismatch(img_a, img_b, rel=0.0, floor=0.0) <= 0.01  # same closest hits whatever the tree
    assert depth >= 1


@pytest.mark.parametrize("name", ["tiny", "cornell", "heightfield"])
def test_lbvh_tree_invariants_and_same_hits(port_oracle, scene_cache, name):
    """The Morton-sort builder (CAVEPT_BVH_LBVH): same invariants, depth ceil(log2 N), and the
    same closest hits as the SAH tree (an image is a function of the hits only)."""
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    img_sah, _ = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    built = []
    for inst in sc.instances:
        if inst.mesh not in built:
            built.append(inst.mesh)
    emu.build_world(asm, sc, built, lbvh=True)  # binary radix tree
    depth = _world_tree_check(emu, asm, sc, holes_ok=True)
    assert depth == emu.fast_levels  # the height the bottom-up pass reports is the real one
    img_lbvh, _ = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    assert H.image_mismatch(img_lbvh, img_sah, rel=0.0, floor=0.0) <= 0.01  # exact-tie pixels only
    emu.build_world(asm, sc, built, lbvh=True, force_median=True)  # balanced fallback
    depth = _world_tree_check(emu, asm, sc, holes_ok=True)
    assert depth <= int(np.ceil(np.log2(max(emu.n_prims, 2)))) + 1
    img_bal, _ = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    assert H.image_mismatch(img_bal, img_sah, rel=0.0, floor=0.0) <= 0.01


@pytest.mark.parametrize("name", ["tiny", "cornell", "heightfield"])
def test_hybrid_tree_invariants_and_same_hits(port_oracle, scene_cache, name):
    """CAVEPT_BVH_HYBRID: the radix tree above, the SAH builder's subtree phase below (treelets of <= 768 prims rebuilt in
    the dead pair slots of their own range): the tree invariants hold with the pair array no larger than the LBVH's, and
    the closest hits are the SAH tree's."""
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    img_sah, st_sah = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    built = []
    for inst in sc.instances:
        if inst.mesh not in built:
            built.append(inst.mesh)
    emu.build_world(asm, sc, built, lbvh=True)
    _, st_lbvh = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    emu.build_world(asm, sc, built, hybrid=True)
    assert emu.wrc == 0 and emu.wb_error_bits == 0
    assert emu.n_pairs == emu.n_prims - 1
    depth = _world_tree_check(emu, asm, sc, holes_ok=True)
    assert depth <= emu.fast_levels
    img_h, st_h = emu.render(sc.camera, 40, 24, 0, 2, 4, fast=True)
    assert H.image_mismatch(img_h, img_sah, rel=0.0, floor=0.0) <= 0.01  # exact-tie pixels only
    # tree quality: no worse than the plain radix tree (node records per segment)
    assert st_h["node_visits"] <= 1.02 * st_lbvh["node_visits"], (st_h["node_visits"], st_lbvh["node_visits"], st_sah["node_visits"])


def test_world_tree_with_mirrored_instance(port_oracle):
    """det(transform) < 0 flips the winding the reference's mesh-space back-face test sees; the
    world-space triangles are stored (A, C, B) so the same faces are culled."""
    hf = scenes.heightfield_mesh(6, 6, 3, cell=0.5, height=0.4)
    mirror = scenes.mat_trs((0.0, 0.0, 0.0))
    mirror[0] = -1.0  # x -> -x
    sc = scenes.Scene("mirror", [hf], [scenes.Instance(0, 0, mirror)], [scenes.Material((0.7, 0.7, 0.7), 0.8, 0.0, 0.0)],
                      scenes.look_at((-1.5, 2.5, -2.0), (-1.5, 0.0, 1.5), fov=50.0))
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    _world_tree_check(emu, asm, sc)
    a, sa = emu.render(sc.camera, 48, 27, 0, 2, 4, fast=False)
    b, sb = emu.render(sc.camera, 48, 27, 0, 2, 4, fast=True)
    assert sa["segments"] > 48 * 27 * 2  # the surface is seen (not culled away)
    assert abs(sa["segments"] - sb["segments"]) <= max(2, sa["segments"] // 200)
    assert H.image_mismatch(b, a, rel=1e-4) <= 0.02


def test_inverse_matches_glm_golden():
    import ctypes as C

    lib = E.load()
    g = H.load_golden("inverse_golden.npz")
    for m, want in zip(g["mats"], g["inverses"]):
        out = np.zeros(16, dtype=np.float32)
        lib.emu_mat4_inverse(np.ascontiguousarray(m).ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert out.tobytes() == want.tobytes()


def test_intersection_bodies_match_the_engines_cpu_twins():
    """Row a20: the product's HitTriangle / HitBvh source (stepped on the host) against golden outputs of
    TestIntersection::RayTriangle / RayAabb (cave/engine/math/intersection.cpp:18-70, compiled in place)."""
    import ctypes as C

    import emu_loader

    g, rays = H.isect_golden_inputs()
    lib = emu_loader.load()
    n = rays.shape[0]
    out = np.zeros((n, 8), dtype=np.float32)
    tris, boxes = np.ascontiguousarray(g["tris"]), np.ascontiguousarray(g["boxes"])
    lib.emu_probe(C.c_int(n), rays.ctypes.data_as(C.c_void_p), tris.ctypes.data_as(C.c_void_p), boxes.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    H.check_probe_against_reference_twins(out, g)


@pytest.mark.parametrize("name,w,h", [("tiny", 61, 35), ("heightfield", 64, 36)])
def test_job_queue_bodies_equal_the_per_pixel_body_bitwise(port_oracle, scene_cache, name, w, h):
    """The persistent job-queue body (trace_pool_body.h: two pixel streams per lane, records parked in shared
    memory) runs the same walk_step / walk_test_triangle / shade_hit sequence per pixel as body_trace_pixel_fast:
    same accumulator bytes, same segment / node / triangle counts -- also on a ragged image (61 x 35 is not a
    multiple of the 8x4 job tile) and across two launches (progressive accumulation)."""
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, w, h, 0, 3, 5, fast=1)
    for body in (2, 3):  # 2: two pixel streams per lane (trace_pool_body.h); 3: the per-pixel body drawing jobs from the queue (k_trace_queue)
        b, sb = emu.render(sc.camera, w, h, 0, 3, 5, fast=body)
        assert a.tobytes() == b.tobytes(), body
        assert sa == sb, body
        a2, _ = emu.render(sc.camera, w, h, 3, 2, 5, fast=1, accum=a.copy(), reset_first=False)
        b2, _ = emu.render(sc.camera, w, h, 3, 2, 5, fast=body, accum=b.copy(), reset_first=False)
        assert a2.tobytes() == b2.tobytes() and np.all(b2[..., 3] == 5.0), body


@pytest.mark.timeout(600)
@pytest.mark.parametrize("name,w,h,n_warps", [("tiny", 61, 35, 1), ("tiny", 61, 35, 3), ("cornell", 64, 36, 2), ("tiny", 8, 4, 4), ("tiny", 200, 3, 2)])
def test_queue_kernel_on_real_warps_terminates_and_equals_the_per_pixel_body(port_oracle, scene_cache, name, w, h, n_warps):
    """k_trace_queue's body on warps of 32 host threads (tests/emu/emu.cpp, -DCV_WARP_EMU: __any_sync / __ballot_sync
    and the warp-aggregated job draw go through a barrier).  This is the test that sees the warp-level scheduling:
    lanes that finish their pixel after another lane closed the queue must retire, lanes that drew padding jobs of a
    ragged rectangle must draw again, and the loop must end -- with the same accumulator bytes and segment count as
    the per-pixel body, however many warps share the queue (more warps than jobs included: 8x4 image, 4 warps)."""
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, w, h, 0, 3, 5, fast=1)
    b, sb = emu.render_warps(sc.camera, w, h, 0, 3, 5, n_warps=n_warps)
    assert a.tobytes() == b.tobytes()
    assert sa["segments"] == sb["segments"]
    # a tile launch (TILED mode) over part of the image, continuing the accumulator
    tile = (8, 4, min(w, 29), min(h, 19))
    if tile[2] > tile[0] and tile[3] > tile[1]:
        b2, _ = emu.render_warps(sc.camera, w, h, 3, 2, 5, n_warps=n_warps, tile=tile, accum=b.copy(), reset_first=False)
        n = b2[..., 3]
        inside = np.zeros((h, w), dtype=bool)
        inside[tile[1]:tile[3], tile[0]:tile[2]] = True
        assert np.all(n[inside] == 5.0) and np.all(n[~inside] == 3.0)
        assert b2[~inside].tobytes() == b[~inside].tobytes()


@pytest.mark.parametrize("name,w,h,mb", [("tiny", 61, 35, 5), ("cornell", 64, 36, 5), ("heightfield", 48, 27, 5)])
def test_four_wide_walk_finds_the_same_hits_in_half_the_steps(port_oracle, scene_cache, name, w, h, mb):
    """CAVEPT_KERNEL_WIDE4: Q4Node k = sibling pair k with its interior children replaced by their pairs
    (wbvh_body.h wb_body_q4), walked by walk_step_wide (trace_body.h).  Same closest hits as the pair walk -> the same
    image and segment count; every record's child boxes contain the fp32 boxes of the nodes they were made from; the
    number of dependent node fetches (steps) per segment drops to about half while the boxes tested stay about equal."""
    sc = scene_cache(name)
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, w, h, 0, 2, mb, fast=1)
    b, sb = emu.render(sc.camera, w, h, 0, 2, mb, fast=4)
    assert sa["segments"] == sb["segments"]
    assert a.tobytes() == b.tobytes()
    steps_pair, steps_wide = sa["node_visits"] / 2, sb["node_visits"] / 4
    assert steps_wide <= 0.62 * steps_pair
    assert sb["node_visits"] <= 1.2 * sa["node_visits"]
    # structure: children of record k are a cut through the subtree of pair k
    q4 = emu.q4nodes[:emu.n_pairs]
    fn = emu.fnodes
    go, gs = emu.grid[:3], emu.grid[3:]
    for k in range(0, emu.n_pairs, max(1, emu.n_pairs // 200)):
        refs = q4[k, 12:].view(np.int32)
        live = [int(r) for r in refs if r != 0x7fffffff]
        assert 2 <= len(live) <= 4
        # expand pair k the same way on the host: the leaves below the record's children partition pair k's leaves
        def leaves(ref):  # ref in the quantised numbering (pair index / leaf ref)
            if ref < 0:
                return [ref]
            out = []
            for s in (0, 1):
                a_ = int(fn[2 * ref + s]["a"])
                out += leaves(a_ >> 1 if a_ >= 0 else a_)
            return out
        want = sorted(leaves(k))
        got = sorted(sum((leaves(r) for r in live), []))
        assert got == want


def test_refit_reproduces_the_tree_and_follows_moved_instances(port_oracle, scene_cache):
    """wbvh_body.h refit_wbvh (what cavept_update_instances runs instead of a rebuild): with nothing moved the
    bottom-up pass reproduces fnodes / qpairs / ftris bit for bit; with two instances moved, the refitted tree returns
    the same closest hits as a tree built from scratch on the moved scene (identical image), every box still contains
    what hangs below it, and the surface-area sum (the quality guard's proxy) is reported."""
    import copy

    from caveengine_b200 import scenes

    sc = scene_cache("tiny")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    before = (emu.fnodes.tobytes(), emu.qpairs.tobytes(), emu.ftris.tobytes(), emu.fnormals.tobytes())
    _, area0 = emu.refit(asm.meshes, np.zeros(len(sc.instances)))
    assert (emu.fnodes.tobytes(), emu.qpairs.tobytes(), emu.ftris.tobytes(), emu.fnormals.tobytes()) == before
    assert area0 > 0 and int(emu.level_items.sum()) == emu.n_pairs
    moved = copy.deepcopy(sc)
    moved.instances[1].transform = scenes.mat_trs((1.3, 1.4, 1.2), yaw_deg=12.0)  # both stay inside the scene box the tree's grid was laid over
    moved.instances[3].transform = scenes.mat_trs((2.8, 0.55, 1.4), yaw_deg=-40.0)
    asm2 = port_oracle.assemble(moved)
    box, area = emu.refit(asm2.meshes, [0, 1, 0, 1, 0])
    assert 0.8 * area0 < area < 1.3 * area0
    emu.asm = asm2
    _world_tree_check(emu, asm2, moved)
    a, sa = emu.render(moved.camera, 64, 36, 0, 2, 5, fast=1)
    fresh = E.EmuScene(asm2, moved)
    b, sb = fresh.render(moved.camera, 64, 36, 0, 2, 5, fast=1)
    assert sa["segments"] == sb["segments"]
    assert H.image_mismatch(a, b, rel=0.0, floor=0.0) <= 0.001
    # the scene box the host checks against the quantisation grid is the union of the root pair
    lo = np.minimum(emu.fnodes["mn"][0], emu.fnodes["mn"][1])
    assert np.array_equal(box[:3], lo)

    def fits_grid(bx):  # cavept_api.cu refit_world: outward rounding must not meet the clamp at 0 / 32767
        g_lo, g_hi = (bx[:3] - emu.grid[:3]) * emu.grid[3:], (bx[3:] - emu.grid[:3]) * emu.grid[3:]
        return bool(np.all(g_lo >= 1.0) and np.all(g_hi <= 32766.0))

    assert fits_grid(box)
    out = copy.deepcopy(sc)
    out.instances[1].transform = scenes.mat_trs((0.4, 1.5, 1.2), yaw_deg=12.0)  # pokes out of the old scene box: the host must rebuild
    box_out, _ = emu.refit(port_oracle.assemble(out).meshes, [0, 1, 0, 0, 0])
    assert not fits_grid(box_out)


def _far_camera(sc, mult):
    """Camera `mult` scene extents away from the scene's centre, looking at it, the view narrowed so the scene still fills
    the image (the shader's image-plane distance is tan(fov), path_tracer.cs.glsl:31-32)."""
    pos = np.concatenate([m.positions.reshape(-1, 3) for m in sc.meshes])
    c = 0.5 * (pos.min(0) + pos.max(0))
    ext = float((pos.max(0) - pos.min(0)).max())
    d = np.array([0.3, 0.5, -1.0])
    eye = c + d / np.linalg.norm(d) * ext * mult
    return scenes.look_at(eye, c, float(np.degrees(np.arctan(1.2 * mult))))


@pytest.mark.parametrize("mult", [1, 30, 100, 1000])
def test_walk_loses_no_hit_with_the_camera_far_outside_the_scene(port_oracle, scene_cache, mult):
    """the quantised slab test measures its planes from the ray origin, so its absolute error grows with the
    camera's distance from the grid while the boxes' outward margin is fixed.  Ground truth = every triangle tested
    with the kernel's own triangle arithmetic (no tree): up to 1000 scene extents away the production walk returns the
    same closest t for every pixel.  (At 30,000 extents ~5 % of the hits differ -- and differ just as much with the
    rays advanced to the scene box in double precision first, which was built, measured and removed: at that distance
    the reference's fp32 triangle test itself is rounding noise, `origin - A` cancels 15 of its 24 bits.)"""
    sc = scene_cache("heightfield")
    emu = E.EmuScene(port_oracle.assemble(sc), sc)
    cam = _far_camera(sc, mult)
    want_slot, want_t = emu.primary_hits(cam, 128, 72, 0, mode=0)
    got_slot, got_t = emu.primary_hits(cam, 128, 72, 0, mode=1)
    assert (want_slot >= 0).sum() > 1000  # the scene is in view
    assert np.array_equal(got_t, want_t)
    assert np.array_equal(got_slot >= 0, want_slot >= 0)


# ---- the reference's switched-off textured RayColor (glsl/path_tracer.cs.glsl:62-112), as built: parity unpinned, so the
#      tests pin it where it must coincide with the live shader, and against a numpy statement of the sampler -------------
def _textured_emu(port_oracle, sc):
    asm = port_oracle.assemble(sc)
    return asm, E.EmuScene(asm, sc)


def test_textured_path_with_constant_maps_equals_the_live_shader(port_oracle, scene_cache):
    """A base-colour map holding the material's own colour, a metal-rough map holding its own (roughness, metallic) and an
    environment map holding the constant sky 0.3 change nothing: bit-identical accumulator."""
    sc = scene_cache("tiny")
    asm, emu = _textured_emu(port_oracle, sc)
    want, sw = emu.render(sc.camera, 64, 36, 0, 2, 5, fast=1)
    case = H.texture_case(sc)
    slots = case["slots"]
    textures = []
    for g, m in enumerate(slots):
        mat = sc.materials[m]
        textures.append(np.tile(np.array([*mat.base_color, 1.0], dtype=np.float32), (4, 4, 1)))
        textures.append(np.tile(np.array([0.0, mat.roughness, mat.metallic, 1.0], dtype=np.float32), (2, 2, 1)))
    n = len(slots)
    emu.bind_textures(case["uvs_global"], textures, np.arange(n) * 2, np.arange(n) * 2 + 1, np.full((4, 8, 3), 0.3, dtype=np.float32))
    got, sg = emu.render(sc.camera, 64, 36, 0, 2, 5, fast=6)
    assert sg["segments"] == sw["segments"]
    assert got.tobytes() == want.tobytes()


def test_fully_transparent_base_maps_let_every_ray_through(port_oracle, scene_cache):
    """alpha <= 0.01 everywhere: every hit is a pass-through that spends a bounce and draws no random number, so a pixel is
    exactly the sky (0.3) when its ray leaves the scene within max_bounce surfaces and exactly 0 otherwise."""
    sc = scene_cache("tiny")
    asm, emu = _textured_emu(port_oracle, sc)
    case = H.texture_case(sc)
    n = len(case["slots"])
    clear = np.zeros((2, 2, 4), dtype=np.float32)
    emu.bind_textures(case["uvs_global"], [clear], np.zeros(n, dtype=np.int32), np.full(n, -1, dtype=np.int32), None)
    got, _ = emu.render(sc.camera, 64, 36, 0, 1, 5, fast=6)
    rgb = got[..., :3]
    assert np.all((rgb == np.float32(0.3)) | (rgb == 0.0))
    assert (rgb == np.float32(0.3)).any() and (rgb == 0.0).any()


def test_textured_path_changes_the_image_and_is_deterministic(port_oracle, scene_cache):
    sc = scene_cache("tiny")
    asm, emu = _textured_emu(port_oracle, sc)
    plain, _ = emu.render(sc.camera, 64, 36, 0, 2, 5, fast=1)
    case = H.texture_case(sc)
    emu.bind_textures(case["uvs_global"], case["textures"], *case["maps_gpu"], case["env"])
    a, sa = emu.render(sc.camera, 64, 36, 0, 2, 5, fast=6)
    b, sb = emu.render(sc.camera, 64, 36, 0, 2, 5, fast=6)
    assert a.tobytes() == b.tobytes() and sa == sb
    assert H.image_mismatch(a, plain, rel=1e-3) > 0.5  # textures, cut-outs and the environment are all in the picture
    assert np.isfinite(a).all()


def test_texture_sampler_is_bilinear_with_repeat_wrap():
    """tex_sample against a float64 numpy statement of GL_LINEAR / GL_REPEAT (texel centres at (i + 0.5) / size)."""
    import ctypes as C

    lib = E.load()
    rng = np.random.default_rng(11)
    w, h = 7, 5
    tex = np.ascontiguousarray(rng.uniform(0, 1, (h, w, 4)), dtype=np.float32)
    uv = np.ascontiguousarray(rng.uniform(-2.5, 3.5, (2000, 2)), dtype=np.float32)
    out = np.zeros((2000, 4), dtype=np.float32)
    lib.emu_tex_sample(tex.ctypes.data_as(C.c_void_p), C.c_int(w), C.c_int(h), uv.ctypes.data_as(C.c_void_p), C.c_int(2000), out.ctypes.data_as(C.c_void_p))
    x = uv[:, 0].astype(np.float32) * np.float32(w) - np.float32(0.5)
    y = uv[:, 1].astype(np.float32) * np.float32(h) - np.float32(0.5)
    x0, y0 = np.floor(x), np.floor(y)
    fx, fy = (x - x0).astype(np.float64)[:, None], (y - y0).astype(np.float64)[:, None]
    x0, y0 = x0.astype(np.int64), y0.astype(np.int64)
    t = tex.astype(np.float64)
    a, b = t[y0 % h, x0 % w], t[y0 % h, (x0 + 1) % w]
    c, d = t[(y0 + 1) % h, x0 % w], t[(y0 + 1) % h, (x0 + 1) % w]
    want = (a + fx * (b - a)) * (1 - fy) + (c + fx * (d - c)) * fy
    assert np.abs(out - want).max() < 2e-6
