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
    assert H.image_mismatch(img, want, rel=1e-5) <= 0.002


def test_fast_traversal_equals_reference_walk_bitwise(port_oracle, scene_cache):
    sc = scene_cache("tiny")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    a, sa = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=False)
    b, sb = emu.render(sc.camera, 64, 36, 0, 3, 5, fast=True)
    assert sa["segments"] == sb["segments"]
    assert H.image_mismatch(b, a, rel=0.0, floor=0.0) <= 0.001  # exact except exact-tie pixels
    oracle_img, so = port_oracle.render(asm, sc.camera, 64, 36, 0, 3, 5)
    assert so["segments"] == sa["segments"] and so["node_visits"] == sa["node_visits"] and so["tri_tests"] == sa["tri_tests"]
    assert a.tobytes() == oracle_img.tobytes()


def test_compact_layout_invariants(port_oracle, scene_cache):
    sc = scene_cache("heightfield")
    asm = port_oracle.assemble(sc)
    emu = E.EmuScene(asm, sc)
    blas = emu.fnodes[: emu.tlas_base]
    refs = blas["a"]
    leaf = refs < 0
    counts = ((refs[leaf] >> 29) & 3) + 1
    assert int(counts.sum()) == sc.n_tris  # every triangle in exactly one leaf
    assert np.array_equal(counts, blas["b"][leaf])
    ids = emu.ftris["v0"][:, 3].view(np.int32)
    assert sorted(ids.tolist()) == list(range(sc.n_tris))
    interior = refs[~leaf]
    assert np.all(interior % 2 == 0) and np.all(interior + 1 < emu.tlas_base)
    assert sorted(interior.tolist() + [int(emu.mesh_root_ref[0])]) == list(range(0, emu.tlas_base, 2))  # every pair referenced once
    # pre-gathered normals follow the triangles
    tri = asm.indices["tri"][ids]
    assert np.array_equal(emu.fnormals["v1"][:, :3], asm.vertices["normal"][tri[:, 1]])


def test_inverse_matches_glm_golden():
    import ctypes as C

    lib = E.load()
    g = H.load_golden("inverse_golden.npz")
    for m, want in zip(g["mats"], g["inverses"]):
        out = np.zeros(16, dtype=np.float32)
        lib.emu_mat4_inverse(np.ascontiguousarray(m).ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert out.tobytes() == want.tobytes()
