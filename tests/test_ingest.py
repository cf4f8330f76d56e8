"""Asset ingest: the engine's `.mesh` / `.scene` files through the
library's parser.  Host-side only, so these run without a GPU (the render of an ingested scene
is in test_gpu_parity.py).  Everything is bit-exact: file bytes -> arrays, and TRS chains ->
world matrices against the reference's own TransformComponent (golden from oracle/_ref)."""
import hashlib
import os

import numpy as np
import pytest

import helpers as H

REF_MONKEY = "/root/reference/projects/sample3d/resources/models/monkey"


@pytest.fixture(scope="module")
def cb():
    import caveengine_b200 as cb

    cb.load_library()
    return cb


def _f(x):
    return "%.9g" % float(x)  # what the engine's YAML emitter writes for a float (e.g. 0.600000024)


def write_scene(path, entities):
    """entities: dicts with id, and optionally parent, trs (10 floats), mesh_guid, materials, material (6 floats)."""
    out = ["version: 19", "seed: %d" % (max(e["id"] for e in entities) + 1), "root: 1", "physics_mode: 0", "entities:"]
    for e in entities:
        out.append("  - id: %d" % e["id"])
        out += ["    NameComponent:", "      name: entity-%d" % e["id"]]
        if "parent" in e:
            out += ["    HierarchyComponent:", "      parent_id: %d" % e["parent"]]
        if "trs" in e:
            t = e["trs"]
            out += ["    TransformComponent:", "      translation: [%s]" % ", ".join(_f(x) for x in t[0:3]), "      rotation: [%s]" % ", ".join(_f(x) for x in t[3:7]),
                    "      scale: [%s]" % ", ".join(_f(x) for x in t[7:10])]
        if "mesh_guid" in e:
            out += ["    MeshRendererComponent:", "      is_visible: true", "      cast_shadow: true", "      transparency: false", "      mesh_id: %s" % e["mesh_guid"]]
            mats = e.get("materials", [])
            if e.get("block_materials"):
                out += ["      materials:"] + ["        - %d" % m for m in mats]
            else:
                out.append("      materials: [%s]" % ", ".join(str(m) for m in mats))
        if "material" in e:
            m = e["material"]
            out += ["    MaterialComponent:", "      base_color: [%s, 1]" % ", ".join(_f(x) for x in m[0:3]), "      metallic: %s" % _f(m[4]), "      roughness: %s" % _f(m[3]),
                    "      emissive: %s" % _f(m[5]), "      material_id: 00000000-0000-0000-0000000000000000"]
    with open(path, "w") as f:
        f.write("\n".join(out))


def write_meta(mesh_path, guid):
    with open(mesh_path + ".meta", "w") as f:
        f.write('type: mesh\nguid: %s\nname: m\nimport_path: "@res://%s"\ndependencies: []\nsource_created_time: ""\nsource_last_modified: ""\nimport_settings:\n  {}'
                % (guid, os.path.basename(mesh_path)))


GUID_A = "935F026E-C054-6046-913DC55FA22DE59C"
GUID_B = "1426528D-D731-A34A-B688E84232179ACF"


def test_mesh_file_round_trip(cb, tmp_path):
    from caveengine_b200 import scenes

    m = scenes.heightfield_mesh(13, 7, 99)
    p = str(tmp_path / "h.mesh")
    cb.write_mesh_file(p, m)
    raw = open(p, "rb").read()
    # the layout MeshAsset::SerializeBinary writes: subset_count, vector<MeshSubset>, flags, then the vectors
    assert np.frombuffer(raw, "<u8", 2).tolist() == [1, 1]
    assert np.frombuffer(raw, "<u4", 2, 16).tolist() == [0, m.indices.size]
    assert np.array_equal(np.frombuffer(raw, "<f4", 3, 24), m.positions.min(axis=0)) and np.array_equal(np.frombuffer(raw, "<f4", 3, 36), m.positions.max(axis=0))
    assert len(raw) == 16 + 32 + 4 + 8 + m.indices.size * 4 + 2 * (8 + m.positions.size * 4) + 6 * 8
    with cb.SceneFile(p) as sf:
        assert (sf.mesh_count, sf.instance_count, sf.material_count, sf.skipped_renderers) == (1, 1, 1, 0)
        sc = sf.to_scene()
    got = sc.meshes[0]
    assert got.positions.tobytes() == m.positions.tobytes() and got.normals.tobytes() == m.normals.tobytes() and got.indices.tobytes() == m.indices.tobytes()
    assert np.array_equal(sc.instances[0].transform, np.eye(4, dtype=np.float32).ravel())


def test_world_matrices_match_reference_transform_component(cb, tmp_path):
    """TRS chains -> world matrices, bit for bit against the reference's TransformComponent::GetLocalMatrix
    and the UpdateHierarchy product (golden generated from oracle/_ref by make_ingest_golden.py)."""
    from caveengine_b200 import scenes

    g = H.load_golden("ingest_golden.npz")
    trs, lens, want = g["chains/trs"], g["chains/len"], g["chains/world"]
    mesh_path = str(tmp_path / "q.mesh")
    cb.write_mesh_file(mesh_path, scenes._quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1)))
    write_meta(mesh_path, GUID_A)
    entities, next_id, leaf_ids = [], 1, []
    mat_id = 1000
    for c in range(trs.shape[0]):
        ids = list(range(next_id, next_id + int(lens[c])))
        next_id += int(lens[c])
        for k, eid in enumerate(ids):  # k = 0 is the renderable entity, k+1 its parent ...
            e = {"id": eid, "trs": trs[c, k]}
            if k + 1 < len(ids):
                e["parent"] = ids[k + 1]
            if k == 0:
                e.update(mesh_guid=GUID_A, materials=[mat_id], block_materials=bool(c % 2))
                leaf_ids.append(eid)
            entities.append(e)
    entities.append({"id": mat_id, "material": [0.25, 0.5, 0.75, 0.3, 0.9, 2.0]})
    scene_path = str(tmp_path / "chains.scene")
    write_scene(scene_path, entities)
    with cb.SceneFile(scene_path) as sf:
        assert (sf.mesh_count, sf.instance_count, sf.material_count, sf.skipped_renderers) == (1, trs.shape[0], 1, 0)
        sc = sf.to_scene()
    got = np.stack([i.transform for i in sc.instances])
    assert got.tobytes() == want.tobytes(), f"{(got.view(np.uint32) != want.view(np.uint32)).any(axis=1).sum()} of {len(want)} world matrices differ from glm's"
    m = sc.materials[0]
    assert np.array_equal(np.float32([*m.base_color, m.roughness, m.metallic, m.emissive]), np.float32([0.25, 0.5, 0.75, 0.3, 0.9, 2.0]))


def test_scene_semantics(cb, tmp_path):
    """Unique meshes in first-appearance order, material = materials[0], unresolved or transform-less
    renderers skipped (DEV_VERIFY, path_tracer.cpp:216), entity ids need not be dense."""
    from caveengine_b200 import scenes

    a, b = str(tmp_path / "sub" / "a.mesh"), str(tmp_path / "b.mesh")
    os.makedirs(tmp_path / "sub")
    cb.write_mesh_file(a, scenes._open_box(0.8, 0.8, 0.8))
    cb.write_mesh_file(b, scenes._sphere(12, 8, 0.5, 8))
    write_meta(a, GUID_A)
    write_meta(b, GUID_B.lower())  # GUID text compares case-insensitively
    ident = [0, 0, 0, 0, 0, 0, 1, 1, 1, 1]
    ents = [
        {"id": 1, "trs": ident},
        {"id": 7, "parent": 1, "trs": ident, "mesh_guid": GUID_B, "materials": [9, 8]},
        {"id": 8, "material": [1, 0, 0, 1, 0, 0]},
        {"id": 9, "material": [0, 1, 0, 0.5, 0.5, 3]},
        {"id": 12, "parent": 1, "trs": [1, 2, 3, 0, 0, 0, 1, 2, 2, 2], "mesh_guid": GUID_A, "materials": [8]},
        {"id": 13, "parent": 1, "trs": ident, "mesh_guid": "00000000-0000-0000-0000000000000004", "materials": [8]},  # engine built-in: no file
        {"id": 14, "parent": 1, "mesh_guid": GUID_A, "materials": [8]},  # no TransformComponent
        {"id": 15, "parent": 1, "trs": ident, "mesh_guid": GUID_B, "materials": [8]},
    ]
    p = str(tmp_path / "s.scene")
    write_scene(p, ents)
    with cb.SceneFile(p) as sf:
        assert (sf.mesh_count, sf.instance_count, sf.material_count, sf.skipped_renderers) == (2, 3, 2, 2)
        sc = sf.to_scene()
    assert [i.mesh for i in sc.instances] == [0, 1, 0] and [i.material for i in sc.instances] == [0, 1, 1]
    assert sc.meshes[0].n_tris == scenes._sphere(12, 8, 0.5, 8).n_tris and sc.meshes[1].n_tris == scenes._open_box(0.8, 0.8, 0.8).n_tris
    assert sc.materials[0].base_color == (0.0, 1.0, 0.0) and float(sc.materials[0].emissive) == 3.0
    t = sc.instances[1].transform.reshape(4, 4).T
    assert np.array_equal(t, np.float32([[2, 0, 0, 1], [0, 2, 0, 2], [0, 0, 2, 3], [0, 0, 0, 1]]))


@pytest.mark.parametrize("case", ["truncated_mesh", "no_material_component", "no_materials", "cycle", "bad_yaml", "missing_file", "nothing_resolves"])
def test_errors_are_reported_not_trapped(cb, tmp_path, case):
    from caveengine_b200 import scenes

    mesh = str(tmp_path / "a.mesh")
    cb.write_mesh_file(mesh, scenes._quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1)))
    write_meta(mesh, GUID_A)
    ident = [0, 0, 0, 0, 0, 0, 1, 1, 1, 1]
    ents = [{"id": 1, "trs": ident}, {"id": 2, "parent": 1, "trs": ident, "mesh_guid": GUID_A, "materials": [3]}, {"id": 3, "material": [1, 1, 1, 1, 0, 0]}]
    p = str(tmp_path / "s.scene")
    if case == "truncated_mesh":
        raw = open(mesh, "rb").read()
        open(mesh, "wb").write(raw[: len(raw) // 2])
    elif case == "no_material_component":
        ents[2] = {"id": 3, "trs": ident}
    elif case == "no_materials":
        ents[1]["materials"] = []
    elif case == "cycle":
        ents[0]["parent"] = 2
    elif case == "nothing_resolves":
        ents[1]["mesh_guid"] = GUID_B
    write_scene(p, ents)
    if case == "bad_yaml":
        open(p, "a").write("\n   stray: [1, 2\n")
    if case == "missing_file":
        p = str(tmp_path / "absent.scene")
    with pytest.raises(cb.CaveptError) as ei:
        cb.SceneFile(p)
    assert ei.value.code == 1 and len(str(ei.value)) > 20


@pytest.mark.skipif(not os.path.isdir(REF_MONKEY), reason="needs /root/reference (build container only)")
def test_reference_assets_live(cb):
    """The reference's own monkey.scene + Suzanne.mesh + .meta side-cars, straight from /root/reference."""
    g = H.load_golden("ingest_golden.npz")
    with cb.SceneFile(os.path.join(REF_MONKEY, "monkey.scene")) as sf:
        assert (sf.mesh_count, sf.instance_count, sf.material_count, sf.skipped_renderers) == (1, 1, 1, 0)
        sc = sf.to_scene()
    m = sc.meshes[0]
    assert [m.positions.shape[0], m.indices.size] == g["suzanne/counts"].tolist()
    for k, a in (("indices", m.indices), ("positions", m.positions), ("normals", m.normals)):
        assert hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest() == g[f"suzanne/sha_{k}"].tobytes(), k
    mat = sc.materials[0]
    assert np.array_equal(np.float32([*mat.base_color, mat.roughness, mat.metallic, mat.emissive]), g["suzanne/material"])
    assert np.array_equal(sc.instances[0].transform, np.eye(4, dtype=np.float32).ravel())
    with cb.SceneFile(os.path.join(REF_MONKEY, "Suzanne.mesh")) as sf2:
        assert sf2.to_scene().meshes[0].positions.tobytes() == m.positions.tobytes()
