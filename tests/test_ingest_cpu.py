"""CPU suite: OBJ / MTL / texture ingestion (monte-carlo-raytracer_b200/ingest.py)."""
import os

import numpy as np
import pytest

from mcrt_b200 import ingest, scene_library, layouts as L

REF_CORNELL = "/root/reference/assets/meshes/cornell-box/CornellBox-Original.obj"

TINY_OBJ = """mtllib tiny.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0 0 1
vt 0 0
vt 1 0
vt 1 1
vt 0 1
g quad
usemtl textured
f 1/1 2/2 3/3 4/4
g tri
usemtl lamp
f 1 2 5
"""
TINY_MTL = """newmtl textured
Kd 0.5 0.25 1.0
Ks 0.1 0.1 0.1
Ns 18
map_Kd checker.png
map_d checker.png
newmtl lamp
Kd 0 0 0
Ke 3 2 1
"""


@pytest.fixture()
def tiny(tmp_path):
    from PIL import Image
    (tmp_path / "tiny.obj").write_text(TINY_OBJ)
    (tmp_path / "tiny.mtl").write_text(TINY_MTL)
    img = np.zeros((4, 8, 4), np.uint8)
    img[..., 3] = 255
    img[0, 0] = (255, 0, 0, 255)          # first texel of the file = first texel of the buffer
    Image.fromarray(img, "RGBA").save(tmp_path / "checker.png")
    return ingest.load_obj(str(tmp_path / "tiny.obj"), scale=2.0)


def test_tiny_scene_layout(tiny):
    s = tiny
    assert [m.name for m in s.meshes] == ["quad", "tri"]
    assert s.num_triangles == 3 and len(s.shapes) == 2 and len(s.materials) == 2
    q = s.meshes[0]
    # quad: fan (0,1,2),(0,2,3), winding reversed, z negated, scaled
    assert q.indices.tolist() == [[0, 2, 1], [0, 3, 2]]
    assert np.allclose(q.positions, [[0, 0, 0], [2, 0, 0], [2, 2, 0], [0, 2, 0]])
    assert np.allclose(q.uvs, [[0, 0], [1, 0], [1, 1], [0, 1]])
    # flat normal of the OBJ quad is +z; left-handed flip makes it -z, and the reversed winding agrees
    assert np.allclose(q.normals, [[0, 0, -1]] * 4)
    p = q.positions
    gn = np.cross(p[q.indices[0, 0]] - p[q.indices[0, 2]], p[q.indices[0, 1]] - p[q.indices[0, 2]])
    assert gn[2] < 0
    t = s.meshes[1]
    assert np.allclose(t.positions[2], [0, 0, -2])
    # materials and textures
    m = s.materials[0]
    assert np.allclose(m["uber_kd"][:3], [0.5, 0.25, 1.0]) and np.allclose(m["uber_ks"][:3], 0.1)
    assert np.isclose(m["uber_roughness"][0], np.sqrt(2.0 / 20.0))
    assert m["uber_diffuseTexId"] == 0 and m["uber_opacityTexId"] == 0 and m["uber_normalMapId"] == L.RT_INVALID_ID
    assert len(s.textures) == 1 and s.textures[0]["width"] == 8 and s.textures[0]["height"] == 4
    assert s.texdata.size == 8 * 4 * 4 and s.texdata[:4].tolist() == [255, 0, 0, 255]
    # the emissive group is a triangle-mesh area light bound to its shape
    assert len(s.lights) == 1 and s.lights[0]["type"] == L.RT_TRIANGLE_MESH_AREA_LIGHT
    assert s.lights[0]["shapeId"] == 1 and s.shapes[1]["lightID"] == 0 and s.shapes[0]["lightID"] == L.RT_INVALID_ID
    assert np.allclose(s.lights[0]["intensity"][:3], [3, 2, 1]) and np.isclose(s.lights[0]["area"], 2.0)
    assert s.shapes[1]["startIdx"] == 6 and s.shapes[1]["startVertex"] == 4


@pytest.mark.skipif(not os.path.exists(REF_CORNELL), reason="reference assets not present")
def test_reference_cornell_obj_reproduces_the_builtin_scene():
    """Importing the reference's own CornellBox-Original.obj (scale 5, PathTracingApp.cpp:371) gives the
    scene the C1 / golden tests use: same shapes in the same order, same world-space triangles, same
    materials and light."""
    a = ingest.load_obj(REF_CORNELL, scale=5.0, camera=scene_library.CORNELL_CAMERA)
    b = scene_library.cornell_box()
    # the file declares the boxes' faces before their `g` line (under the previous group, with their own
    # `usemtl`), so sub-meshes are identified by (group, material) in order of appearance: 8 of them
    assert len(a.meshes) == len(b.meshes) == 8
    def world_tris(scene, sid):
        m = scene.meshes[sid]
        return m.positions[m.indices].astype(np.float64)        # identity transforms

    assert a.num_triangles == b.num_triangles == 36
    for sid in range(len(b.meshes)):
        sa, sb = world_tris(a, sid), world_tris(b, sid)
        # same triangles with the same orientation (vertex order up to rotation)
        key = lambda t: sorted(map(tuple, np.round(t, 5).tolist()))
        assert sorted(key(t) for t in sa) == sorted(key(t) for t in sb)
        na = np.cross(sa[:, 1] - sa[:, 0], sa[:, 2] - sa[:, 0]).sum(0)
        nb = np.cross(sb[:, 1] - sb[:, 0], sb[:, 2] - sb[:, 0]).sum(0)
        assert np.dot(na, nb) > 0
    assert np.allclose(a.materials["uber_kd"], b.materials["uber_kd"])
    assert np.allclose(a.materials["uber_roughness"], b.materials["uber_roughness"])
    assert len(a.lights) == 1 and a.lights[0]["shapeId"] == b.lights[0]["shapeId"]
    assert np.allclose(a.lights[0]["intensity"], b.lights[0]["intensity"]) and np.isclose(a.lights[0]["area"], b.lights[0]["area"])


def test_second_use_of_an_asset_becomes_an_instance(tiny, tmp_path):
    """RTScene::attachMesh (RTScene.cpp:570-598): the second use of a mesh adds RTShapes that SHARE the template's
    index / vertex ranges and material and carry their own transform; the Radeon Rays side gets CreateInstance.
    Geometry, materials and textures are stored once."""
    T1 = np.eye(4, dtype=np.float32)
    T1[:3, 3] = (5.0, 0.0, 0.0)
    T2 = np.eye(4, dtype=np.float32)
    T2[:3, :3] *= 2.0
    obj = str(tmp_path / "tiny.obj")
    s = ingest.load_entities([(obj, None, 2.0), (obj, T1, 2.0), (obj, T2, 2.0)])
    assert len(s.meshes) == 6 and len(s.shapes) == 6
    assert [m.instance_of for m in s.meshes] == [None, None, 0, 1, 0, 1]
    # geometry pooled once
    assert s.indices.size == tiny.indices.size and s.positions.shape == tiny.positions.shape
    assert len(s.materials) == len(tiny.materials) and len(s.textures) == len(tiny.textures)
    for i, t in ((2, 0), (3, 1), (4, 0), (5, 1)):
        for k in ("startIdx", "startVertex", "numTriangles", "materialId"):
            assert s.shapes[i][k] == s.shapes[t][k]
    assert np.array_equal(s.shapes[2]["toWorldTransform"], T1) and np.array_equal(s.shapes[4]["toWorldTransform"], T2)
    assert np.allclose(s.shapes[4]["toWorldInverseTranspose"][:3, :3], np.eye(3) * 0.5)
    # the emissive group is a light once per import, linked to the template's shape
    assert len(s.lights) == 1 and int(s.lights[0]["shapeId"]) == 1 and s.shapes[1]["lightID"] == 0
    # a different scale is a different asset, not an instance
    s2 = ingest.load_entities([(obj, None, 2.0), (obj, None, 1.0)])
    assert [m.instance_of for m in s2.meshes] == [None, None, None, None]
    assert s2.indices.size == 2 * tiny.indices.size
