"""GPU suite: the two-level structure.  Radeon Rays switches to a top-level tree over
shapes + one bottom-level tree per mesh whenever the world holds an instance, or when "bvh.force2level" is set
(src/device/calc_intersection_device.cpp:68-106); its kernel takes the ray into the shape's object space and
intersects object-space triangles there (intersect_bvh2level_skiplinks.cl:123-131,205-222).  The oracle here is the
brute-force restatement of exactly that (oracle/port: oracle_intersect_2level / oracle_occluded_2level).

What the oracle pins: ids, t and barycentrics as the reference's TWO-LEVEL kernel computes them (ray transformed,
object-space triangles) -- bit for bit.  Against the FLAT build (vertices transformed, world-space triangles) the
same scene agrees bit for bit only where every transform is the identity; elsewhere the two formulations round
differently (that is true inside Radeon Rays itself) and t agrees to ~1e-6 relative.
"""
import time

import numpy as np
import pytest

from mcrt_b200 import rr, pt, camera, scene_library, layouts as L
from mcrt_b200.scenes import Mesh, Scene, CameraSpec, lambert, tri_mesh_light, mesh_area
from oracle import port, ref
from raygen import random_rays

pytestmark = pytest.mark.gpu


def rot_y(deg):
    c, s = np.cos(np.deg2rad(deg)), np.sin(np.deg2rad(deg))
    m = np.eye(4, dtype=np.float32)
    m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
    return m


def placement(k):
    m = rot_y(37.0 * k)
    m[:3, :3] *= np.float32(0.6 + 0.15 * k)
    m[:3, 3] = (3.0 * (k % 3) - 3.0, 0.4 * k, 3.0 * (k // 3) - 1.5)
    return m


def build_world(api, n_inst=5):
    """A sphere mesh attached once as itself and n_inst times as an instance, plus a floor mesh."""
    sphere = scene_library.uv_sphere("ball", (0.0, 0.0, 0.0), 1.0, 24, 12, 0)
    floor = scene_library.floor_mesh("floor", 8.0, -1.5, 0)
    shapes, table_src = [], []
    s0 = api.CreateMesh(sphere.positions, sphere.positions.shape[0], 12, sphere.indices, 0, None, sphere.indices.shape[0])
    s0.SetId(10)
    api.AttachShape(s0)
    shapes.append(s0)
    table_src.append(("ball", sphere.positions, sphere.indices, 10, None))
    for k in range(n_inst):
        inst = api.CreateInstance(s0)
        inst.SetId(20 + k)
        m = placement(k + 1)
        minv = np.linalg.inv(m.astype(np.float64)).astype(np.float32)
        inst.SetTransform(m, minv)
        api.AttachShape(inst)
        shapes.append(inst)
        table_src.append(("ball", sphere.positions, sphere.indices, 20 + k, minv))
    f = api.CreateMesh(floor.positions, floor.positions.shape[0], 12, floor.indices, 0, None, floor.indices.shape[0])
    f.SetId(3)
    api.AttachShape(f)
    shapes.append(f)
    table_src.append(("floor", floor.positions, floor.indices, 3, None))
    return shapes, table_src


def rays_for_world(n=20000, seed=5):
    rays = random_rays(n, np.array([-6.0, -2.0, -5.0]), np.array([6.0, 4.0, 6.0]), seed)
    rays["o"][:, 3] = 50.0
    return rays


def assert_hits_equal(got, want):
    assert np.array_equal(got["shapeid"], want["shapeid"])
    assert np.array_equal(got["primid"], want["primid"])
    hit = want["shapeid"] >= 0
    assert hit.mean() > 0.3
    assert np.array_equal(got["uvwt"][hit].view(np.uint32), want["uvwt"][hit].view(np.uint32))


def test_instances_commit_two_level_and_match_the_bruteforce_two_level_oracle():
    api = rr.IntersectionApi.Create(0)
    try:
        shapes, src = build_world(api)
        api.Commit()
        st = api.GetStats()
        assert st["two_level"] == 1 and st["num_instances"] == len(src)
        # bottom-level trees exist once per MESH: 2 meshes, not 7 shapes
        sphere_tris = src[0][2].shape[0]
        assert st["num_triangles"] == sphere_tris + src[-1][2].shape[0]
        tris, table = port.two_level_world(src)
        rays = rays_for_world()
        assert_hits_equal(api.QueryIntersectionHost(rays), port.intersect_2level(tris, table, rays))
        assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded_2level(tris, table, rays))
        # ray mask: a ray whose mask equals a shape's id does not see that shape (intersect_bvh2level_skiplinks.cl:236)
        masked = rays.copy()
        masked["extra"][:, 0] = 21
        got = api.QueryIntersectionHost(masked)
        assert not (got["shapeid"] == 21).any()
        assert_hits_equal(got, port.intersect_2level(tris, table, masked))
    finally:
        rr.IntersectionApi.Delete(api)


def test_transform_change_rebuilds_only_the_top_level():
    """SetTransform + Commit on a two-level world: the bottom-level trees stay where they are on the device, the top
    level and the shape table are rebuilt -- and the hits follow (RTScene::updateDynamicEntities, RTScene.cpp:317-368;
    the reference rebuilds everything on the host, intersector_2level.cpp:257-520)."""
    api = rr.IntersectionApi.Create(0)
    try:
        shapes, src = build_world(api)
        api.Commit()
        st0 = api.GetStats()
        rays = rays_for_world(8000, 9)
        times = []
        for step in range(3):
            for k in range(1, 6):
                m = placement(k)
                m[:3, 3] += np.float32(0.25 * (step + 1))
                minv = np.linalg.inv(m.astype(np.float64)).astype(np.float32)
                shapes[k].SetTransform(m, minv)
                src[k] = (src[k][0], src[k][1], src[k][2], src[k][3], minv)
            t0 = time.perf_counter()
            api.Commit()
            times.append(time.perf_counter() - t0)
            tris, table = port.two_level_world(src)
            assert_hits_equal(api.QueryIntersectionHost(rays), port.intersect_2level(tris, table, rays))
        st = api.GetStats()
        assert st["two_level_rebuilds"] == st0["two_level_rebuilds"] and st["top_level_updates"] == st0["top_level_updates"] + 3
        assert min(times) < 0.05, times
    finally:
        rr.IntersectionApi.Delete(api)


def test_forced_two_level_equals_the_flat_tree_bit_for_bit_under_identity_transforms(cornell):
    """"bvh.force2level" on a world without instances and with identity transforms: transforming the ray by the
    identity is exact, so ids, t and barycentrics equal the flat tree's -- and "bvh.forceflat" brings the flat tree back."""
    rays = random_rays(30000, *[np.asarray(b, np.float64) for b in cornell.world_bounds()], 3)
    rays["o"][:, 3] = 100.0
    res = {}
    for mode in ("flat", "two"):
        api = rr.IntersectionApi.Create(0)
        try:
            rr.attach_scene(api, cornell)
            if mode == "two":
                api.SetOption("bvh.force2level", 1.0)
            api.Commit()
            assert api.GetStats()["two_level"] == (1 if mode == "two" else 0)
            res[mode] = (api.QueryIntersectionHost(rays), api.QueryOcclusionHost(rays))
            if mode == "two":
                api.SetOption("bvh.force2level", 0.0)
                api.SetOption("bvh.forceflat", 1.0)
                api.Commit()
                assert api.GetStats()["two_level"] == 0
                res["back"] = (api.QueryIntersectionHost(rays), api.QueryOcclusionHost(rays))
        finally:
            rr.IntersectionApi.Delete(api)
    for other in ("two", "back"):
        assert_hits_equal(res[other][0], res["flat"][0])
        assert np.array_equal(res[other][1], res["flat"][1])


def instanced_scene():
    """Three uses of one sphere mesh (the second and third are instances) over a floor, one emissive panel."""
    ball = scene_library.uv_sphere("ball", (0.0, 0.0, 0.0), 1.0, 32, 16, 0)
    floor = scene_library.floor_mesh("floor", 8.0, -1.0, 1)
    panel = scene_library.floor_mesh("panel", 1.5, 6.0, 2)
    panel.indices = panel.indices[:, [0, 2, 1]].copy()      # faces down
    meshes = [ball, floor, panel]
    for k in (1, 2):
        m = placement(k)
        meshes.append(Mesh(ball.name, ball.positions, ball.normals, ball.uvs, ball.indices, 0, m, instance_of=0))
    mats = np.concatenate([lambert((0.7, 0.3, 0.3)), lambert((0.6, 0.6, 0.6)), lambert((0.0, 0.0, 0.0))])
    sc = Scene("instanced", meshes, mats, np.zeros(0, dtype=L.LIGHT),
               CameraSpec(position=(0.0, 2.5, -9.0), euler_deg=(10.0, 0.0, 0.0)))
    sc.lights = np.concatenate([tri_mesh_light(2, (25.0, 25.0, 25.0), mesh_area(panel))])
    return sc.finalize()


def test_path_tracer_over_an_instanced_scene_vs_reference():
    """The importer's instancing end to end: RTShape entries that share index / vertex ranges, CreateInstance on the
    Radeon Rays side, the two-level trace kernels inside the fused wavefront.  The oracle renders the same RTShape
    table over a FLAT BVH2 (every use of the mesh flattened with its transform), i.e. world-space triangles: the hit
    distances differ in the last bits from the object-space intersection, so this is a tolerance test."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = instanced_scene()
    assert scene.shapes[3]["startIdx"] == scene.shapes[0]["startIdx"] and scene.shapes[3]["numTriangles"] == scene.shapes[0]["numTriangles"]
    w, h, depth = 192, 128, 4
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_pt_pipeline(api, scene, w, h, depth)
        assert api.GetStats()["two_level"] == 1
        got = []
        for _ in range(2):
            pl.update()
            got.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    o = ref.RefOracle()
    rp = ref.RefPathTracer(o, scene, o.build_scene_bvh(scene), w, h)
    for f in range(2):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        want = rp.render_sample(cam, f, depth)
        e = (np.abs(got[f][:, :3] - want[:, :3]) / np.maximum(np.abs(want[:, :3]), 1e-3)).max(axis=1)
        assert (e <= 1e-4).mean() >= 0.97, float((e <= 1e-4).mean())
        assert np.allclose(got[f][:, :3].mean(0), want[:, :3].mean(0), rtol=2e-2)
        assert got[f][:, :3].max() > 0


def test_bdpt_refuses_a_two_level_world_and_runs_on_the_flattened_one():
    scene = instanced_scene()
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_bdpt_pipeline(api, scene, 64, 48, 3)
        with pytest.raises(pt.PathTracerError, match="forceflat"):
            pl.update()
        api.SetOption("bvh.forceflat", 1.0)
        api.Commit()
        pl.update()
        assert pl.read_buffer("RadianceBufferCL").reshape(-1, 4)[:, :3].max() > 0
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
