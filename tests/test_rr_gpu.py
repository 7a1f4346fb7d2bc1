"""GPU suite: the CUDA traversal behind the IntersectionApi mirror (mcrt_b200.rr -> C ABI), checked
against the brute-force oracle and -- where it was built -- the compiled reference kernel.
The cases restate the reference's own API tests (UnitTest/radeon_rays_apitest_cl.h,
radeon_rays_conformance_test_cl.h, radeon_rays_test_cl.h)."""
import os

import numpy as np
import pytest

from mcrt_b200 import rr, layouts as L, scene_library
from oracle import port
from raygen import random_rays, camera_rays

pytestmark = pytest.mark.gpu

TRI_VERTS = np.array([[-1, -1, 0], [0, 1, 0], [1, -1, 0]], np.float32)
TRI_IDX = np.array([[0, 1, 2]], np.int32)


@pytest.fixture()
def api():
    a = rr.IntersectionApi.Create(0)
    yield a
    rr.IntersectionApi.Delete(a)


def one_ray(o, d, maxt=10000.0):
    r = L.make_rays(1)
    r["o"][0, :3] = o
    r["o"][0, 3] = maxt
    r["d"][0, :3] = d
    return r


def single_triangle(api):
    mesh = api.CreateMesh(TRI_VERTS, 3, 12, TRI_IDX, 0, None, 1)
    api.AttachShape(mesh)
    api.Commit()
    return mesh


def test_device_info():
    assert rr.IntersectionApi.GetDeviceCount() >= 1
    assert "NVIDIA" in rr.IntersectionApi.GetDeviceInfo(0)["vendor"]


def test_empty_scene_commit_throws(api):
    """EmptyScene (radeon_rays_apitest_cl.h:140): Commit() on an empty world throws."""
    assert api.IsWorldEmpty()
    with pytest.raises(rr.RadeonRaysError, match="Scene is empty"):
        api.Commit()


def test_intersection_1ray(api):
    """Intersection_1Ray (:203-249) through device buffers, map/unmap and events."""
    mesh = single_triangle(api)
    ray = one_ray((0, 0, -10), (0, 0, 1))
    ray_buf = api.CreateBuffer(48, ray)
    isect_buf = api.CreateBuffer(32, None)
    ev = api.QueryIntersection(ray_buf, 1, isect_buf)
    ev.Wait()
    assert ev.Complete()
    data, mev = api.MapBuffer(isect_buf, rr.MAP_READ, 0, 32, dtype=L.INTERSECTION)
    mev.Wait()
    assert data["shapeid"][0] == mesh.GetId() and data["primid"][0] == 0
    assert data["uvwt"][0, 3] == np.float32(10.0)
    api.UnmapBuffer(isect_buf).Wait()
    api.DeleteEvent(ev)
    api.DeleteBuffer(ray_buf)
    api.DeleteBuffer(isect_buf)


def test_intersection_1ray_inactive_not_written(api):
    """Intersection_1Ray_Active (:413-489): inactive ray leaves the pre-filled kNullId hit untouched."""
    single_triangle(api)
    ray = one_ray((0, 0, -10), (0, 0, 1))
    ray["extra"][0, 1] = 0
    pre = np.zeros(1, dtype=L.INTERSECTION)
    pre["shapeid"], pre["primid"] = 4242, 77
    pre["uvwt"] = 3.5
    got = api.QueryIntersectionHost(ray, pre.copy())
    assert got.tobytes() == pre.tobytes()
    assert api.QueryOcclusionHost(ray, np.array([55], np.int32))[0] == 55


def test_miss_leaves_uvwt_untouched(api):
    single_triangle(api)
    pre = np.zeros(1, dtype=L.INTERSECTION)
    pre["uvwt"] = 9.25
    pre["pad"] = 31
    got = api.QueryIntersectionHost(one_ray((5, 5, -10), (0, 0, 1)), pre.copy())
    assert got["shapeid"][0] == -1 and got["primid"][0] == -1
    assert (got["uvwt"] == 9.25).all() and (got["pad"] == 31).all()


def test_intersection_3rays(api):
    """Intersection_3Rays (:491-553)."""
    mesh = single_triangle(api)
    rays = np.concatenate([one_ray((0, 0, -10), (0, 0, 1)), one_ray((0, 0.5, -10), (0, 0, 1)),
                           one_ray((0.5, -0.5, -10), (0, 0, 1))])
    hits = api.QueryIntersectionHost(rays)
    assert (hits["shapeid"] == mesh.GetId()).all()
    assert (api.QueryOcclusionHost(rays) == 1).all()


def test_masked_ray_skips_shape(api):
    """Intersection_1Ray_Masked_* (:251-300)."""
    mesh = single_triangle(api)
    ray = one_ray((0, 0, -10), (0, 0, 1))
    ray["extra"][0, 0] = mesh.GetId()
    assert api.QueryIntersectionHost(ray)["shapeid"][0] == rr.K_NULL_ID
    assert api.QueryOcclusionHost(ray)[0] == -1


def test_transform_and_dynamic_geometry(api):
    """Intersection_1Ray_Transformed / _DynamicGeo (:556,:646): SetTransform + Commit moves the shape."""
    mesh = single_triangle(api)
    ray0, ray5 = one_ray((0, 0, -10), (0, 0, 1)), one_ray((5, 0, -10), (0, 0, 1))
    assert api.QueryIntersectionHost(ray0)["shapeid"][0] == mesh.GetId()
    m = np.eye(4, dtype=np.float32)
    m[0, 3] = 5.0
    mesh.SetTransform(m)
    api.Commit()
    assert api.QueryIntersectionHost(ray0)["shapeid"][0] == rr.K_NULL_ID
    assert api.QueryIntersectionHost(ray5)["shapeid"][0] == mesh.GetId()
    got_m, _ = mesh.GetTransform()
    assert np.array_equal(got_m, m)


def test_instances(api):
    """Intersection_1Ray_TransformedInstance* / _InstanceNoShape (:844-1082): an instance carries its
    own id and transform; the base mesh need not be attached."""
    mesh = api.CreateMesh(TRI_VERTS, 3, 12, TRI_IDX, 0, None, 1)
    inst = api.CreateInstance(mesh)
    m = np.eye(4, dtype=np.float32)
    m[1, 3] = 4.0
    inst.SetTransform(m)
    inst.SetId(91)
    api.AttachShape(inst)
    api.Commit()
    assert api.QueryIntersectionHost(one_ray((0, 0, -10), (0, 0, 1)))["shapeid"][0] == rr.K_NULL_ID
    assert api.QueryIntersectionHost(one_ray((0, 4, -10), (0, 0, 1)))["shapeid"][0] == 91
    api.AttachShape(mesh)
    api.AttachShape(mesh)       # idempotent (world.cpp:30)
    api.Commit()
    assert api.QueryIntersectionHost(one_ray((0, 0, -10), (0, 0, 1)))["shapeid"][0] == mesh.GetId()
    api.DetachShape(mesh)
    api.Commit()
    assert api.QueryIntersectionHost(one_ray((0, 0, -10), (0, 0, 1)))["shapeid"][0] == rr.K_NULL_ID


def test_mesh_strided(api):
    """MeshStrided (:146): byte strides on vertices and indices."""
    v = np.zeros((3, 5), np.float32)
    v[:, :3] = TRI_VERTS
    ix = np.array([[0, 1, 2, -1]], np.int32)
    mesh = api.CreateMesh(v, 3, 20, ix, 16, None, 1)
    api.AttachShape(mesh)
    api.Commit()
    assert api.QueryIntersectionHost(one_ray((0, 0, -10), (0, 0, 1)))["shapeid"][0] == mesh.GetId()


def test_options_and_id_counter(api):
    for k, v in [("acc.type", "fatbvh"), ("bvh.builder", "sah"), ("bvh.force2level", 0.0), ("bvh.sah.num_bins", 64.0),
                 ("bvh.sah.traversal_cost", 10.0), ("some.unknown.key", "x")]:
        api.SetOption(k, v)
    a = api.CreateMesh(TRI_VERTS, 3, 12, TRI_IDX, 0, None, 1)
    b = api.CreateMesh(TRI_VERTS, 3, 12, TRI_IDX, 0, None, 1)
    assert b.GetId() == a.GetId() + 1
    api.ResetIdCounter()
    assert api.CreateMesh(TRI_VERTS, 3, 12, TRI_IDX, 0, None, 1).GetId() == 1
    with pytest.raises(rr.RadeonRaysError):
        api.CreateMesh(TRI_VERTS, 3, 12, np.array([[0, 1, 2, 2]], np.int32), 0, np.array([4], np.int32), 1)


def test_interop_wrap_and_indirect_count(api):
    """ApiCl interop cases (radeon_rays_test_cl.h:110-540): caller-owned device memory wrapped without
    a copy, ray count read from device memory, dependency events."""
    import torch
    mesh = single_triangle(api)
    rays = np.concatenate([one_ray((0, 0, -10), (0, 0, 1))] * 8)
    t_rays = torch.from_numpy(rays.view(np.uint8).copy()).cuda()
    t_hits = torch.full((8 * 8,), -7, dtype=torch.int32, device="cuda")
    t_cnt = torch.tensor([5], dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rb = api.CreateFromDevicePtr(t_rays.data_ptr(), t_rays.numel())
    hb = api.CreateFromDevicePtr(t_hits.data_ptr(), t_hits.numel() * 4)
    cb = api.CreateFromDevicePtr(t_cnt.data_ptr(), 4)
    ev = api.QueryIntersectionIndirect(rb, cb, 8, hb)
    ev2 = api.QueryOcclusion(rb, 8, api.CreateBuffer(32), waitevent=ev)
    ev2.Wait()
    h = t_hits.cpu().numpy().view(L.INTERSECTION)
    assert (h["shapeid"][:5] == mesh.GetId()).all()
    assert (h["shapeid"][5:] == -7).all()          # indices >= *numrays are not written


@pytest.mark.parametrize("n", [1, 100, 1000, 10000])
def test_conformance_cornell(api, cornell, n):
    """ApiConformanceCL (radeon_rays_conformance_test_cl.h:118-560): random rays against brute force.
    Bar: ids and t bit-exact, ties resolved to the smallest (shape, prim) -- i.e. identical to the
    oracle, which applies the same documented rule."""
    rr.attach_scene(api, cornell)
    api.SetOption("acc.type", "fatbvh")
    api.SetOption("bvh.builder", "sah")
    api.Commit()
    lo, hi = cornell.world_bounds()
    rays = random_rays(n, lo - 1.0, hi + 1.0, seed=0xABCDEF12 % (2 ** 31))
    tris = port.scene_triangles(cornell)
    got = api.QueryIntersectionHost(rays)
    want = port.intersect(tris, rays)
    res = port.compare_hits(got, want, tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))


def test_primary_hits_bit_exact_vs_oracle_and_reference(api, cornell, request):
    """BASELINE config 1 primary hits, 512x512: primitive ids bit-exact against the oracle; against
    the reference's own kernel the only differences allowed are ties."""
    rr.attach_scene(api, cornell)
    api.Commit()
    rays, _ = camera_rays(cornell, 512, 512)
    tris = port.scene_triangles(cornell)
    got = api.QueryIntersectionHost(rays)
    want, ties = port.intersect(tris, rays, want_ties=True)
    res = port.compare_hits(got, want, tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    from oracle import ref
    if ref.available():
        o = ref.RefOracle()
        ref_hits = o.intersect(o.build_scene_bvh(cornell), rays)
        r2 = port.compare_hits(got, ref_hits, tris, rays)
        assert r2["bad"] == 0, r2
        # the reference's slab test may cull a leaf its own triangle test accepts (E4): < 1e-5 of rays
        assert r2["b_farther"] <= max(1, rays.shape[0] // 100000), r2
        # every id difference sits on a pixel where the oracle found more than one triangle at that t
        assert all(ties[i] > 1 for i in r2["tie_idx"])


def test_large_soup_bit_exact(api):
    """200k-triangle random soup, 20k incoherent rays: closest hit and any hit against brute force."""
    rng = np.random.RandomState(7)
    n = 200000
    c = rng.rand(n, 1, 3) * 20.0
    v = (c + (rng.rand(n, 3, 3) - 0.5) * 0.5).reshape(-1, 3).astype(np.float32)
    idx = np.arange(3 * n, dtype=np.int32).reshape(n, 3)
    mesh = api.CreateMesh(v, 3 * n, 12, idx, 12, None, n)
    mesh.SetId(3)
    api.AttachShape(mesh)
    api.Commit()
    rays = random_rays(20000, np.zeros(3), np.full(3, 20.0), 99, maxt=8.0)
    tris = port.mesh_triangles([(v, idx, 3, None)])
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))
    st = api.GetStats()
    assert st["num_triangles"] == n and st["num_nodes"] > 0


def _rot_y(deg, t=(0.0, 0.0, 0.0)):
    a = np.deg2rad(deg)
    m = np.eye(4, dtype=np.float32)
    m[0, 0], m[0, 2], m[2, 0], m[2, 2] = np.cos(a), np.sin(a), -np.sin(a), np.cos(a)
    m[:3, 3] = t
    return m


def test_device_refit_equals_rebuild(api, cornell):
    """Dynamic updates: with option "bvh.refit" a transform-only Commit refits the tree on
    the device.  Hits must equal the brute-force oracle on the moved geometry AND a from-scratch build, the
    device pointers of the tree must not move, and a topology change must fall back to a full build."""
    shapes = rr.attach_scene(api, cornell)
    inst = api.CreateInstance(shapes[len(shapes) - 1])
    inst.SetId(77)
    inst.SetTransform(_rot_y(10.0, (0.2, 0.5, -0.1)))
    api.AttachShape(inst)
    api.SetOption("bvh.refit", 1.0)
    api.Commit()
    st0 = api.GetStats()
    assert st0["refits"] == 0 and st0["last_commit_refit"] == 0
    ptrs0 = api.GetBvh()[:2]
    lo, hi = cornell.world_bounds()
    rays = random_rays(20000, lo - 1.0, hi + 1.0, seed=4711)
    xf = {}
    for step, (deg, t) in enumerate([(25.0, (0.3, 0.0, 0.2)), (-40.0, (-0.4, 0.3, 0.0)), (0.0, (0.0, 0.0, 0.0))]):
        for k in (len(shapes) - 1, len(shapes) - 2):          # the two boxes move, the room stays
            xf[k] = _rot_y(deg * (1 if k % 2 else -1), t)
            shapes[k].SetTransform(xf[k])
        inst_m = _rot_y(3.0 * deg, (0.2 + 0.1 * step, 0.5, -0.1))
        inst.SetTransform(inst_m)
        inst.SetId(77 + step)
        api.Commit()
        st = api.GetStats()
        assert st["refits"] == step + 1 and st["last_commit_refit"] == 1
        assert api.GetBvh()[:2] == ptrs0
        meshes = [(m.positions, m.indices, i, xf.get(i, m.transform)) for i, m in enumerate(cornell.meshes)]
        last = cornell.meshes[len(shapes) - 1]
        meshes.append((last.positions, last.indices, 77 + step, inst_m))
        tris = port.mesh_triangles(meshes)
        got = api.QueryIntersectionHost(rays)
        res = port.compare_hits(got, port.intersect(tris, rays), tris, rays)
        assert res["bad"] == 0 and res["tie_ok"] == 0, res
        assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))
        # a from-scratch build of the same world gives the same answers
        fresh = rr.IntersectionApi.Create(0)
        try:
            for pos, idx, sid, m in meshes:
                sh = fresh.CreateMesh(pos, pos.shape[0], 12, idx, 0, None, idx.shape[0])
                sh.SetId(sid)
                if m is not None:
                    sh.SetTransform(m)
                fresh.AttachShape(sh)
            fresh.Commit()
            ref_hits = fresh.QueryIntersectionHost(rays)
        finally:
            rr.IntersectionApi.Delete(fresh)
        for f in ("shapeid", "primid"):
            assert np.array_equal(got[f], ref_hits[f])
        hit = got["shapeid"] >= 0
        assert np.array_equal(got["uvwt"][hit].view(np.uint32), ref_hits["uvwt"][hit].view(np.uint32))
    # topology change -> full build again, and refits resume afterwards
    api.DetachShape(inst)
    api.Commit()
    assert api.GetStats()["last_commit_refit"] == 0
    shapes[0].SetTransform(_rot_y(0.0, (0.0, 0.01, 0.0)))
    api.Commit()
    assert api.GetStats()["last_commit_refit"] == 1


def test_device_refit_large_soup(api):
    """Refit at a size where every level of the tree has many nodes: 100k-triangle soup in 16 shapes, each shape
    given its own rigid motion."""
    rng = np.random.RandomState(11)
    shapes, geo = [], []
    per = 6250
    for k in range(16):
        c = rng.rand(per, 1, 3) * 10.0 + np.array([[(k % 4) * 10.0, (k // 4) * 10.0, 0.0]])
        v = (c + (rng.rand(per, 3, 3) - 0.5) * 0.4).reshape(-1, 3).astype(np.float32)
        idx = np.arange(3 * per, dtype=np.int32).reshape(per, 3)
        sh = api.CreateMesh(v, 3 * per, 12, idx, 12, None, per)
        sh.SetId(k)
        api.AttachShape(sh)
        shapes.append(sh)
        geo.append((v, idx))
    api.SetOption("bvh.refit", 1.0)
    api.Commit()
    xfs = [_rot_y(rng.uniform(-60, 60), rng.uniform(-3, 3, 3)) for _ in range(16)]
    for sh, m in zip(shapes, xfs):
        sh.SetTransform(m)
    api.Commit()
    assert api.GetStats()["last_commit_refit"] == 1
    rays = random_rays(20000, np.array([-5.0, -5.0, -5.0]), np.array([45.0, 45.0, 15.0]), 5, maxt=30.0)
    tris = port.mesh_triangles([(v, idx, k, xfs[k]) for k, (v, idx) in enumerate(geo)])
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))


def test_golden_intersections(api, cornell):
    """Committed vectors from the reference's own kernel (tests/golden/cornell_isect_4096.npz)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornell_isect_4096.npz"))
    rays = np.ascontiguousarray(g["rays"]).view(L.RAY)
    rr.attach_scene(api, cornell)
    api.Commit()
    ref_hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
    ref_hits["shapeid"], ref_hits["primid"], ref_hits["uvwt"] = g["shapeid"], g["primid"], g["uvwt"]
    tris = port.scene_triangles(cornell)
    res = port.compare_hits(api.QueryIntersectionHost(rays), ref_hits, tris, rays)
    assert res["bad"] == 0 and res["b_farther"] <= 1, res
    assert np.array_equal(api.QueryOcclusionHost(rays), g["occluded"])


@pytest.mark.parametrize("how", ["acc.type=hlbvh", "bvh.builder=lbvh"])
def test_device_built_tree_cornell(api, cornell, how):
    """acc.type = "hlbvh" (Radeon Rays' GPU-built hierarchy, intersector_hlbvh.cpp) selects the device builder
    (lbvh_build.cu).  Hit results do not depend on the tree: ids, t and barycentrics equal the brute-force oracle's
    bit for bit, any-hit booleans are equal, also for directions with zero components."""
    rr.attach_scene(api, cornell)
    k, v = how.split("=")
    api.SetOption(k, v)
    api.Commit()
    st = api.GetStats()
    assert st["last_build_on_device"] == 1 and st["num_triangles"] == 36 and 1 <= st["num_levels"] <= 19
    tris = port.scene_triangles(cornell)
    lo, hi = cornell.world_bounds()
    rays = random_rays(10000, lo - 1.0, hi + 1.0, seed=1234)
    rays["d"][::7, 1] = np.float32(-0.0)
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))
    cam, _ = camera_rays(cornell, 128, 128)
    res = port.compare_hits(api.QueryIntersectionHost(cam), port.intersect(tris, cam), tris, cam)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res


def test_device_built_tree_large_soup_and_duplicates(api):
    """Device build of a 200k-triangle soup in two meshes (one transformed, one instance-free duplicate of a block of
    triangles, so equal Morton codes occur): closest hit and any hit against brute force; then the same tree is
    refitted after a transform change (the device build leaves the refit inputs in place) and checked again."""
    rng = np.random.RandomState(11)
    n = 120000
    c = rng.rand(n, 1, 3) * 20.0
    v = (c + (rng.rand(n, 3, 3) - 0.5) * 0.5).reshape(-1, 3).astype(np.float32)
    idx = np.arange(3 * n, dtype=np.int32).reshape(n, 3)
    dup = 4000                                         # exact duplicates: identical boxes and Morton codes
    v2 = np.concatenate([v[:3 * dup], (rng.rand(3 * (80000 - dup), 3) * 20.0).astype(np.float32)])
    n2 = v2.shape[0] // 3
    idx2 = np.arange(3 * n2, dtype=np.int32).reshape(n2, 3)
    a, b = api.CreateMesh(v, 3 * n, 12, idx, 12, None, n), api.CreateMesh(v2, 3 * n2, 12, idx2, 12, None, n2)
    a.SetId(3)
    b.SetId(5)
    m = _rot_y(20.0, (0.5, -0.25, 1.0))
    b.SetTransform(m, np.linalg.inv(m))
    api.AttachShape(a)
    api.AttachShape(b)
    api.SetOption("acc.type", "hlbvh")
    api.SetOption("bvh.refit", 1.0)
    api.Commit()
    st = api.GetStats()
    assert st["last_build_on_device"] == 1 and st["num_triangles"] == n + n2 and st["num_levels"] <= 19
    rays = random_rays(20000, np.zeros(3), np.full(3, 20.0), 99, maxt=8.0)
    tris = port.mesh_triangles([(v, idx, 3, None), (v2, idx2, 5, m)])
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0, res          # duplicated triangles tie with different ids: smallest id wins in both
    assert res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))
    m2 = _rot_y(-35.0, (1.0, 0.5, -2.0))
    b.SetTransform(m2, np.linalg.inv(m2))
    api.Commit()
    st = api.GetStats()
    assert st["last_commit_refit"] == 1
    tris = port.mesh_triangles([(v, idx, 3, None), (v2, idx2, 5, m2)])
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res


def test_device_build_falls_back_below_four_triangles(api):
    api.SetOption("acc.type", "hlbvh")
    single_triangle(api)                 # commits: the device builder declines, the host builder takes over
    assert api.GetStats()["last_build_on_device"] == 0
    h = api.QueryIntersectionHost(one_ray((0.0, 0.0, -10.0), (0.0, 0.0, 1.0)))
    assert h["shapeid"][0] != -1


def test_axis_aligned_and_negative_zero_directions(api, cornell):
    """Directions with exact +0.0 / -0.0 components (directional-light shadow rays are like this):
    the sign used for the slab order must be the one used for the inverse direction."""
    rr.attach_scene(api, cornell)
    api.Commit()
    tris = port.scene_triangles(cornell)
    rng = np.random.RandomState(3)
    n = 6000
    lo, hi = cornell.world_bounds()
    rays = random_rays(n, lo + 0.1, hi - 0.1, 17)
    d = rays["d"][:, :3].copy()
    axis = rng.randint(0, 3, n)
    zero = np.where(rng.rand(n) < 0.5, np.float32(0.0), np.float32(-0.0))
    d[np.arange(n), axis] = zero
    second = (axis + 1 + rng.randint(0, 2, n)) % 3
    twice = rng.rand(n) < 0.3
    d[np.arange(n)[twice], second[twice]] = np.float32(-0.0)
    d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-6)
    rays["d"][:, :3] = d
    res = port.compare_hits(api.QueryIntersectionHost(rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0 and res["tie_ok"] == 0, res
    assert np.array_equal(api.QueryOcclusionHost(rays), port.occluded(tris, rays))


def test_cpp_adaptor_binary_on_device():
    """The engine-side C++ calling sequence through include/radeon_rays_b200.hpp (binary prebuilt
    against the reference's radeon_rays.h by __graft_entry__.build(), tests/cpp/adaptor_check.cpp)."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "adaptor_check")
    if not os.path.exists(exe):
        pytest.skip("adaptor_check not prebuilt (needs the reference headers)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr
