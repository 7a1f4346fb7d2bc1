"""CPU suite: the intersection oracle (oracle/port) pinned against (a) the known-answer cases of the
reference's own tests (UnitTest/radeon_rays_apitest_cl.h) and (b) the reference's own kernel,
compiled here (oracle/_ref: intersect_bvh2_lds.cl over Radeon Rays' Bvh2)."""
import numpy as np
import pytest

from mcrt_b200 import layouts as L
from oracle import port
from raygen import random_rays, camera_rays

# ApiBackendOpenCL fixture (radeon_rays_apitest_cl.h:82-99)
TRI_VERTS = np.array([[-1, -1, 0], [0, 1, 0], [1, -1, 0]], np.float32)
TRI_IDX = np.array([[0, 1, 2]], np.int32)


def one_ray(o, d, maxt=10000.0):
    r = L.make_rays(1)
    r["o"][0, :3] = o
    r["o"][0, 3] = maxt
    r["d"][0, :3] = d
    return r


def test_struct_sizes_match_compiled_reference(ref_oracle):
    import ctypes
    a = (ctypes.c_int * 16)()
    n = ref_oracle.lib.ref_pt_struct_sizes(a, 16)
    assert list(a)[:n] == [L.SHAPE.itemsize, L.MATERIAL.itemsize, L.LIGHT.itemsize, L.THROUGHPUT.itemsize, 192, 240,
                           L.CAMERA.itemsize, L.RAYDIFF.itemsize, L.TEXDESC.itemsize, L.RAY.itemsize,
                           L.INTERSECTION.itemsize]


def test_intersection_1ray_known_answer():
    """Intersection_1Ray (radeon_rays_apitest_cl.h:203-249): o=(0,0,-10) d=(0,0,1) hits the mesh."""
    tris = port.mesh_triangles([(TRI_VERTS, TRI_IDX, 7, None)])
    h = port.intersect(tris, one_ray((0, 0, -10), (0, 0, 1)))
    assert h["shapeid"][0] == 7 and h["primid"][0] == 0
    assert h["uvwt"][0, 3] == np.float32(10.0)


def test_inactive_ray_not_written():
    """Intersection_1Ray_Active (:413-489): an inactive ray leaves the pre-filled hit untouched."""
    tris = port.mesh_triangles([(TRI_VERTS, TRI_IDX, 7, None)])
    r = one_ray((0, 0, -10), (0, 0, 1))
    r["extra"][0, 1] = 0
    pre = np.zeros(1, dtype=L.INTERSECTION)
    pre["shapeid"], pre["primid"] = 1234, 5678
    h = port.intersect(tris, r, pre.copy())
    assert h.tobytes() == pre.tobytes()
    occ = port.occluded(tris, r, np.array([99], np.int32))
    assert occ[0] == 99


def test_mask_skips_shape():
    """Intersection_1Ray_Masked (:251-300): a ray whose mask equals the shape id skips that shape."""
    tris = port.mesh_triangles([(TRI_VERTS, TRI_IDX, 7, None)])
    r = one_ray((0, 0, -10), (0, 0, 1))
    r["extra"][0, 0] = 7
    assert port.intersect(tris, r)["shapeid"][0] == -1
    assert port.occluded(tris, r)[0] == -1


def test_hit_at_exactly_maxt_is_rejected():
    """`t < closest_t` after `temp <= t_max` (intersect_bvh2_lds.cl:174): t == maxt is a miss."""
    tris = port.mesh_triangles([(TRI_VERTS, TRI_IDX, 1, None)])
    assert port.intersect(tris, one_ray((0, 0, -10), (0, 0, 1), maxt=10.0))["shapeid"][0] == -1
    assert port.intersect(tris, one_ray((0, 0, -10), (0, 0, 1), maxt=10.000001))["shapeid"][0] == 1
    assert port.occluded(tris, one_ray((0, 0, -10), (0, 0, 1), maxt=10.0))[0] == -1


def test_transformed_mesh():
    """Intersection_1Ray_Transformed (:556): a translated mesh is hit where it was moved to."""
    m = np.eye(4, dtype=np.float32)
    m[0, 3] = 5.0
    tris = port.mesh_triangles([(TRI_VERTS, TRI_IDX, 3, m)])
    assert port.intersect(tris, one_ray((0, 0, -10), (0, 0, 1)))["shapeid"][0] == -1
    assert port.intersect(tris, one_ray((5, 0, -10), (0, 0, 1)))["shapeid"][0] == 3


@pytest.mark.parametrize("n", [1, 100, 1000, 10000])
def test_port_matches_reference_kernel_on_cornell(ref_oracle, cornell, n):
    """ApiConformanceCL (radeon_rays_conformance_test_cl.h:118-560) with the roles swapped: the
    reference's BVH kernel is the truth, the brute-force port is what is being pinned.  Criterion is
    stricter than the reference's (same shapeid, dt^2 <= 1e-5): bit-equal t, and identical ids
    except where the reference's first-found winner lies in the tied set."""
    lo, hi = cornell.world_bounds()
    rays = random_rays(n, lo - 1.0, hi + 1.0, seed=0xABCDEF12 % (2 ** 31))
    nodes = ref_oracle.build_scene_bvh(cornell)
    tris = port.scene_triangles(cornell)
    ref_hits = ref_oracle.intersect(nodes, rays)
    ora_hits = port.intersect(tris, rays)
    res = port.compare_hits(ref_hits, ora_hits, tris, rays)
    assert res["bad"] == 0, res
    ref_occ = ref_oracle.occluded(nodes, rays)
    ora_occ = port.occluded(tris, rays)
    assert np.array_equal(ref_occ, ora_occ)


def test_port_matches_reference_kernel_on_primary_rays(ref_oracle, cornell):
    """Camera rays through the Cornell box hit shared edges and corners all the time: this is the
    tie-heavy case."""
    rays, _ = camera_rays(cornell, 128, 128)
    nodes = ref_oracle.build_scene_bvh(cornell)
    tris = port.scene_triangles(cornell)
    res = port.compare_hits(ref_oracle.intersect(nodes, rays), port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0, res
    assert res["exact"] + res["tie_ok"] + res["b_farther"] == res["active"]


def test_rng_restatement_known_values():
    """wang_hash / xorshift / randFloat (rng.cl:48-66,104-107) against hand-evaluated values."""
    lib = port.load()
    assert lib.oracle_wang_hash(0) == 3232319850
    assert lib.oracle_wang_hash(1) == 663891101
    import ctypes
    s = ctypes.c_uint32(0)
    u = lib.oracle_rand_float(ctypes.byref(s))
    # xorshift(0): s = 2463534242; s ^= s<<13; s ^= s>>17; s ^= s<<5
    x = 2463534242
    x ^= (x << 13) & 0xFFFFFFFF
    x ^= x >> 17
    x ^= (x << 5) & 0xFFFFFFFF
    assert s.value == x
    assert u == np.float32(x) / np.float32(4294967296.0)
