"""CPU suite: the committed golden vectors (tests/golden/, made by tools/make_golden.py from the
compiled reference) against (a) the oracle port and (b) a fresh run of the compiled reference."""
import os

import numpy as np

from mcrt_b200 import layouts as L, camera
from oracle import port

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_port_matches_golden_intersections(cornell):
    g = np.load(os.path.join(GOLDEN, "cornell_isect_4096.npz"))
    rays = g["rays"].view(L.RAY)
    tris = port.scene_triangles(cornell)
    ref_hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
    ref_hits["shapeid"], ref_hits["primid"], ref_hits["uvwt"] = g["shapeid"], g["primid"], g["uvwt"]
    res = port.compare_hits(ref_hits, port.intersect(tris, rays), tris, rays)
    assert res["bad"] == 0, res
    assert np.array_equal(port.occluded(tris, rays), g["occluded"])


def test_golden_is_reproducible_from_compiled_reference(ref_oracle, cornell):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "tools"))
    import make_golden
    fresh = make_golden.cornell_pt()
    g = np.load(os.path.join(GOLDEN, "cornell_pt_64x64_d5.npz"))
    for k in g.files:
        assert np.array_equal(fresh[k], g[k]), k


def test_bdpt_golden_is_reproducible_from_compiled_reference(ref_oracle):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "tools"))
    import make_golden
    fresh = make_golden.cornell_spheres_bdpt()
    g = np.load(os.path.join(GOLDEN, "cornell_spheres_bdpt_64x64_d5.npz"))
    for k in g.files:      # float atomics: summation order is not fixed (BDPT.cl:654-669)
        assert np.allclose(fresh[k], g[k], rtol=1e-5, atol=1e-7), k
    # BDPT and PT estimate the same image
    assert 0.05 < g["radiance_f0"][:, 0].mean() < 2.0


def test_golden_image_is_plausible():
    g = np.load(os.path.join(GOLDEN, "cornell_pt_64x64_d5.npz"))
    r = g["radiance_f0"]
    assert r.shape == (64 * 64, 4) and np.isfinite(r).all()
    assert (g["primary_shapeid"] >= 0).mean() > 0.95
    # the light (shape 7) is seen directly and emits (17, 12, 4)
    lit = g["primary_shapeid"] == 7
    assert lit.any() and np.allclose(r[lit][:, :3], [17.0, 12.0, 4.0])
    assert 0.05 < r[:, 0].mean() < 2.0


def test_taa_offsets_follow_sobol_dims_0_and_1():
    # dimension 0 = van der Corput, dimension 1 = second Sobol dimension; lerp(-2, 2, u)
    assert camera.taa_pixel_offset(0) == (-2.0, -2.0)
    assert camera.taa_pixel_offset(1) == (0.0, 0.0)
    assert camera.taa_pixel_offset(2) == (-1.0, 1.0)
    assert camera.taa_pixel_offset(3) == (1.0, -1.0)
