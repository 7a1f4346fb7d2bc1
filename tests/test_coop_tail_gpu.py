"""GPU suite: the cooperative tail of the traversal kernels (ptb_set_cooperative_tail, csrc/pt_wavefront.cuh).  A ray
that is still being walked after its launch's queue ran dry is split over the idle lanes of its warp, each lane
restricted to an interval of t; the warp's smallest (t, shape, prim) must be the single walk's answer.  With
min_iters = 1 every ray that outlives the queue is split, several times over, in all three traversal kernels
(closest hit, any hit, mixed), so the equality below exercises the interval logic on hundreds of thousands of rays:
boundaries (a hit exactly on one), ties between duplicate triangles (the Cornell file carries one face twice),
shadow rays that end on their emitter (t ~ tmax), rays that miss everything."""
import numpy as np
import pytest

import parity_util as pu
from mcrt_b200 import rr, pt, scene_library
from oracle import ref

pytestmark = pytest.mark.gpu


def render(scene, w, h, depth, frames, coop, halves=2, partition=None):
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_pt_pipeline(api, scene, w, h, depth)
        pl.ensure_size()
        pl.set_overlap(halves)
        pl.set_cooperative_tail(coop)
        if partition is not None:
            pl.set_partition(*partition)
        pl.set_path_trace(depth)
        rad, tr = [], []
        for _ in range(frames):
            pl.update()
            rad.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            tr.append(pl.path_trace(depth).copy())
        img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
        cnt = pl.counters()
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    return rad, tr, img, cnt


@pytest.mark.parametrize("case", ["cornell", "cornell_spheres", "sponza", "knot", "tiles"])
def test_cooperative_tail_equals_one_ray_per_lane_bitwise(case):
    scene, w, h, depth, kw = {
        "cornell": (scene_library.cornell_box(), 256, 256, 5, {}),
        "cornell_spheres": (scene_library.cornell_box(True), 192, 192, 6, {"halves": 1}),
        "sponza": (scene_library.sponza_class(64), 384, 240, 5, {}),
        "knot": (scene_library.dragon_class(240, 240), 320, 200, 4, {}),
        "tiles": (scene_library.sponza_class(64), 512, 320, 4, {"partition": (2, 8)}),     # small launches: mostly tail
    }[case]
    base = render(scene, w, h, depth, 3, 0, **kw)
    for iters in (1, 6):
        got = render(scene, w, h, depth, 3, iters, **kw)
        for f in range(3):
            assert np.array_equal(base[0][f].view(np.uint32), got[0][f].view(np.uint32)), (case, iters, f)
            assert np.array_equal(base[1][f], got[1][f]), (case, iters, f)
        assert np.array_equal(base[2].view(np.uint32), got[2].view(np.uint32))
        for k in ("samples", "closest_rays", "shadow_rays", "shaded_vertices"):
            assert base[3][k] == got[3][k], (case, iters, k)


def test_cooperative_tail_keeps_parity_with_the_reference_kernels(cornell):
    """And against the oracle directly: the per-sample classification of tests/parity_util.py with every tail ray split."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    w = h = 192
    depth = 5
    got, got_tr, _, _ = render(cornell, w, h, depth, 2, 1)
    want, want_tr, raw = pu.oracle_frames(cornell, w, h, depth, [0, 1], keep_rays=True)
    for f in range(2):
        rep = pu.classify(got[f], want[f], got_tr[f], want_tr[f], scene=cornell)
        rep["causes"] = pu.explain_divergences(cornell, got_tr[f], want_tr[f], raw[f]["closest_rays"], raw[f]["shadow_rays"])
        pu.assert_parity(rep, 2e-3, "cooperative tail, frame %d" % f)
