"""GPU suite, collected LAST on purpose (a failure here must not hide other tests):
rrb_query_intersection_noalloc (include/rrb.h, csrc/rrb_noalloc.cu) -- the closest-hit query whose scratch (work
counter + compacted active-ray list) lives in a static device array, so that the call allocates nothing.  No
counterpart in the reference.

Checked through the C ABI against the existing entry point rrb_query_intersection on the same device buffers:
byte-identical Intersection records (the tolerance of the repo's traversal tests: none), including the bytes that
must stay untouched (inactive rays, uvwt of misses) -- on two calls in a row with different ray sets and sizes, the
second smaller than the first, so stale scratch contents of the first call would show; the size limit is enforced
with a clear error; and a two-level world goes through the same entry point.
"""
import ctypes

import numpy as np
import pytest

from mcrt_b200 import rr, scene_library, layouts as L
from raygen import random_rays

pytestmark = pytest.mark.gpu


def ray_set(scene_bounds, n, seed, inactive_every):
    lo, hi = (np.asarray(b, np.float64) for b in scene_bounds)
    rays = random_rays(n, lo, hi, seed)
    rays["o"][:, 3] = 100.0
    if inactive_every:
        rays["extra"][::inactive_every, 1] = 0       # inactive rays: nothing may be written for them
    return rays


def sentinel_hits(n, value):
    h = np.zeros(n, dtype=L.INTERSECTION)
    h.view(np.uint8)[:] = value                       # every byte a known pattern: "untouched" is observable
    return h


def both_ways(api, rays, n_query=None):
    """The same device ray buffer through rrb_query_intersection and rrb_query_intersection_noalloc, each into its
    own sentinel-filled device hit buffer; returns the two host copies as bytes."""
    n = rays.shape[0]
    nq = n if n_query is None else n_query
    rb = api.CreateBuffer(n * 48, rays)
    h_old = api.CreateBuffer(n * 32, sentinel_hits(n, 0xA5))
    h_new = api.CreateBuffer(n * 32, sentinel_hits(n, 0xA5))
    try:
        api.QueryIntersection(rb, nq, h_old).Wait()
        api.QueryIntersectionNoAlloc(rb, nq, h_new)            # asynchronous on the API's stream ...
        out = []
        for hb in (h_old, h_new):
            arr, ev = api.MapBuffer(hb, rr.MAP_READ, 0, n * 32)    # ... which the map is ordered behind
            ev.Wait()
            out.append(np.array(arr, copy=True))
            api.UnmapBuffer(hb, arr)
        return out[0], out[1]
    finally:
        for b in (rb, h_old, h_new):
            api.DeleteBuffer(b)


def test_noalloc_query_equals_the_existing_entry_point_on_two_calls_in_a_row(cornell):
    api = rr.IntersectionApi.Create(0)
    try:
        rr.attach_scene(api, cornell)
        api.Commit()
        bounds = cornell.world_bounds()
        # call 1: 150,000 rays, every 3rd inactive; call 2 right after: 40,000 other rays, every 5th inactive,
        # only the first 33,333 queried (the tail of the hit buffer must stay untouched too)
        for n, seed, skip, nq in ((150000, 11, 3, None), (40000, 12, 5, 33333)):
            rays = ray_set(bounds, n, seed, skip)
            want, got = both_ways(api, rays, nq)
            assert np.array_equal(got, want), "noalloc query differs from rrb_query_intersection (n=%d)" % n
            rec = got.view(L.INTERSECTION)
            nq_ = n if nq is None else nq
            active = rays["extra"][:nq_, 1] != 0
            assert (rec["shapeid"][:nq_][active] >= 0).mean() > 0.5            # the kernel really ran
            assert (got.reshape(n, 32)[:nq_][~active] == 0xA5).all()            # inactive rays untouched
            assert (got.reshape(n, 32)[nq_:] == 0xA5).all()                     # beyond numrays untouched
    finally:
        rr.IntersectionApi.Delete(api)


def test_noalloc_query_rejects_sizes_beyond_the_static_scratch_and_uncommitted_worlds(cornell):
    api = rr.IntersectionApi.Create(0)
    try:
        lib = rr.load_library()
        with pytest.raises(rr.RadeonRaysError, match="Commit"):
            api.QueryIntersectionNoAlloc(0x1000, 16, 0x1000)
        rr.attach_scene(api, cornell)
        api.Commit()
        rb = api.CreateBuffer(48 * 64, ray_set(cornell.world_bounds(), 64, 1, 0))
        hb = api.CreateBuffer(32 * 64)
        with pytest.raises(rr.RadeonRaysError, match="RRB_NOALLOC_MAX_RAYS"):
            api.QueryIntersectionNoAlloc(rb, rr.NOALLOC_MAX_RAYS + 1, hb)
        assert lib.rrb_query_intersection_noalloc(api._h, rb.device_ptr, -1, hb.device_ptr) != 0
        assert lib.rrb_query_intersection_noalloc(api._h, None, 8, hb.device_ptr) != 0
        api.QueryIntersectionNoAlloc(rb, 0, hb)                 # zero rays: accepted, nothing launched
        api.QueryIntersectionNoAlloc(rb, 64, hb)                # and the limit itself is about numrays, not buffers
        arr, ev = api.MapBuffer(hb, rr.MAP_READ, 0, 32 * 64, dtype=L.INTERSECTION)
        ev.Wait()
        assert (arr["shapeid"] >= 0).any()
        api.UnmapBuffer(hb, arr)
        api.DeleteBuffer(rb)
        api.DeleteBuffer(hb)
    finally:
        rr.IntersectionApi.Delete(api)


def test_noalloc_query_on_a_two_level_world_and_at_the_size_limit(cornell):
    """force2level takes the object-space kernel variant; RRB_NOALLOC_MAX_RAYS rays fill the static list exactly."""
    api = rr.IntersectionApi.Create(0)
    try:
        rr.attach_scene(api, cornell)
        api.SetOption("bvh.force2level", 1.0)
        api.Commit()
        assert api.GetStats()["two_level"] == 1
        rays = ray_set(cornell.world_bounds(), rr.NOALLOC_MAX_RAYS, 21, 7)
        want, got = both_ways(api, rays)
        assert np.array_equal(got, want)
    finally:
        rr.IntersectionApi.Delete(api)
