"""GPU suite: BDPT per STRATEGY.  The image-level BDPT tests compare sums that
both sides build with float atomics in an unspecified order; here the comparison is made before any summation:
every (pixel, t, s) connection record of the CUDA path ("BDPTConnection*" buffers, include/ptb.h) against the
reference's own connection buffers -- tempRadianceBuffer as PrepareConnections leaves it (BDPT.cl:460-652), the s = 0
entries as ConnectVertices fills them in (:722-730), connectionVisibilities as Radeon Rays answers them -- on C4's
scene (Cornell box + glass and mirror spheres, depth 5), two frames (the second reads the first one's sampled light
vertices, E12).

The reference numbers a pixel's connections consecutively over the (t, s) loop of ITS vertex counts
(BDPT.cl:708-719: skipped strategies do not advance the index), so the slot of (t, s) depends on the pixel:
reference_slots() restates that loop.
"""
import numpy as np
import pytest

from mcrt_b200 import rr, pt, camera, scene_library
from oracle import port, ref

pytestmark = pytest.mark.gpu
W = H = 256
D = 5
REL_TOL = 1e-4
BLACK_EPS = np.float32(1e-6)        # isNotBlack (bxdfs.cl): what both sides treat as a contribution at all


def reference_slots(cam_counts, light_counts, depth):
    """{(t, s): int32[P] slot of that strategy in the pixel's run of maxConnections entries, -1 where the pixel's
    subpaths are too short} -- the loop of PrepareConnections / ConnectVertices (BDPT.cl:497-508, 708-719)."""
    slot = np.zeros(cam_counts.shape[0], np.int32)
    out = {}
    for t in range(1, depth + 3):
        for s in range(0, depth + 2):
            d = t + s - 2
            if (t == 1 and s == 1) or d < 0 or d > depth:
                continue
            valid = (t <= cam_counts) & (s <= light_counts)
            out[(t, s)] = np.where(valid, slot, -1).astype(np.int32)
            slot += valid
    return out


def render_both(scene, frames):
    """[(cuda records, cuda radiance, oracle trace, oracle connection rays)] per frame."""
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_bdpt_pipeline(api, scene, W, H, D)
        got = []
        for _ in range(frames):
            pl.update()
            got.append((pl.bdpt_connections(), pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy()))
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    o = ref.RefOracle()
    rb = ref.RefBDPT(o, scene, o.build_scene_bvh(scene), W, H, D)
    out = []
    for f in range(frames):
        cam = camera.make_camera(scene.camera, W, H, camera.taa_pixel_offset(f))
        tr = {}
        rb.render_sample(cam, f, trace=tr)
        out.append((got[f][0], got[f][1], tr, rb.conn_rays.copy()))
    return out, rb.max_conn


def test_every_connection_record_equals_the_references_connection_buffers():
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = scene_library.cornell_box(True)
    frames, max_conn = render_both(scene, 2)
    tris = port.scene_triangles(scene)
    P = W * H
    for f, (rec, radiance, tr, conn_rays) in enumerate(frames):
        slots = reference_slots(tr["cam_counts"], tr["light_counts"], D)
        n = rec["pixel"].shape[0]
        assert n > P                                              # several strategies per pixel survive
        # ---- the reference's record of every CUDA record
        slot = np.full(n, -1, np.int64)
        for (t, s), tab in slots.items():
            m = (rec["t"] == t) & (rec["s"] == s)
            slot[m] = tab[rec["pixel"][m]]
        assert (slot >= 0).all(), "a CUDA record names a strategy the reference's vertex counts do not allow"
        flat = rec["pixel"].astype(np.int64) * max_conn + slot
        assert np.unique(flat).size == n                          # one record per (pixel, t, s)
        is_s0 = rec["s"] == 0
        want_C = np.where(is_s0[:, None], tr["connected"][flat, :3], tr["unweighted"][flat, :3])
        err = (np.abs(rec["C"] - want_C) / np.maximum(np.abs(want_C), 1e-6)).max(axis=1)
        exact = (rec["C"].view(np.uint32) == np.ascontiguousarray(want_C).view(np.uint32)).all(axis=1)
        bad = err > REL_TOL
        print("frame %d: %d records, %.6f bit-identical, %d outside 1e-4 (max rel err %.3g)" % (
            f, n, exact.mean(), int(bad.sum()), float(err.max())))
        assert bad.mean() <= 1e-4, "unweighted contributions differ: %d of %d records, e.g. t,s = %s" % (
            int(bad.sum()), n, sorted(set(zip(rec["t"][bad].tolist(), rec["s"][bad].tolist())))[:6])
        # ---- and the other way round: every non-black connection of the reference is a CUDA record
        ref_live = np.zeros(P * max_conn, bool)
        for (t, s), tab in slots.items():
            px = np.nonzero(tab >= 0)[0]
            idx = px.astype(np.int64) * max_conn + tab[px]
            if s == 0:
                ref_live[idx] = (tr["connected"][idx, :3] > BLACK_EPS).any(axis=1)
            else:
                ref_live[idx] = tr["conn_active"][idx] & (tr["unweighted"][idx, :3] > BLACK_EPS).any(axis=1)
        have = np.zeros(P * max_conn, bool)
        have[flat] = True
        missing, extra = int((ref_live & ~have).sum()), int((have & ~ref_live).sum())
        print("frame %d: reference connections without a CUDA record %d, CUDA records the reference leaves black %d of %d" % (
            f, missing, extra, int(ref_live.sum())))
        assert missing <= 1e-4 * ref_live.sum() and extra <= 1e-4 * ref_live.sum()
        # ---- visibility: identical rays, so a difference is a traversal answer; the brute force decides whose
        rayed = ~is_s0 & have[flat] & ref_live[flat]
        ref_vis = tr["conn_vis"][flat] == -1                      # -1 = nothing hit (BDPT.cl:735)
        differ = rayed & (ref_vis != rec["visible"])
        print("frame %d: visibility differs on %d of %d connection rays" % (f, int(differ.sum()), int(rayed.sum())))
        assert differ.sum() <= 1e-3 * rayed.sum()
        if differ.any():
            rays = np.ascontiguousarray(conn_rays[flat[differ]])
            brute_vis = port.occluded(tris, rays) == -1
            assert np.array_equal(brute_vis, rec["visible"][differ]), \
                "connection rays on which the CUDA walk, not the reference's BVH2 cull, disagrees with the brute force"
        # ---- the records ARE the image: sum of C * weight over the records, by target pixel
        img = np.zeros((P, 3), np.float64)
        np.add.at(img, rec["target"], rec["C"].astype(np.float64) * rec["weight"].astype(np.float64)[:, None])
        assert np.allclose(img, radiance[:, :3], rtol=2e-5, atol=1e-6)
        assert (rec["weight"][~rec["visible"]] == 0).all() and (rec["weight"][rec["visible"]] >= 0).all()
        assert (rec["weight"] <= 1.0).all()
