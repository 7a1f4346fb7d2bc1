"""Parity helpers shared by the full-size GPU parity tests and tools/parity_report.py.

The bar (BASELINE.json north_star): per-sample radiance within 1e-4 relative of the oracle's with the identical
per-pixel RNG stream, *documented* exceptions only.  What can differ at all between the CUDA path and the strict
host build of the reference's kernels (oracle/_ref) is CUDA's libm (sinf / cosf / logf / powf / atanf / tanf are
within 2 ulp of glibc's, not bit-equal): a sampled direction or a disk-light sample can differ in its last bit.
That perturbation is harmless (relative radiance error ~1e-6) unless it flips a DISCRETE event later along the path:
the next ray lands on the other side of a geometric edge (different triangle, or hit / miss), a shadow ray's
visibility flips at an occluder's edge, or a path's `pdf ~ 0` / `black throughput` termination test flips.  After
such a flip the two paths are different samples and nothing bounds their difference.

So every pixel outside the tolerance must SHOW such a flip: both sides record, per bounce, the hit ids, whether a
shadow ray was traced and its visibility, whether the path went on and the sampled lobe flags
("PathTraceBuffer", include/ptb.h; oracle.ref.RefPathTracer.render_sample(trace=...)), and classify() names the
first bounce at which the records differ.  A pixel that leaves the tolerance with identical records at every
bounce is `unexplained` -- a shading bug -- and the tests allow none.
"""
import numpy as np

REL_TOL = 1e-4

# flags of a path-trace record (include/ptb.h)
F_SHADED, F_SHADOW, F_EMITTED, F_SPAWNED = 1, 2, 4, 8

CLASSES = ("closest_hit_flip", "shadow_visibility_flip", "shadow_ray_flip", "termination_flip", "lobe_flip",
           "emission_flip", "unexplained")


def rel_err(a, b):
    """Per-pixel max-channel relative error with an absolute floor for near-black pixels."""
    return (np.abs(a - b) / np.maximum(np.abs(b), 1e-3)).max(axis=1)


def same_triangle(scene, shape_a, prim_a, shape_b, prim_b):
    """E1: coplanar DUPLICATE triangles (the Cornell boxes carry one face twice, the file's own
    quirk) are hit at the identical t; the reference reports whichever its tree meets first, the CUDA walk the
    smallest (shape, prim).  Two records name the same surface point when their triangles have the same three
    world-space vertices; everything downstream (position, normal, material of the shared shape) is then equal."""
    out = np.zeros(shape_a.shape[0], bool)
    cand = np.nonzero(shape_a == shape_b)[0]          # duplicates live inside one shape
    for sid in np.unique(shape_a[cand]):
        if sid < 0 or sid >= len(scene.meshes):
            continue
        m = scene.meshes[int(sid)]
        sel = cand[shape_a[cand] == sid]
        T = m.indices.shape[0]
        sel = sel[(prim_a[sel] >= 0) & (prim_a[sel] < T) & (prim_b[sel] >= 0) & (prim_b[sel] < T)]
        if sel.size == 0:
            continue
        va = np.sort(m.positions[m.indices[prim_a[sel]]].reshape(len(sel), 3, 3).view([("x", "f4"), ("y", "f4"), ("z", "f4")]), axis=1)
        vb = np.sort(m.positions[m.indices[prim_b[sel]]].reshape(len(sel), 3, 3).view([("x", "f4"), ("y", "f4"), ("z", "f4")]), axis=1)
        out[sel] = (va == vb).all(axis=(1, 2))
    return out


def first_divergence(got_trace, want_trace, scene=None):
    """For every pixel: (bounce, class index) of the first discrete difference between the two records, or
    (-1, index of 'unexplained') when the records agree at every bounce.  Traces are (D, P, 4) int32.  With
    `scene`, hit ids that differ only by an E1 tie between duplicate triangles count as equal."""
    D, P, _ = got_trace.shape
    bounce = np.full(P, -1, np.int32)
    cls = np.full(P, CLASSES.index("unexplained"), np.int32)
    open_ = np.ones(P, bool)
    for b in range(D):
        g, w = got_trace[b], want_trace[b]
        gs, ws = (g[:, 2] & F_SHADED) != 0, (w[:, 2] & F_SHADED) != 0
        both = gs & ws
        ids_differ = both & ((g[:, 0] != w[:, 0]) | (g[:, 1] != w[:, 1]))
        if scene is not None and ids_differ.any():
            k = np.nonzero(ids_differ)[0]
            ids_differ[k[same_triangle(scene, g[k, 0], g[k, 1], w[k, 0], w[k, 1])]] = False
        tests = [
            ("closest_hit_flip", (gs != ws) | ids_differ),
            ("emission_flip", both & (((g[:, 2] ^ w[:, 2]) & F_EMITTED) != 0)),
            ("shadow_ray_flip", both & (((g[:, 2] ^ w[:, 2]) & F_SHADOW) != 0)),
            ("shadow_visibility_flip", both & ((g[:, 2] & w[:, 2] & F_SHADOW) != 0) & (g[:, 3] != w[:, 3])),
            ("termination_flip", both & (((g[:, 2] ^ w[:, 2]) & F_SPAWNED) != 0)),
            ("lobe_flip", both & ((g[:, 2] & w[:, 2] & F_SPAWNED) != 0) & ((g[:, 2] >> 8) != (w[:, 2] >> 8))),
        ]
        for name, m in tests:
            hit = open_ & m
            bounce[hit] = b
            cls[hit] = CLASSES.index(name)
            open_ &= ~hit
    return bounce, cls


def classify(got, want, got_trace, want_trace, pixels=None, tol=REL_TOL, scene=None):
    """Parity report of one frame.  got / want: (P, >=3) radiance; traces (D, P, 4); `pixels`: restrict to these
    global indices (tile partition).  Returns a dict of plain numbers (JSON-able)."""
    e = rel_err(got[:, :3], want[:, :3])
    sel = np.ones(e.shape[0], bool) if pixels is None else np.isin(np.arange(e.shape[0]), pixels)
    out_tol = sel & ~(e <= tol)              # NaN counts as out of tolerance
    bounce, cls = first_divergence(got_trace, want_trace, scene)
    diverged = sel & (bounce >= 0)
    exact = sel & (got[:, :3].view(np.uint32) == want[:, :3].view(np.uint32)).all(axis=1)
    rep = {
        "pixels": int(sel.sum()),
        "within_tol": int((sel & (e <= tol)).sum()),
        "out_of_tol": int(out_tol.sum()),
        "frac_within": float((sel & (e <= tol)).sum() / max(int(sel.sum()), 1)),
        "max_rel_err_of_undiverged": float(e[sel & ~diverged].max()) if (sel & ~diverged).any() else 0.0,
        "bit_identical": int(exact.sum()),              # pixels whose radiance equals the oracle's bit for bit
        "diverged_paths": int(diverged.sum()),          # paths with a discrete flip, inside the tolerance or not
        "diverged_by_class": {c: int((diverged & (cls == i)).sum()) for i, c in enumerate(CLASSES) if c != "unexplained"},
        "bounce0_record_mismatches": int((sel & (bounce == 0)).sum()),
        "by_class": {c: int((out_tol & (cls == i)).sum()) for i, c in enumerate(CLASSES)},
        "by_bounce": {int(b): int((out_tol & (bounce == b)).sum()) for b in np.unique(bounce[out_tol])},
        "mean_got": [float(x) for x in got[sel, :3].mean(0)],
        "mean_want": [float(x) for x in want[sel, :3].mean(0)],
    }
    rep["unexplained"] = rep["by_class"]["unexplained"]
    rep["unexplained_pixels"] = [int(i) for i in np.nonzero(out_tol & (cls == CLASSES.index("unexplained")))[0][:16]]
    return rep


def explain_divergences(scene, got_trace, want_trace, closest_rays, shadow_rays, pixels=None):
    """WHY the first record difference of every diverged pixel happened.  With the shading arithmetic bit-identical
    (csrc/pt_libm.h), both sides trace the identical ray at that bounce, so the only thing that can differ is the
    answer of the two TRAVERSALS: the reference walks Radeon Rays' BVH2 with an fp32 slab test that can cull a leaf
    whose triangle test would accept (SURVEY App. A E4 -- common for shadow rays that end on the light they were
    aimed at, t ~ tmax), and reports the first of several equal-t hits in its tree order (E1).  The arbiter is the
    brute-force restatement of the reference's triangle test over every triangle (oracle/port): a divergence is
    `reference_cull` when the CUDA answer equals the brute-force answer and the reference kernel's does not,
    `cuda_wrong` when it is the other way round -- a bug --, `other` when neither or when the first difference is
    not a traversal result at all.  closest_rays / shadow_rays: per-bounce RTRay arrays of the oracle run."""
    from oracle import port
    bounce, cls = first_divergence(got_trace, want_trace, scene)
    sel = bounce >= 0
    if pixels is not None:
        keep = np.zeros(sel.shape[0], bool)
        keep[pixels] = True
        sel &= keep
    out = {"reference_cull": 0, "cuda_wrong": 0, "other": 0, "examples": []}
    idx = np.nonzero(sel)[0]
    if idx.size == 0:
        return out
    tris = port.scene_triangles(scene)
    for kind, name in ((CLASSES.index("closest_hit_flip"), "closest"), (CLASSES.index("shadow_visibility_flip"), "shadow")):
        for b in np.unique(bounce[idx]):
            px = idx[(cls[idx] == kind) & (bounce[idx] == b)]
            if px.size == 0:
                continue
            g, w = got_trace[b][px], want_trace[b][px]
            if name == "closest":
                rays = np.ascontiguousarray(closest_rays[b][px])
                bf = port.intersect(tris, rays)
                bf_hit = bf["shapeid"] >= 0
                g_hit, w_hit = (g[:, 2] & F_SHADED) != 0, (w[:, 2] & F_SHADED) != 0

                def agrees(rec, rec_hit):
                    same = rec_hit == bf_hit
                    both = rec_hit & bf_hit
                    ids = (rec[:, 0] == bf["shapeid"]) & (rec[:, 1] == bf["primid"])
                    k = np.nonzero(both & ~ids)[0]
                    if k.size:      # E1: another triangle at the identical t (duplicate geometry)
                        ids[k[same_triangle(scene, rec[k, 0], rec[k, 1], bf["shapeid"][k], bf["primid"][k])]] = True
                    return same & (~both | ids)
                g_ok, w_ok = agrees(g, g_hit), agrees(w, w_hit)
            else:
                rays = np.ascontiguousarray(shadow_rays[b][px])
                bf_occ = port.occluded(tris, rays) != -1
                g_ok, w_ok = (g[:, 3] != 0) == bf_occ, (w[:, 3] != 0) == bf_occ
            out["reference_cull"] += int((g_ok & ~w_ok).sum())
            out["cuda_wrong"] += int((~g_ok & w_ok).sum())
            out["other"] += int((g_ok == w_ok).sum())
            for i in np.nonzero(~(g_ok & ~w_ok))[0][:6]:
                ex = {"pixel": int(px[i]), "bounce": int(b), "kind": name, "cuda": [int(x) for x in g[i]],
                      "reference": [int(x) for x in w[i]], "ray_o": [float(x) for x in rays["o"][i]],
                      "ray_d": [float(x) for x in rays["d"][i]], "ray_extra": [int(x) for x in rays["extra"][i]]}
                if name == "closest":
                    ex["brute_force"] = [int(bf["shapeid"][i]), int(bf["primid"][i]), float(bf["uvwt"][i, 3])]
                else:
                    ex["brute_force_occluded"] = bool(bf_occ[i])
                    h = port.intersect(tris, rays[i:i + 1])
                    ex["brute_force_closest"] = [int(h["shapeid"][0]), int(h["primid"][0]), float(h["uvwt"][0, 3])]
                out["examples"].append(ex)
    rest = idx[(cls[idx] != CLASSES.index("closest_hit_flip")) & (cls[idx] != CLASSES.index("shadow_visibility_flip"))]
    out["other"] += int(rest.size)
    for i in rest[:4]:
        out["examples"].append({"pixel": int(i), "bounce": int(bounce[i]), "kind": CLASSES[cls[i]]})
    return out


def assert_parity(rep, max_out_frac, what=""):
    """The test bar.  (1) Every pixel whose path shows no discrete divergence from the oracle's is BIT-IDENTICAL to
    it.  (2) Every divergence is a traversal answer on an identical ray where the CUDA walk agrees with the
    brute-force restatement of the reference's triangle test and the reference's BVH2 kernel does not (E4 / E1):
    none where the CUDA walk is the one that is wrong, none of any other kind.  (3) The fraction of pixels that
    leave 1e-4 because of (2) stays below `max_out_frac` (2x the observed one).  (4) Image means agree."""
    assert rep["unexplained"] == 0, "%s: %d pixels leave the tolerance without any discrete divergence: %s" % (
        what, rep["unexplained"], rep["unexplained_pixels"])
    assert rep["max_rel_err_of_undiverged"] == 0.0, (what, rep["max_rel_err_of_undiverged"])
    assert rep["bit_identical"] >= rep["pixels"] - rep["diverged_paths"], (what, rep["bit_identical"], rep["diverged_paths"])
    c = rep["causes"]
    assert c["cuda_wrong"] == 0 and c["other"] == 0, "%s: divergences not explained by the reference's own traversal: %s" % (what, c)
    assert c["reference_cull"] == rep["diverged_paths"], (what, c, rep["diverged_paths"])
    frac_out = rep["out_of_tol"] / max(rep["pixels"], 1)
    assert frac_out <= max_out_frac, "%s: %.6f of the pixels outside %g (allowed %.6f): %s" % (
        what, frac_out, REL_TOL, max_out_frac, rep["by_class"])
    assert np.allclose(rep["mean_got"], rep["mean_want"], rtol=2e-3), (what, rep["mean_got"], rep["mean_want"])


# ---------------------------------------------------------------------------------------------- drivers
def cuda_frames(scene, w, h, depth, frames, part=None, device=0, more_frames=(), images=None):
    """Fused wavefront through the C ABI with the diagnostic record on.  frames: iterable of frame indices.
    Returns ([radiance (P,4)], [trace (D,P,4)], final image).  `more_frames` are rendered and accumulated after
    `frames` with the record off (nothing kept but the image); `images`, a dict, receives the accumulated image after
    `frames` ("after_frames") and after `more_frames` ("final")."""
    from mcrt_b200 import rr, pt
    api = rr.IntersectionApi.Create(device)
    try:
        pl = pt.make_pt_pipeline(api, scene, w, h, depth)
        pl.ensure_size()
        if part is not None:
            pl.set_partition(*part)
        pl.set_path_trace(depth)
        integ = pl.passes[0]
        rad, tr = [], []
        for f in frames:
            integ.m_frameIndex = f
            pl.update()
            rad.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            tr.append(pl.path_trace(depth).copy())
        img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
        if images is not None:
            images["after_frames"] = img
        if len(more_frames):
            pl.set_path_trace(0)
            for f in more_frames:
                integ.m_frameIndex = f
                pl.update()
            img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
        if images is not None:
            images["final"] = img
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    return rad, tr, img


def oracle_frames(scene, w, h, depth, frames, pixels=None, keep_rays=False):
    """The compiled reference kernels (oracle/_ref) on the host, same frames; ([radiance], [trace]) -- with
    keep_rays also the list of the raw trace dicts (per-bounce "closest_rays" / "shadow_rays")."""
    from mcrt_b200 import camera
    from oracle import ref
    o = ref.RefOracle()
    rp = ref.RefPathTracer(o, scene, o.build_scene_bvh(scene), w, h)
    rad, tr, raw = [], [], []
    for f in frames:
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        t = {"keep_rays": True} if keep_rays else {}
        rad.append(rp.render_sample(cam, f, depth, trace=t, pixels=pixels).copy())
        tr.append(t["path_trace"])
        raw.append(t)
    return (rad, tr, raw) if keep_rays else (rad, tr)


# ---------------------------------------------------------------------------------------------- the benchmarked configs
def config(name):
    """(scene, width, height, depth, frames, partition) of BASELINE.json's configs as bench.py / tools/bench_configs.py
    render them.  Frame lists are bounded so that the oracle finishes in seconds to a few minutes on the host cores."""
    from mcrt_b200 import scene_library
    if name == "C1":      # in full: 512^2, 16 spp, depth 5
        return scene_library.cornell_box(), 512, 512, 5, list(range(16)), None
    if name == "C2":      # one full 1080p frame at 871,202 triangles, depth 8
        return scene_library.dragon_class(), 1920, 1080, 8, [0], None
    if name == "C3":      # one full 1080p frame with the 1024^2 textures bench.py uses
        return scene_library.sponza_class(1024), 1920, 1080, 5, [0], None
    if name == "C5":      # 10 M triangles at 4K: rank 0's tiles of a 64-way partition, frame 127 -- the seed
        # idx + (frame + 1) * W * H * (bounce + 1) wraps in uint32 from bounce 3 on (E14), indices stay global
        return scene_library.traversal_bench_scene(2236), 3840, 2160, 5, [127], (0, 64)
    raise KeyError(name)


def run_config(name, device=0, keep=None, more_frames=()):
    """Per-frame parity reports of one config: CUDA fused wavefront vs oracle/_ref.  `keep`, a dict, receives the
    CUDA path's accumulated image after the config's frames ("cuda_image") and after `more_frames` as well
    ("cuda_image_final"), the oracle's per-frame radiance of the config's frames ("oracle_radiance"), the mask of
    pixels whose path diverged in any of them ("diverged_any") and the oracle renderer ("oracle_renderer")."""
    from mcrt_b200 import multi_gpu
    scene, w, h, depth, frames, part = config(name)
    pixels = None if part is None else multi_gpu.partition_pixels(w, h, part[0], part[1])
    images = {}
    got, got_tr, img = cuda_frames(scene, w, h, depth, frames, part=part, device=device, more_frames=more_frames,
                                   images=images)
    if keep is not None:
        keep["cuda_image"], keep["cuda_image_final"], keep["oracle_radiance"] = images["after_frames"], images["final"], []
        keep["diverged_any"] = np.zeros(w * h, bool)
    reps = []
    # the oracle one frame at a time: its per-bounce ray copies (for the brute-force arbitration) are large
    from mcrt_b200 import camera
    from oracle import ref
    o = ref.RefOracle()
    rp = ref.RefPathTracer(o, scene, o.build_scene_bvh(scene), w, h)
    for i, f in enumerate(frames):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        t = {"keep_rays": True}
        want = rp.render_sample(cam, f, depth, trace=t, pixels=pixels)
        if keep is not None:
            keep["oracle_radiance"].append(want.copy())
            keep["diverged_any"] |= first_divergence(got_tr[i], t["path_trace"], scene)[0] >= 0
            keep["oracle_renderer"] = rp
        r = classify(got[i], want, got_tr[i], t["path_trace"], pixels=pixels, scene=scene)
        r["causes"] = explain_divergences(scene, got_tr[i], t["path_trace"], t["closest_rays"], t["shadow_rays"], pixels)
        r["config"], r["frame"] = name, f
        if pixels is not None:      # pixels outside the partition stay exactly black on both sides
            outside = np.ones(w * h, bool)
            outside[pixels] = False
            r["outside_partition_nonzero"] = int(np.count_nonzero(got[i][outside, :3])) + int(np.count_nonzero(want[outside, :3]))
        reps.append(r)
    return reps
