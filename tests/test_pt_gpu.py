"""GPU suite: the CUDA path-tracer stages against the compiled reference kernels (oracle/_ref) and the
committed golden fixtures generated from them (tests/golden/, tools/make_golden.py).

Bar (BASELINE.json north_star): per-sample radiance within 1e-4 relative error with the identical
per-pixel RNG stream.  fp32 differences come only from CUDA's libm (sinf/cosf/logf are within 2 ulp
of glibc's, not bit-equal), which perturbs sampled directions in the last bit; a path whose ray then
lands on the other side of a geometric edge diverges completely, so the assertion is on the
fraction of pixels inside the tolerance plus the image mean.
"""
import os

import numpy as np
import pytest

from mcrt_b200 import rr, pt, camera, scene_library, layouts as L
from oracle import ref

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-4


def rel_err(a, b):
    """Per-pixel max-channel relative error with an absolute floor for near-black pixels."""
    return (np.abs(a - b) / np.maximum(np.abs(b), 1e-3)).max(axis=1)


@pytest.fixture()
def api():
    a = rr.IntersectionApi.Create(0)
    yield a
    rr.IntersectionApi.Delete(a)


def render_frames(api, scene, w, h, depth, frames, fused=True):
    pl = pt.make_pt_pipeline(api, scene, w, h, depth, fused=fused)
    out = []
    for _ in range(frames):
        pl.update()
        out.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
    img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
    cnt = pl.counters()
    pl.scene.destroy()
    return out, img, cnt


def ref_frames(scene, w, h, depth, frames):
    o = ref.RefOracle()
    nodes = o.build_scene_bvh(scene)
    rp = ref.RefPathTracer(o, scene, nodes, w, h)
    out = []
    for f in range(frames):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        out.append(rp.render_sample(cam, f, depth).copy())
    return out, rp


def check_against(got, want, min_frac):
    e = rel_err(got[:, :3], want[:, :3])
    frac = float((e <= REL_TOL).mean())
    assert frac >= min_frac, "only %.5f of pixels within %g (max %.3g)" % (frac, REL_TOL, e.max())
    m_got, m_want = got[:, :3].mean(0), want[:, :3].mean(0)
    assert np.allclose(m_got, m_want, rtol=2e-2), (m_got, m_want)
    return frac


@pytest.mark.parametrize("fused", [True, False])
def test_cornell_per_sample_radiance_vs_reference(api, cornell, fused):
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    w = h = 128
    got, _, _ = render_frames(api, cornell, w, h, 5, 3, fused=fused)
    want, _ = ref_frames(cornell, w, h, 5, 3)
    for f in range(3):
        check_against(got[f], want[f], 0.995)


def test_fused_wavefront_equals_stagewise_bitwise(api, cornell):
    """The queue-based wavefront and the reference-shaped stage-by-stage dispatch run the same
    arithmetic in the same order: per-sample radiance must be bit-identical."""
    a, img_a, cnt_a = render_frames(api, cornell, 96, 64, 5, 2, fused=True)
    api2 = rr.IntersectionApi.Create(0)
    try:
        b, img_b, _ = render_frames(api2, cornell, 96, 64, 5, 2, fused=False)
    finally:
        rr.IntersectionApi.Delete(api2)
    for f in range(2):
        assert np.array_equal(a[f].view(np.uint32)[:, :3], b[f].view(np.uint32)[:, :3])
    assert np.array_equal(img_a.view(np.uint32)[:, :3], img_b.view(np.uint32)[:, :3])
    assert cnt_a["samples"] == 2 * 96 * 64 and cnt_a["closest_rays"] > cnt_a["samples"]


def test_golden_cornell(api, cornell):
    """Committed vectors from the compiled reference (64x64, depth 5, frames 0 and 1)."""
    path = os.path.join(GOLDEN, "cornell_pt_64x64_d5.npz")
    g = np.load(path)
    got, _, _ = render_frames(api, cornell, 64, 64, 5, 2)
    for f in range(2):
        check_against(got[f], g["radiance_f%d" % f], 0.99)


def test_depth1_direct_light_only_matches_tightly(api, cornell):
    """maxDepth = 1: one vertex, NEE only, no sampled directions re-traced -> no chaotic divergence;
    every pixel must be inside the tolerance."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    got, _, _ = render_frames(api, cornell, 128, 128, 1, 1)
    want, _ = ref_frames(cornell, 128, 128, 1, 1)
    e = rel_err(got[0][:, :3], want[0][:, :3])
    assert (e <= REL_TOL).mean() >= 0.9995, e.max()


def test_uber_material_scene_vs_reference(api):
    """Glossy dielectric BRDF + glossy BTDF, directional + disk lights (the C2 material set) on a
    coarse knot: exercises Trowbridge-Reitz sampling / evaluation and both delta light types."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = scene_library.dragon_class(60, 24)
    w, h = 160, 96
    got, _, _ = render_frames(api, scene, w, h, 4, 2)
    want, _ = ref_frames(scene, w, h, 4, 2)
    for f in range(2):
        check_against(got[f], want[f], 0.97)


def test_accumulation_matches_reference_formula(api, cornell):
    """ReconstructionPass: clamp to [0,1000], triangle-filter weight of the frame's jitter, running
    weighted mean (reconstruction.cl:6-60), recomputed on the host from the per-sample radiance."""
    frames = 4
    got, img, _ = render_frames(api, cornell, 64, 64, 3, frames)
    acc = np.zeros_like(got[0], dtype=np.float32)
    wsum = np.float32(0.0)
    for f in range(frames):
        ox, oy = camera.taa_pixel_offset(f)
        wgt = np.float32(max(0.0, np.float32(2.0) - abs(np.float32(ox)))) * np.float32(max(0.0, np.float32(2.0) - abs(np.float32(oy))))
        r = np.clip(got[f], 0.0, 1000.0).astype(np.float32)
        acc = (acc + r * wgt).astype(np.float32) if f else (r * wgt).astype(np.float32)
        wsum = np.float32(wsum + wgt) if f else wgt
    want = (acc / wsum).astype(np.float32)
    assert np.allclose(img[:, :3], want[:, :3], rtol=1e-6, atol=1e-7)


def test_tile_partition_union_is_bit_identical(api, cornell):
    """Multi-GPU partition A: rendering the tiles of part 0/2 and part 1/2 separately and
    summing the (zero elsewhere) buffers reproduces the single-context image bit for bit."""
    w, h = 192, 128
    full, _, _ = render_frames(api, cornell, w, h, 4, 1)
    parts = []
    for part in range(2):
        a2 = rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a2, cornell, w, h, 4)
            pl.ensure_size()
            pl.set_partition(part, 2)
            pl.update()
            parts.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            pl.scene.destroy()
        finally:
            rr.IntersectionApi.Delete(a2)
    merged = parts[0] + parts[1]
    assert np.array_equal(merged.view(np.uint32)[:, :3], full[0].view(np.uint32)[:, :3])


def test_textured_normal_mapped_scene_vs_reference(api):
    """C3 geometry (262,144 triangles, 24 textured + normal-mapped materials, two alpha-masked; point +
    directional + triangle-mesh area light) with 64^2 textures: texture fetch, normal mapping, opacity
    lobe and the three light types against the compiled reference."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = scene_library.sponza_class(64)
    w, h = 240, 136
    got, _, _ = render_frames(api, scene, w, h, 5, 2)
    want, _ = ref_frames(scene, w, h, 5, 2)
    for f in range(2):
        check_against(got[f], want[f], 0.97)


def test_sobol_sampler_variant_vs_reference(api, cornell):
    """samplers.cl's other compile-time variant (RT_SAMPLER_SOBOL): the compiled reference built with
    -DRT_SAMPLER=0 against the run-time switch ptb_set_sampler, same generator matrices."""
    from mcrt_b200 import sobol
    if not ref.available(sobol=True):
        pytest.skip("oracle/_ref Sobol variant not built")
    table = sobol.matrices()
    w = h = 96
    pl = pt.make_pt_pipeline(api, cornell, w, h, 4)
    pl.set_sampler(sobol.RT_SAMPLER_SOBOL, table)
    got = []
    for _ in range(3):
        pl.update()
        got.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
    pl.set_sampler(sobol.RT_SAMPLER_RANDOM)
    pl.update()
    rnd = pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy()
    pl.scene.destroy()
    o = ref.RefOracle(sobol_table=table)
    rp = ref.RefPathTracer(o, cornell, o.build_scene_bvh(cornell), w, h)
    for f in range(3):
        cam = camera.make_camera(cornell.camera, w, h, camera.taa_pixel_offset(f))
        check_against(got[f], rp.render_sample(cam, f, 4), 0.995)
    # and it is a different estimate from the random sampler's frame 3
    assert not np.array_equal(rnd, got[2])
    with pytest.raises(pt.PathTracerError):
        pl2 = pt.make_pt_pipeline(api, cornell, 32, 32, 2)
        try:
            pl2.set_sampler(sobol.RT_SAMPLER_SOBOL)      # no table given
        finally:
            pl2.scene.destroy()


def test_all_five_pixel_filters_vs_reference_reconstruction_kernel(api, cornell):
    """ReconstructionPass with box / triangle / Gaussian / Mitchell / Lanczos-sinc filters (filters.cl:12-69)
    against the reference's own ReconstructionPassAllFilters kernel compiled for the host: the per-sample
    radiance comes from the CUDA path, both accumulate it over 4 jittered frames."""
    import ctypes
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    lib = ref.RefOracle().lib
    lib.ref_reconstruction_all_filters.argtypes = [ctypes.c_int] * 3 + [ctypes.c_void_p] * 4
    w, h, frames = 64, 48, 4
    P = w * h
    images = {}
    samples = None
    for ftype in range(5):
        a2 = rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a2, cornell, w, h, 3)
            pl.GI.filterType = ftype
            rad = []
            for _ in range(frames):
                pl.update()
                rad.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            images[ftype] = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
            props0 = pl.GI.filter_props()
            pl.scene.destroy()
        finally:
            rr.IntersectionApi.Delete(a2)
        if samples is None:
            samples = rad
        else:      # the filter only changes the weights, never the samples
            assert all(np.array_equal(x, y) for x, y in zip(samples, rad))
    weighted = np.zeros((5 * P, 4), np.float32)
    weights = np.zeros(5 * P, np.float32)
    for f in range(frames):
        ox, oy = camera.taa_pixel_offset(f)
        props = np.array([props0.radius[0], props0.radius[1], props0.mitchell_b, props0.mitchell_c,
                          props0.lanczos_sinc_tau, props0.gaussian_alpha, props0.gaussian_exp_x, props0.gaussian_exp_y,
                          ox, oy], np.float32)
        r = samples[f].copy()
        lib.ref_reconstruction_all_filters(w, h, f, props.ctypes.data, r.ctypes.data, weighted.ctypes.data,
                                           weights.ctypes.data)
    for ftype in range(5):
        want = weighted[ftype * P:(ftype + 1) * P] / weights[ftype * P:(ftype + 1) * P, None]
        ok = np.isfinite(want).all(axis=1)      # a filter whose weights sum to 0 gives 0/0 in both
        assert ok.mean() > 0.99
        assert np.allclose(images[ftype][ok, :3], want[ok, :3], rtol=2e-6, atol=1e-7), ftype


def test_denoise_and_tonemap_passes_vs_reference_kernels(api, cornell):
    """RTDenoisePass + RTToneMappingPass on device buffers against the reference's BilateralDenoise /
    ReinhardToneMapping kernels compiled for the host, chained like the reference pipeline
    (reconstruction -> denoise -> tone map)."""
    import ctypes
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    lib = ref.RefOracle().lib
    lib.ref_bilateral_denoise.argtypes = [ctypes.c_int] * 3 + [ctypes.c_float] * 2 + [ctypes.c_void_p] * 2
    lib.ref_reinhard_tonemap.argtypes = [ctypes.c_int] * 2 + [ctypes.c_float] + [ctypes.c_void_p] * 2
    w, h = 100, 70      # not a multiple of the tile size: exercises the clamped halo
    for radius in (1, 4):
        a2 = rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a2, cornell, w, h, 3)
            pl.GI.useDenoise, pl.GI.useTonemapping = True, True
            pl.GI.denoiseKernelRadius = radius
            pl.GI.bilateralDenoiseSigmaRange = 0.5
            for _ in range(2):
                pl.update()
            img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
            den = pl.read_buffer("DenoisedImageCL").reshape(-1, 4).copy()
            tm = pl.read_buffer("TonemappedImageCL").reshape(-1, 4).copy()
            assert pl.fetch("NextFrameImageName") == "TonemappedImageCL"
            pl.scene.destroy()
        finally:
            rr.IntersectionApi.Delete(a2)
        want_den = np.zeros_like(img)
        lib.ref_bilateral_denoise(w, h, radius, 1.0, 0.5, img.ctypes.data, want_den.ctypes.data)
        assert np.allclose(den, want_den, rtol=1e-5, atol=1e-7)
        want_tm = np.zeros_like(img)
        lib.ref_reinhard_tonemap(w, h, 2.0, want_den.ctypes.data, want_tm.ctypes.data)
        ok = np.isfinite(want_tm).all(axis=1)      # black pixels: tL / L = 0 / 0 in both
        assert np.allclose(tm[ok], want_tm[ok], rtol=1e-5, atol=1e-7)
        assert (np.isfinite(tm).all(axis=1) == ok).all()


def test_ingested_obj_scene_renders_like_the_reference(api, tmp_path):
    """OBJ + MTL + PNG through ingest.load_obj, rendered by the CUDA path and by the compiled reference:
    texture addressing (row order, uv orientation), opacity map and the emissive-group light."""
    from PIL import Image
    from mcrt_b200 import ingest
    from mcrt_b200.scenes import CameraSpec
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.RandomState(7)
    img = rng.randint(0, 256, size=(16, 32, 4)).astype(np.uint8)
    img[..., 3] = 255
    Image.fromarray(img, "RGBA").save(tmp_path / "noise.png")
    (tmp_path / "room.mtl").write_text(
        "newmtl wall\nKd 1 1 1\nKs 0.2 0.2 0.2\nNs 30\nmap_Kd noise.png\nnewmtl lamp\nKd 0 0 0\nKe 20 20 20\n")
    (tmp_path / "room.obj").write_text(
        "mtllib room.mtl\n"
        "v -2 0 -2\nv 2 0 -2\nv 2 0 2\nv -2 0 2\nv -2 3 -2\nv 2 3 -2\nv 2 3 2\nv -2 3 2\n"
        "v -0.5 2.9 -0.5\nv 0.5 2.9 -0.5\nv 0.5 2.9 0.5\nv -0.5 2.9 0.5\n"
        "vt 0 0\nvt 3 0\nvt 3 2\nvt 0 2\n"
        "g floor\nusemtl wall\nf 1/1 4/4 3/3 2/2\n"
        "g back\nusemtl wall\nf 1/1 2/2 6/3 5/4\n"
        "g left\nusemtl wall\nf 4/1 1/2 5/3 8/4\n"
        "g right\nusemtl wall\nf 2/1 3/2 7/3 6/4\n"
        "g lamp\nusemtl lamp\nf 9 10 11 12\n")
    scene = ingest.load_obj(str(tmp_path / "room.obj"), camera=CameraSpec(position=(0.0, 1.5, -6.0), euler_deg=(0, 0, 0)))
    assert scene.num_triangles == 10 and len(scene.lights) == 1 and len(scene.textures) == 1
    w, h = 96, 64
    got, _, _ = render_frames(api, scene, w, h, 3, 2)
    want, _ = ref_frames(scene, w, h, 3, 2)
    for f in range(2):
        check_against(got[f], want[f], 0.99)
    assert got[0][:, :3].max() > 0


def test_dynamic_scene_refit_equals_fresh_scene(api):
    """Dynamic updates (RTScene::updateDynamicEntities): moving shapes through SetTransform +
    the shapes-array hot update + a refit Commit renders, bit for bit, what a scene built from scratch with those
    transforms renders; a light and a material hot update likewise; and the moved scene matches the compiled
    reference."""
    def rot_y(deg, t):
        a = np.deg2rad(deg)
        m = np.eye(4, dtype=np.float32)
        m[0, 0], m[0, 2], m[2, 0], m[2, 2] = np.cos(a), np.sin(a), -np.sin(a), np.cos(a)
        m[:3, 3] = t
        return m

    w, h, depth = 96, 96, 4
    scene = scene_library.cornell_box()
    n = len(scene.meshes)
    pl = pt.make_pt_pipeline(api, scene, w, h, depth, refit=True)
    pl.update()
    moves = {n - 1: rot_y(20.0, (0.15, 0.0, 0.1)), n - 2: rot_y(-35.0, (-0.1, 0.2, 0.0))}
    lights = scene.lights.copy()
    lights["intensity"][0, :3] *= 0.5
    mats = scene.materials.copy()
    mats["uber_kd"][0, :3] = (0.2, 0.7, 0.3)
    pl.scene.updateDynamicEntities(moves, lights=lights, materials=mats)
    assert api.GetStats()["last_commit_refit"] == 1
    got = []
    assert pl.frameIndex == 1
    for f in range(2):                   # frames 1 and 2
        pl.update()
        got.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
    pl.scene.destroy()

    fresh = scene_library.cornell_box()
    for i, m in moves.items():
        fresh.meshes[i].transform = m
    fresh = fresh.finalize()
    fresh.lights, fresh.materials = lights, mats
    api2 = rr.IntersectionApi.Create(0)
    try:
        pl2 = pt.make_pt_pipeline(api2, fresh, w, h, depth)
        pl2.update()                     # frame 0, to put the frame counter where the moved pipeline's was
        want = []
        for f in range(2):
            pl2.update()
            want.append(pl2.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
        pl2.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api2)
    for f in range(2):
        assert np.array_equal(got[f].view(np.uint32)[:, :3], want[f].view(np.uint32)[:, :3])
    if ref.available():
        o = ref.RefOracle()
        rp = ref.RefPathTracer(o, fresh, o.build_scene_bvh(fresh), w, h)
        cam = camera.make_camera(fresh.camera, w, h, camera.taa_pixel_offset(1))
        check_against(got[0], rp.render_sample(cam, 1, depth), 0.99)


def test_two_stream_half_frames_equal_single_wavefront_bitwise(api, cornell):
    """ptb_set_overlap: a frame issued as two half-frame wavefronts on two streams gives, bit for bit, the radiance
    and the counters of the same frame issued as one wavefront."""
    w, h, depth = 512, 320, 5          # 163,840 pixels: above the split threshold
    res = []
    for halves in (2, 1):
        a = api if halves == 2 else rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a, cornell, w, h, depth)
            pl.ensure_size()
            pl.set_overlap(halves)
            frames = []
            for _ in range(3):
                pl.update()
                frames.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
            cnt = pl.counters()
            pl.scene.destroy()
        finally:
            if halves == 1:
                rr.IntersectionApi.Delete(a)
        res.append((frames, img, cnt))
    for f in range(3):
        assert np.array_equal(res[0][0][f].view(np.uint32), res[1][0][f].view(np.uint32))
    assert np.array_equal(res[0][1].view(np.uint32), res[1][1].view(np.uint32))
    for k in ("samples", "closest_rays", "shadow_rays", "shaded_vertices"):
        assert res[0][2][k] == res[1][2][k], k
    assert res[0][2]["kernel_launches"] > res[1][2]["kernel_launches"]


@pytest.mark.parametrize("halves", [1, 2])
def test_coherence_bucketed_queues_equal_push_order_bitwise(api, halves):
    """ptb_set_coherence: ordering the live-path queues by the sampler's coming light / lobe decisions changes which
    lanes shade which path, not what any path computes -- radiance, image and counters are bit-identical.  Mode 4 is
    the consumer-side sort with the scene tables staged in shared memory by bulk copies (shade_sorted)."""
    scene = scene_library.sponza_class(64)     # three light types, textured two-lobe materials, alpha masks
    w, h, depth = 512, 320, 5
    res = []
    for mode in (0, 1, 2, 3, 4):
        a = api if mode == 0 else rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a, scene, w, h, depth)
            pl.ensure_size()
            pl.set_overlap(halves)
            pl.set_coherence(mode)
            frames = []
            for _ in range(3):
                pl.update()
                frames.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
            cnt = pl.counters()
            pl.scene.destroy()
        finally:
            if mode != 0:
                rr.IntersectionApi.Delete(a)
        res.append((frames, img, cnt))
    assert res[0][2]["shaded_vertices"] > 2 * w * h      # paths do go several bounces deep
    for mode in (1, 2, 3, 4):
        for f in range(3):
            assert np.array_equal(res[0][0][f].view(np.uint32), res[mode][0][f].view(np.uint32)), (mode, f)
        assert np.array_equal(res[0][1].view(np.uint32), res[mode][1].view(np.uint32))
        for k in ("samples", "closest_rays", "shadow_rays", "shaded_vertices"):
            assert res[0][2][k] == res[mode][2][k], (mode, k)


@pytest.mark.parametrize("coherence", [0, 3])
def test_split_shade_stage_equals_the_single_kernel_bitwise(coherence):
    """ptb_set_shade_split: the shade stage as k_shade_surface (geometry, normal map, emission, material textures,
    a 112-byte record per live vertex) + k_shade_scatter (light sampling, BSDF evaluation, lobe sampling) evaluates
    the expressions of the single kernel on the same values: radiance, accumulated image, counters and the
    per-bounce path record are bit-identical -- on textured two-lobe materials with alpha masks and three light
    types, on the glass / mirror Cornell box (specular chains: emission after a specular bounce ends the path in the
    first kernel), with half-frame wavefronts, frames in flight and a tile partition."""
    cases = [(scene_library.sponza_class(64), 512, 320, 5, {}),
             (scene_library.cornell_box(True), 256, 256, 6, {}),
             (scene_library.sponza_class(64), 512, 320, 4, {"in_flight": 3, "partition": (1, 3)})]
    for scene, w, h, depth, opt in cases:
        res = []
        for split in (False, True):
            a = rr.IntersectionApi.Create(0)
            try:
                pl = pt.make_pt_pipeline(a, scene, w, h, depth)
                pl.ensure_size()
                pl.set_coherence(coherence)
                pl.set_shade_split(split)
                if "partition" in opt:
                    pl.set_partition(*opt["partition"])
                if "in_flight" in opt:
                    pl.set_frames_in_flight(opt["in_flight"])
                else:
                    pl.set_path_trace(depth)
                frames, traces = [], []
                for _ in range(4):
                    pl.update()
                    if "in_flight" not in opt:
                        frames.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
                        traces.append(pl.path_trace(depth).copy())
                img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
                cnt = pl.counters()
                pl.scene.destroy()
            finally:
                rr.IntersectionApi.Delete(a)
            res.append((frames, traces, img, cnt))
        one, two = res
        assert one[3]["shaded_vertices"] > w * h // 4
        for f in range(len(one[0])):
            assert np.array_equal(one[0][f].view(np.uint32), two[0][f].view(np.uint32)), f
            assert np.array_equal(one[1][f], two[1][f]), f
        assert np.array_equal(one[2].view(np.uint32), two[2].view(np.uint32))
        for k in ("samples", "closest_rays", "shadow_rays", "shaded_vertices"):
            assert one[3][k] == two[3][k], k


@pytest.mark.parametrize("depth_in_flight", [2, 4])
def test_frames_in_flight_equal_one_at_a_time_bitwise(api, depth_in_flight):
    """ptb_set_frames_in_flight: consecutive frames rendered concurrently on side streams (own wavefront state and
    camera copy each, accumulation in render order) give bit for bit the images and counters of frames rendered one
    at a time -- with reads in between (which join the frames) and without, with a tile partition, after a hot material
    update, and when the frames-in-flight depth changes in the middle of a run."""
    w, h, depth, frames = 512, 320, 5, 7

    def run(d, partition=None, read_each=False):
        scene = scene_library.sponza_class(64)      # fresh: the hot update below replaces the host scene's materials
        a = rr.IntersectionApi.Create(0)
        try:
            pl = pt.make_pt_pipeline(a, scene, w, h, depth)
            pl.ensure_size()
            if partition:
                pl.set_partition(*partition)
            pl.set_frames_in_flight(d)
            per_frame = []
            for f in range(frames):
                if f == 3:      # hot update in the middle: later frames must see it, earlier ones must not
                    mats = scene.materials.copy()
                    mats["uber_kd"][:, :3] *= 0.5
                    pl.scene.updateDynamicEntities(materials=mats)
                if f == 5 and d > 1:
                    pl.set_frames_in_flight(d - 1)
                pl.update()
                if read_each:
                    per_frame.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
            img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
            wr = pl.read_buffer("WeightedRadianceBuffer").reshape(-1, 4).copy()
            cnt = pl.counters()
            pl.scene.destroy()
        finally:
            rr.IntersectionApi.Delete(a)
        return per_frame, img, wr, cnt

    for partition in (None, (1, 3)):
        base = run(1, partition, read_each=True)
        for read_each in (True, False):
            got = run(depth_in_flight, partition, read_each)
            if read_each:
                for f in range(frames):
                    assert np.array_equal(base[0][f].view(np.uint32), got[0][f].view(np.uint32)), (partition, f)
            assert np.array_equal(base[2].view(np.uint32), got[2].view(np.uint32)), (partition, read_each)
            assert np.array_equal(base[1].view(np.uint32), got[1].view(np.uint32), equal_nan=False) or \
                np.array_equal(np.nan_to_num(base[1]).view(np.uint32), np.nan_to_num(got[1]).view(np.uint32))
            for k in ("samples", "closest_rays", "shadow_rays", "shaded_vertices"):
                assert base[3][k] == got[3][k], (partition, read_each, k)


def test_division_shortcut_is_ieee_division(api, cornell):
    """pt_math.cuh:fdiv (zero numerators answered without the hardware sequence's slow path) returns, bit for bit,
    what IEEE division returns -- for signed zeros, infinities, NaNs, denormals and 4 M random bit patterns."""
    pl = pt.make_pt_pipeline(api, cornell, 64, 64, 2)
    rng = np.random.RandomState(5)
    special = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 1e-39, 3e38, -3e38, 1.17549435e-38,
                        2.0 ** -40, 2.0 ** -100, 2.0 ** 100, 0.1, -7.5], np.float32)
    a = np.concatenate([np.repeat(special, special.size), rng.randint(0, 2 ** 32, 1 << 22, dtype=np.uint64).astype(np.uint32).view(np.float32),
                        np.zeros(1 << 16, np.float32), -np.zeros(1 << 16, np.float32)])
    b = np.concatenate([np.tile(special, special.size), rng.randint(0, 2 ** 32, 1 << 22, dtype=np.uint64).astype(np.uint32).view(np.float32),
                        rng.randint(0, 2 ** 32, 1 << 17, dtype=np.uint64).astype(np.uint32).view(np.float32)])
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    got, want = np.empty_like(a), np.empty_like(a)
    pt._check(pt._lib().ptb_selftest_fdiv(pl.scene._h, a.ctypes.data, b.ctypes.data, got.ctypes.data, want.ctypes.data, a.size))
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got[~nan].view(np.uint32), want[~nan].view(np.uint32))
    with np.errstate(all="ignore"):
        host = a / b                       # the host agrees with the device on everything but NaN payloads
    assert np.array_equal(host[~nan].view(np.uint32), want[~nan].view(np.uint32))
    pl.scene.destroy()


def _mean_radiance(scene, w, h, depth, frames, **estimator):
    a = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_pt_pipeline(a, scene, w, h, depth, use_taa=False)
        pl.ensure_size()
        pl.set_estimator(**estimator)
        acc = np.zeros((w * h, 3), np.float64)
        for _ in range(frames):
            pl.update()
            acc += pl.read_buffer("RadianceBufferCL").reshape(-1, 4)[:, :3]
        cnt = pl.counters()
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(a)
    return acc / frames, cnt


@pytest.mark.parametrize("scene_name", ["cornell", "spheres", "glossy"])
def test_opt_in_russian_roulette_and_mis_keep_the_expectation(scene_name):
    """ptb_set_estimator: Russian roulette and NEE/BSDF MIS (BASELINE north_star; absent from the reference, so
    opt-in) are unbiased re-weightings of the reference's estimator -- the converged images agree.  Bound: block means
    (8 x 8 pixels, 96 frames) within 6 % of max(block mean, image mean), global mean within 1 %."""
    scene = scene_library.cornell_box(scene_name == "spheres")
    if scene_name == "glossy":
        # glossy walls and a second, large, dim emitter (the back wall): light pdf and BSDF pdf are of the same order
        # over most of the image, so both halves of the MIS estimator carry weight and an error in either shows
        from mcrt_b200.scenes import Scene, tri_mesh_light, mesh_area
        c = scene
        idx = [m.name for m in c.meshes].index("backWall")
        lights = np.concatenate([c.lights, tri_mesh_light(idx, (1.5, 1.5, 1.5), mesh_area(c.meshes[idx]))])
        mats = c.materials.copy()
        mats["uber_ks"][:, :3] = 0.5
        mats["uber_kd"][:, :3] *= 0.5
        mats["uber_roughness"][:] = 0.15
        scene = Scene("cornell_two_lights", c.meshes, mats, lights, c.camera).finalize()
    w, h, depth, frames = 96, 96, 5, 96
    base, cnt0 = _mean_radiance(scene, w, h, depth, frames)
    for kw in ({"russian_roulette": True}, {"nee_mis": True}, {"russian_roulette": True, "nee_mis": True}):
        got, cnt = _mean_radiance(scene, w, h, depth, frames, **kw)
        assert np.isfinite(got).all()
        gm, bm = got.mean(), base.mean()
        assert abs(gm - bm) <= 0.01 * bm, (kw, gm, bm)
        blk = lambda im: im.reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3))
        d = (np.abs(blk(got) - blk(base)) / np.maximum(blk(base), bm)).max()
        assert d <= 0.06, (kw, d, bm)
        if kw.get("russian_roulette"):
            assert cnt["closest_rays"] < cnt0["closest_rays"]      # paths do get cut
        else:
            assert cnt["closest_rays"] == cnt0["closest_rays"]     # MIS re-weights, it does not change the paths
    if scene_name == "glossy":      # MIS is active: per-sample values differ from the reference estimator's
        f0, _ = _mean_radiance(scene, w, h, depth, 1)
        f0m, _ = _mean_radiance(scene, w, h, depth, 1, nee_mis=True)
        assert (np.abs(f0 - f0m).max(axis=1) > 1e-3).mean() > 0.2
    # the reference's estimator itself is untouched by the switch being available
    again, _ = _mean_radiance(scene, w, h, depth, 2)
    base2, _ = _mean_radiance(scene, w, h, depth, 2, russian_roulette=False, nee_mis=False)
    assert np.array_equal(again, base2)


def test_pipelined_readback_equals_blocking_read(api, cornell):
    """ptb_read_buffer_async: the snapshot copied behind the next frame is the image of the frame it was issued
    after, even though the next frame's accumulate has overwritten the device buffer in the meantime."""
    import torch
    w = h = 128
    pl = pt.make_pt_pipeline(api, cornell, w, h, 3)
    host = [torch.empty(4 * w * h, dtype=torch.float32, pin_memory=True).numpy() for _ in range(3)]
    want = []
    for f in range(6):
        pl.update()
        pl.read_buffer_async("NextFrameImageCL", host[f % 3])
        want.append(pl.read_buffer("NextFrameImageCL").copy())      # blocking read of the same frame
        if f >= 2:
            # two copies may be in flight: the call above waited for the one issued two calls ago (frame f - 2)
            assert np.array_equal(host[(f - 2) % 3].view(np.uint32), want[f - 2].view(np.uint32))
    pl.read_wait()
    for f in (4, 5):
        assert np.array_equal(host[f % 3].view(np.uint32), want[f].view(np.uint32))
    pl.scene.destroy()
