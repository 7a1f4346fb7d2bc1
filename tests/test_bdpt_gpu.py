"""GPU suite: the CUDA bidirectional path tracer against the reference's own BDPT.cl kernels compiled
for the host (oracle/_ref) and the committed golden vectors generated from them.

BDPT sums its strategies into pixels with float atomics in an unspecified order (in the reference too,
BDPT.cl:654-669,888-905), and a light-tracing splat whose projection moves across a pixel border lands
in the neighbouring pixel, so the bar is: per-sample radiance within 1e-4 relative on the stated
fraction of pixels, image mean within 2 %.  Frames must be rendered in order from 0: the s = 1
strategies read the previous frame's sampled light vertex."""
import os

import numpy as np
import pytest

from mcrt_b200 import rr, pt, camera, scene_library
from oracle import ref

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-4


def rel_err(a, b):
    return (np.abs(a - b) / np.maximum(np.abs(b), 1e-3)).max(axis=1)


def gpu_frames(scene, w, h, depth, frames, partition=None):
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_bdpt_pipeline(api, scene, w, h, depth)
        if partition is not None:
            pl.ensure_size()
            pl.set_partition(*partition)
        out = []
        for _ in range(frames):
            pl.update()
            out.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
        cnt = pl.counters()
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    return out, cnt


def ref_frames(scene, w, h, depth, frames):
    o = ref.RefOracle()
    nodes = o.build_scene_bvh(scene)
    rb = ref.RefBDPT(o, scene, nodes, w, h, depth)
    out = []
    for f in range(frames):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        out.append(rb.render_sample(cam, f).copy())
    return out, rb


def check(got, want, min_frac):
    e = rel_err(got[:, :3], want[:, :3])
    frac = float((e <= REL_TOL).mean())
    assert frac >= min_frac, "only %.5f of pixels within %g" % (frac, REL_TOL)
    assert np.allclose(got[:, :3].mean(0), want[:, :3].mean(0), rtol=2e-2), (got[:, :3].mean(0), want[:, :3].mean(0))
    return frac


@pytest.mark.parametrize("depth", [1, 3, 5])
def test_cornell_bdpt_vs_reference(cornell, depth):
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    w = h = 96
    got, cnt = gpu_frames(cornell, w, h, depth, 3)
    want, rb = ref_frames(cornell, w, h, depth, 3)
    for f in range(3):
        check(got[f], want[f], 0.98)
    assert cnt["samples"] == 3 * w * h


def test_glass_and_mirror_spheres_bdpt_vs_reference():
    """C4: Cornell box with a specular-transmission sphere and a mirror sphere -- delta vertices,
    non-connectible vertices and the delta-flag handling in the MIS sums."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = scene_library.cornell_box(True)
    w = h = 128
    got, _ = gpu_frames(scene, w, h, 5, 3)
    want, _ = ref_frames(scene, w, h, 5, 3)
    for f in range(3):
        check(got[f], want[f], 0.97)


def test_all_light_types_bdpt_vs_reference():
    """Directional (infinite, delta) + disk light on the glossy knot scene: the infinite-light subpath
    correction and both area-light emission samplers."""
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    scene = scene_library.dragon_class(40, 16)
    w, h = 128, 72
    got, _ = gpu_frames(scene, w, h, 4, 2)
    want, _ = ref_frames(scene, w, h, 4, 2)
    for f in range(2):
        check(got[f], want[f], 0.95)


def test_bdpt_agrees_with_path_tracer_in_the_mean(cornell):
    """Both integrators estimate the same image: 16-frame means agree within Monte-Carlo noise."""
    w = h = 64
    got, _ = gpu_frames(cornell, w, h, 5, 16)
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_pt_pipeline(api, cornell, w, h, 6)
        acc = np.zeros((w * h, 4), np.float32)
        for _ in range(16):
            pl.update()
            acc += pl.read_buffer("RadianceBufferCL").reshape(-1, 4)
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    m_b = np.mean([g[:, :3].mean(0) for g in got], axis=0)
    m_p = acc[:, :3].mean(0) / 16
    assert np.allclose(m_b, m_p, rtol=0.1), (m_b, m_p)


def test_golden_cornell_spheres_bdpt():
    """Committed vectors of the compiled reference (tools/make_golden.py), frames 0 and 1."""
    g = np.load(os.path.join(GOLDEN, "cornell_spheres_bdpt_64x64_d5.npz"))
    got, _ = gpu_frames(scene_library.cornell_box(True), 64, 64, 5, 2)
    for f in range(2):
        check(got[f], g["radiance_f%d" % f], 0.97)


def test_bdpt_tile_partition_sums_to_full_frame(cornell):
    """Tile partition: each rank traces its tiles' subpaths but splats anywhere; the sum of
    the ranks' full-frame buffers equals the single-context image up to float-add reassociation."""
    w, h = 128, 128
    full, _ = gpu_frames(cornell, w, h, 4, 1)
    parts = [gpu_frames(cornell, w, h, 4, 1, partition=(p, 2))[0][0] for p in range(2)]
    merged = parts[0] + parts[1]
    assert np.allclose(merged[:, :3], full[0][:, :3], rtol=1e-5, atol=1e-6)


def test_bdpt_sobol_sampler_variant_vs_reference(cornell):
    from mcrt_b200 import sobol
    if not ref.available(sobol=True):
        pytest.skip("oracle/_ref Sobol variant not built")
    table = sobol.matrices()
    w = h = 64
    api = rr.IntersectionApi.Create(0)
    try:
        pl = pt.make_bdpt_pipeline(api, cornell, w, h, 3)
        pl.set_sampler(sobol.RT_SAMPLER_SOBOL, table)
        got = []
        for _ in range(2):
            pl.update()
            got.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
        pl.scene.destroy()
    finally:
        rr.IntersectionApi.Delete(api)
    o = ref.RefOracle(sobol_table=table)
    rb = ref.RefBDPT(o, cornell, o.build_scene_bvh(cornell), w, h, 3)
    for f in range(2):
        cam = camera.make_camera(cornell.camera, w, h, camera.taa_pixel_offset(f))
        check(got[f], rb.render_sample(cam, f), 0.97)
