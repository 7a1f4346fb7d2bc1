"""GPU suite, round-2 additions to the C ABI: checkpoint / resume of the accumulation state, the merge entry
point of the multi-GPU path, BDPT under the tile partition through the accumulation buffers, the diagnostic
path record, and the two-rank NCCL merge (needs two devices)."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

from mcrt_b200 import rr, pt, scene_library, multi_gpu

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def new_pipeline(scene, w, h, depth, bdpt=False):
    api = rr.IntersectionApi.Create(0)
    make = pt.make_bdpt_pipeline if bdpt else pt.make_pt_pipeline
    return api, make(api, scene, w, h, depth)


def close(api, pl):
    pl.scene.destroy()
    rr.IntersectionApi.Delete(api)


@pytest.mark.parametrize("bdpt", [False, True])
def test_checkpoint_resume_continues_bit_for_bit(cornell, tmp_path, bdpt):
    """Render 8 frames, save (sum(w L), sum(w), frame index [, BDPT sampled-light records]), load into a NEW context,
    render 8 more: the image and the accumulation buffers equal those of 16 frames rendered straight through."""
    w, h, depth = 96, 64, 4
    api, pl = new_pipeline(cornell, w, h, depth, bdpt)
    for _ in range(16):
        pl.update()
    straight = pl.read_buffer("NextFrameImageCL").copy()
    straight_acc = pl.read_buffer("AccumulationBuffer").copy()
    close(api, pl)

    api, pl = new_pipeline(cornell, w, h, depth, bdpt)
    for _ in range(8):
        pl.update()
    path = str(tmp_path / "ckpt.npz")
    pl.save_checkpoint(path)
    close(api, pl)

    api, pl = new_pipeline(cornell, w, h, depth, bdpt)
    pl.load_checkpoint(path)
    assert pl.passes[0].m_frameIndex == 8
    for _ in range(8):
        pl.update()
    resumed = pl.read_buffer("NextFrameImageCL").copy()
    resumed_acc = pl.read_buffer("AccumulationBuffer").copy()
    close(api, pl)
    if bdpt:
        # BDPT sums its strategies with float atomics in an unspecified order: frames are equal up to reassociation
        assert np.allclose(resumed, straight, rtol=2e-5, atol=1e-6)
    else:
        assert np.array_equal(resumed.view(np.uint32), straight.view(np.uint32))
        assert np.array_equal(resumed_acc.view(np.uint32), straight_acc.view(np.uint32))
    with pytest.raises(pt.PathTracerError):
        api, pl = new_pipeline(cornell, w + 32, h, depth, bdpt)
        try:
            pl.load_checkpoint(path)      # other image size
        finally:
            close(api, pl)


def test_bdpt_frame_resumes_from_restored_sampled_light_records(cornell):
    """The BDPT state a checkpoint carries besides the sums (E12: the s = 1 strategies read the previous frame's
    sampled light vertex before rewriting it): frame 4 rendered in a fresh context after ptb_bdpt_reserve +
    ptb_write_buffer("BDPTSampledLightBuffer") equals frame 4 of the uninterrupted run."""
    w, h, depth = 64, 64, 3
    api, pl = new_pipeline(cornell, w, h, depth, True)
    for _ in range(4):
        pl.update()
    sampled = pl.read_buffer("BDPTSampledLightBuffer").copy()
    assert sampled.size == depth * w * h * 8 and np.count_nonzero(sampled) > 0
    pl.update()
    want = pl.read_buffer("RadianceBufferCL").copy()
    close(api, pl)
    api, pl = new_pipeline(cornell, w, h, depth, True)
    pl.ensure_size()
    with pytest.raises(pt.PathTracerError):
        pl.read_buffer("BDPTSampledLightBuffer")          # exists only once the BDPT buffers are reserved
    pt._check(pt._lib().ptb_bdpt_reserve(pl.scene._h, depth))
    pl.write_buffer("BDPTSampledLightBuffer", sampled)
    with pytest.raises(pt.PathTracerError):
        pl.write_buffer("BDPTSampledLightBuffer", sampled[:-4])      # whole buffers only
    pl.passes[0].m_frameIndex = 4
    pl.update()
    got = pl.read_buffer("RadianceBufferCL").copy()
    close(api, pl)
    assert np.allclose(got, want, rtol=2e-5, atol=1e-6)


def test_merge_entry_point_single_rank_and_accumulation_layout(cornell):
    """ptb_merge_accumulation without a communicator is the division alone; "AccumulationBuffer" is the one
    allocation [sum(w L) float4[P] | sum(w) float[P]] that the named views point into."""
    w, h = 64, 48
    P = w * h
    api, pl = new_pipeline(cornell, w, h, 3)
    for _ in range(3):
        pl.update()
    img = pl.read_buffer("NextFrameImageCL").copy()
    acc = pl.read_buffer("AccumulationBuffer")
    assert acc.size == 5 * P
    assert np.array_equal(acc[:4 * P], pl.read_buffer("WeightedRadianceBuffer"))
    assert np.array_equal(acc[4 * P:], pl.read_buffer("FilterWeightsBuffer"))
    p_acc, _ = pl.device_buffer("AccumulationBuffer")
    p_w, _ = pl.device_buffer("FilterWeightsBuffer")
    assert p_w == p_acc + 16 * P
    pl.write_buffer("NextFrameImageCL", np.zeros(4 * P, np.float32))
    for root in (0, -1):
        pl.merge_accumulation(None, root)
        assert np.array_equal(pl.read_buffer("NextFrameImageCL").view(np.uint32), img.view(np.uint32))
    with pytest.raises(pt.PathTracerError):
        pl.merge_accumulation(None, 1)
    close(api, pl)


def test_bdpt_tile_partition_merged_accumulation_equals_single_context():
    """light-tracing splats land in foreign tiles.  Two partitions accumulate 3 BDPT frames each; the
    sum of their accumulation buffers, divided, must be the single-context image (up to float-add order)."""
    scene = scene_library.cornell_box(True)
    w, h, depth, frames = 192, 128, 4, 3
    P = w * h
    api, pl = new_pipeline(scene, w, h, depth, True)
    for _ in range(frames):
        pl.update()
    full = pl.read_buffer("NextFrameImageCL").reshape(-1, 4).copy()
    full_acc = pl.read_buffer("AccumulationBuffer").copy()
    close(api, pl)
    acc = np.zeros(5 * P, np.float64)
    for part in range(2):
        api, pl = new_pipeline(scene, w, h, depth, True)
        pl.ensure_size()
        pl.set_partition(part, 2)
        for _ in range(frames):
            pl.update()
        acc += pl.read_buffer("AccumulationBuffer")
        close(api, pl)
    assert np.allclose(acc[4 * P:], full_acc[4 * P:], rtol=1e-6)            # one weight per sample and pixel
    merged = (acc[:4 * P].reshape(P, 4) / acc[4 * P:, None]).astype(np.float32)
    assert np.allclose(merged[:, :3], full[:, :3], rtol=1e-4, atol=1e-6)
    # and the energy the old path dropped is there: splats into foreign tiles are a visible share of the image
    assert merged[:, :3].mean() > 0


def test_path_trace_record_is_consistent_with_the_counters(cornell):
    w, h, depth = 96, 64, 5
    api, pl = new_pipeline(cornell, w, h, depth)
    pl.ensure_size()
    pl.set_path_trace(depth)
    pl.reset_counters()
    pl.update()
    tr = pl.path_trace(depth)
    cnt = pl.counters()
    rad_on = pl.read_buffer("RadianceBufferCL").copy()
    shaded = (tr[:, :, 2] & 1) != 0
    shadow = (tr[:, :, 2] & 2) != 0
    spawned = (tr[:, :, 2] & 8) != 0
    assert int(shadow.sum()) == cnt["shadow_rays"]
    assert int(spawned.sum()) + w * h == cnt["closest_rays"]
    assert not spawned[depth - 1].any() and shaded[0].sum() > 0.5 * w * h
    assert (tr[:, :, 3][~shadow] == 0).all()
    pl.set_path_trace(0)
    pl.passes[0].m_frameIndex = 0
    pl.update()
    assert np.array_equal(pl.read_buffer("RadianceBufferCL").view(np.uint32), rad_on.view(np.uint32))
    close(api, pl)


def _device_count():
    import torch
    return torch.cuda.device_count()


def test_two_rank_nccl_merge_through_the_c_abi(tmp_path):
    """Two processes, two GPUs, sample-index partition: after ptb_merge_accumulation(comm, root = 0) every step,
    rank 0's image equals the single-GPU image of the same 8 frames up to float-add reassociation, and the
    non-root buffers are zero."""
    if _device_count() < 2:
        pytest.skip("needs two CUDA devices")
    out = str(tmp_path / "merge")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(ROOT, "tests", "nccl_merge_worker.py"), out]
    subprocess.check_call(cmd, cwd=ROOT, timeout=600)
    merged = np.load(out + "_rank0.npy")
    other = np.load(out + "_rank1_acc.npy")
    assert not other.any()
    scene = scene_library.cornell_box()
    api, pl = new_pipeline(scene, 128, 96, 4)
    for _ in range(8):
        pl.update()
    single = pl.read_buffer("NextFrameImageCL").copy()
    close(api, pl)
    assert np.allclose(merged, single, rtol=1e-5, atol=1e-7)


def test_device_libm_is_bit_identical_to_the_host_c_library(cornell):
    """csrc/pt_libm.h evaluated on the device against this box's libm.so.6 (sinf / cosf / logf called through
    ctypes): every sampled direction of the shading path goes through these.  (The exhaustive comparison of the
    same source on the host is tests/cpp/libm_check.cpp, CPU suite.)"""
    import ctypes
    api, pl = new_pipeline(cornell, 32, 32, 1)
    rng = np.random.RandomState(11)
    xs = np.concatenate([
        (rng.rand(120000) * 7.0).astype(np.float32),                     # [0, 2 pi + ...]: what the samplers pass
        ((rng.rand(30000) - 0.5) * 400.0).astype(np.float32),            # both argument reductions
        (rng.rand(10000).astype(np.float32) * np.float32(1e-3)),
        rng.randint(0, 1 << 32, 40000, dtype=np.uint64).astype(np.uint32).view(np.float32),    # any bit pattern
        np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1.0, 1e-45, 3.4e38, np.pi / 4, 120.0, 119.99999], np.float32)])
    out = np.empty_like(xs)
    libm = ctypes.CDLL("libm.so.6")
    for fn, name in ((0, "sinf"), (1, "cosf"), (2, "logf")):
        f = getattr(libm, name)
        f.argtypes, f.restype = [ctypes.c_float], ctypes.c_float
        pt._check(pt._lib().ptb_selftest_libm(pl.scene._h, fn, xs.ctypes.data_as(ctypes.c_void_p),
                                              out.ctypes.data_as(ctypes.c_void_p), xs.size))
        want = np.array([f(float(x)) for x in xs], np.float32)
        ok = (out.view(np.uint32) == want.view(np.uint32)) | (np.isnan(out) & np.isnan(want))
        assert ok.all(), (name, xs[~ok][:5], out[~ok][:5], want[~ok][:5])
    close(api, pl)
