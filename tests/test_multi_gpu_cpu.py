"""CPU suite (gloo, world_size 2): the N > 1 host logic of monte-carlo-raytracer_b200/multi_gpu.py --
sample-index and tile partitions plus the single all-reduce of the accumulation buffers -- with the
compiled reference (oracle/_ref) standing in for the renderer on each rank."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mcrt_b200 import camera, scene_library, multi_gpu
from oracle import ref

W, H, DEPTH, FRAMES = 64, 64, 3, 4


def filter_weight(frame):
    """Triangle filter of the frame's TAA offset, radius (2,2) (filters.cl:17-20)."""
    ox, oy = camera.taa_pixel_offset(frame)
    return np.float32(max(0.0, np.float32(2.0) - abs(np.float32(ox)))) * np.float32(max(0.0, np.float32(2.0) - abs(np.float32(oy))))


def render(rp, scene, frame):
    cam = camera.make_camera(scene.camera, W, H, camera.taa_pixel_offset(frame))
    return np.clip(rp.render_sample(cam, frame, DEPTH), 0.0, 1000.0).astype(np.float32)


def accumulate(frames, mask=None):
    """reconstruction.cl:6-60 on the host: sum(w * clamp(L)), sum(w) over this rank's frames / pixels."""
    scene = scene_library.cornell_box()
    o = ref.RefOracle()
    rp = ref.RefPathTracer(o, scene, o.build_scene_bvh(scene), W, H)
    wr = np.zeros((W * H, 4), np.float32)
    fw = np.zeros(W * H, np.float32)
    for f in frames:
        r, w = render(rp, scene, f), filter_weight(f)
        if mask is not None:
            r = r * mask[:, None]
            wr += r * w
            fw += w * mask
        else:
            wr += r * w
            fw += w
    return wr, fw


def _worker(rank, world, port, mode, queue):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        if mode == "sample":
            wr, fw = accumulate([f for f in multi_gpu.frames_of_rank(rank, world, FRAMES // world)])
        else:
            mask = np.zeros(W * H, np.float32)
            mask[multi_gpu.partition_pixels(W, H, rank, world)] = 1.0
            wr, fw = accumulate(range(FRAMES), mask)
        if mode == "sample":      # the pipelined form bench.py uses: begin / finish around the collective
            m = multi_gpu.PipelinedMerger(W * H, torch.device("cpu"))
            img = torch.zeros(W * H, 4)
            m.begin(torch.from_numpy(wr), torch.from_numpy(fw))
            m.finish(img)
            assert m.finish(img) is img      # idempotent when nothing is pending
        else:
            img = multi_gpu.merge_accumulation(torch.from_numpy(wr), torch.from_numpy(fw))
        if rank == 0:
            queue.put(img.numpy().copy())
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    img = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return img


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_sample_index_partition_merges_to_the_single_process_image():
    img = _run("sample")
    wr, fw = accumulate(range(FRAMES))
    want = wr / fw[:, None]
    # same samples, sums re-associated across ranks: rounding only
    assert np.allclose(img, want, rtol=1e-5, atol=1e-6)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_tile_partition_merges_bit_identically():
    img = _run("tile")
    wr, fw = accumulate(range(FRAMES))
    want = wr / fw[:, None]
    assert np.array_equal(img.view(np.uint32), want.view(np.uint32))


def test_tile_partitions_are_disjoint_and_cover_the_frame():
    for (w, h, parts) in [(1920, 1080, 8), (200, 130, 3), (64, 64, 2), (63, 65, 4)]:
        seen = np.zeros(w * h, np.int32)
        for p in range(parts):
            px = multi_gpu.partition_pixels(w, h, p, parts)
            assert px.size == np.unique(px).size
            seen[px] += 1
        assert (seen == 1).all()
    # load balance at the bench resolution: tile counts differ by at most one tile
    sizes = [multi_gpu.partition_pixels(1920, 1080, p, 8).size for p in range(8)]
    assert max(sizes) - min(sizes) <= 64 * 64


def test_frames_of_rank_interleave():
    got = sorted(sum((multi_gpu.frames_of_rank(r, 4, 3) for r in range(4)), []))
    assert got == list(range(12))
