#!/usr/bin/env python
"""BASELINE.json configs[4]: the 10 M-triangle scene, regular PT 3840x2160, depth 5, image-TILE partitioned across
N GPUs (ptb_set_partition: 64x64 tiles dealt round-robin), one NCCL all-reduce of the accumulation buffer behind the C ABI
(ptb_merge_accumulation).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/bench_c5_tiles.py --steps K [--reduce-every 1]

Strong scaling: the frame is fixed, every rank renders its tiles of every sample.  Timed with CUDA events on the
launching stream between barriers, max over ranks; the all-reduce(s) and the final division are inside the timed region
(--reduce-every 0 = one reduce at the end of the run, the batch-render case; 1 = every step, progressive display).
Rank 0 prints one JSON line.  Also checks that the merged image equals the single-GPU image of the same samples
bit for bit when --verify is given (rank 0 re-renders the whole frame)."""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200  # noqa: E402,F401
from mcrt_b200 import camera, multi_gpu, pt, rr, scene_library  # noqa: E402


class DevView:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--quads", type=int, default=2236, help="2236 -> 9,999,394 triangles")
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--depth", type=int, default=5)
    ap.add_argument("--reduce-every", type=int, default=0)
    ap.add_argument("--acc-type", default="hlbvh", help="hlbvh = device builder (fast set-up); fatbvh = host SAH")
    ap.add_argument("--verify", action="store_true")
    ap.add_argument("--frames-in-flight", type=int, default=3, help="ptb_set_frames_in_flight (1 = off)")
    ap.add_argument("--partition", default="tiles", choices=["tiles", "samples"],
                    help="tiles: every rank renders its 64x64 tiles of every sample (strong scaling, BASELINE configs[4]); "
                         "samples: rank r renders whole frames r, r+N, ... (weak scaling, like bench.py)")
    ap.add_argument("--parts", type=int, default=0, help="single GPU only: render part 0 of this many tile partitions (what one of N GPUs does)")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    W, H, depth, P = a.width, a.height, a.depth, a.width * a.height
    scene = scene_library.traversal_bench_scene(a.quads)
    stream = torch.cuda.Stream(device=dev)
    api = rr.IntersectionApi.Create(local, cuda_stream=stream.cuda_stream)
    t0 = time.perf_counter()
    pl = pt.make_pt_pipeline(api, scene, W, H, depth, acc_type=a.acc_type)
    setup_s = time.perf_counter() - t0
    pl.ensure_size()
    by_tiles = a.partition == "tiles"
    if by_tiles:
        pl.set_partition(rank, a.parts if (a.parts > 0 and world == 1) else world)
    pl.set_frames_in_flight(a.frames_in_flight)
    lib, h = pt._lib(), pl.scene._h
    fprops = pl.GI.filter_props()
    views = {k: pl.device_buffer(k) for k in ("WeightedRadianceBuffer", "FilterWeightsBuffer", "NextFrameImageCL")}
    t_wr, t_fw, t_img = (torch.as_tensor(DevView(*views[k]), device=dev) for k in views)
    stage = torch.empty(5 * P, dtype=torch.float32, device=dev)
    K, Wm = a.steps, a.warmup
    cams = []
    frame_of = (lambda s: s) if by_tiles else (lambda s: s * world + rank)
    for s_ in range(Wm + K):
        off = camera.taa_pixel_offset(frame_of(s_), pl.GI.filterRadius, True)
        cams.append((camera.make_camera(scene.camera, W, H, off), off))

    def step(s_):
        f = frame_of(s_)
        cam, off = cams[s_]
        fprops.pixel_offset[0], fprops.pixel_offset[1] = off
        pt._check(lib.ptb_set_camera(h, cam.ctypes.data_as(ctypes.c_void_p)))
        pt._check(lib.ptb_pt_render_sample(h, f, depth))
        pt._check(lib.ptb_accumulate(h, f, ctypes.byref(fprops)))

    # the one collective of the path, behind the C ABI: NCCL all-reduce of the 20 B/pixel accumulation buffer + the division
    comm = multi_gpu.create_comm(pl, rank, world) if world > 1 else None

    def merge():
        pt._check(lib.ptb_merge_accumulation(h, comm, -1))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for f in range(Wm):
            step(f)
        merge()
        barrier()
        pt._check(lib.ptb_reset_accumulation(h))
        pl.reset_counters()
        barrier()
        ev0.record(stream)
        for f in range(Wm, Wm + K):
            step(f)
            if a.reduce_every and (f - Wm + 1) % a.reduce_every == 0:
                merge()
        if not a.reduce_every or K % a.reduce_every:
            merge()
        ev1.record(stream)
        barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    cnt = pl.counters()
    tot = torch.tensor([cnt["samples"], cnt["closest_rays"] + cnt["shadow_rays"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms = float(ms.item())
    img = t_img.view(P, 4).cpu().numpy()
    verified = None
    if a.verify and rank == 0 and by_tiles:
        # the whole frame on one GPU, same samples: tile partition A is bit-identical
        with torch.cuda.stream(stream):
            pl.set_partition(0, 1)
            pt._check(lib.ptb_reset_accumulation(h))
            for f in range(Wm, Wm + K):
                step(f)
            merge_local = t_wr.view(P, 4) / t_fw.view(P, 1)
            torch.cuda.synchronize(dev)
        verified = bool(np.array_equal(merge_local.cpu().numpy().view(np.uint32), img.view(np.uint32)))
    if world > 1:
        dist.barrier()
    if rank == 0:
        st = api.GetStats()
        print(json.dumps({
            "config": "C5 knot %d tris %dx%d d%d, image tiles over %d GPU(s)" % (st["num_triangles"], W, H, depth, world),
            "n_gpus": world, "partition": a.partition, "scaling": "strong" if by_tiles else "weak", "steps": K, "warmup": Wm, "ms_per_step": ms / K,
            "paths_per_s": float(tot[0].item()) / (ms * 1e-3), "mrays_per_s": float(tot[1].item()) / (ms * 1e-3) / 1e6,
            "reduce": "every %d step(s)" % a.reduce_every if a.reduce_every else "once per run",
            "frames_in_flight": a.frames_in_flight, "allreduce_bytes": 20 * P, "acc_type": a.acc_type, "bvh_on_device": st["last_build_on_device"],
            "setup_s": setup_s, "commit_s": st["build_seconds"], "image_mean": float(img[:, :3].mean()),
            "finite": bool(np.isfinite(img).all()), "bit_identical_to_one_gpu": verified}), flush=True)
    if comm is not None:
        multi_gpu.destroy_comm(comm)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
