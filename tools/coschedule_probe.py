"""Experiment: aggregate paths/s of TWO path-tracer contexts on two CUDA streams of ONE GPU (frames of the
two contexts interleaved from one host thread) against one context alone.  If kernels with different
bottlenecks (issue-bound traversal, instruction-fetch-bound shading) overlap, the aggregate goes up."""
import os
import sys
import time
import ctypes

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mcrt_b200  # noqa
from mcrt_b200 import rr, pt, camera, scene_library

W, H, D, K = 1920, 1080, 5, 30
scene = scene_library.sponza_class(int(os.environ.get("TEX", "1024")))
n_ctx = int(sys.argv[1]) if len(sys.argv) > 1 else 2
streams = [torch.cuda.Stream() for _ in range(n_ctx)]
pls = []
for s in streams:
    api = rr.IntersectionApi.Create(0, cuda_stream=s.cuda_stream)
    pl = pt.make_pt_pipeline(api, scene, W, H, D)
    pl.ensure_size()
    pls.append(pl)
lib = pt._lib()
cams = [camera.make_camera(scene.camera, W, H, camera.taa_pixel_offset(f)) for f in range(K + 3)]


def run(frames):
    for f in frames:
        for pl in pls:
            pt._check(lib.ptb_set_camera(pl.scene._h, cams[f].ctypes.data_as(ctypes.c_void_p)))
            pt._check(lib.ptb_pt_render_sample(pl.scene._h, f, D))


run(range(3))
torch.cuda.synchronize()
t0 = time.perf_counter()
run(range(3, 3 + K))
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print("contexts %d  trace/SM %s shade/SM %s : %.3f ms per frame-pair, %.1f Mpaths/s aggregate" % (
    n_ctx, os.environ.get("MCRT_TRACE_CTAS_PER_SM", "8"), os.environ.get("MCRT_SHADE_CTAS_PER_SM", "default"),
    dt / K * 1e3, n_ctx * K * W * H / dt / 1e6))
