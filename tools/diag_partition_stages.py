"""Diagnostic: per-stage device time of a frame under a tile partition (one GPU renders part 0 of N)."""
import ctypes, os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200
from mcrt_b200 import camera, pt, rr, scene_library

quads = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
scene = scene_library.traversal_bench_scene(quads)
W, H, depth = 3840, 2160, 5
stream = torch.cuda.Stream()
api = rr.IntersectionApi.Create(0, cuda_stream=stream.cuda_stream)
pl = pt.make_pt_pipeline(api, scene, W, H, depth, acc_type="hlbvh")
pl.ensure_size()
lib, h = pt._lib(), pl.scene._h
fprops = pl.GI.filter_props()
cam = camera.make_camera(scene.camera, W, H, (0.0, 0.0))
pt._check(lib.ptb_set_camera(h, cam.ctypes.data_as(ctypes.c_void_p)))
for parts in (1, 2, 4, 8):
    for halves in (1, 2):
        pl.set_partition(0, parts)
        pl.set_overlap(halves)
        pl.profile_enable(halves == 1)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            for f in range(3):
                pt._check(lib.ptb_pt_render_sample(h, f, depth))
            torch.cuda.synchronize()
            if halves == 1:
                pl.profile_collect()
            ev0.record(stream)
            for f in range(3, 13):
                pt._check(lib.ptb_pt_render_sample(h, f, depth))
            ev1.record(stream)
            torch.cuda.synchronize()
        line = {"parts": parts, "halves": halves, "ms_per_frame": ev0.elapsed_time(ev1) / 10}
        if halves == 1:
            line["stages_ms"] = {k: round(v[0] / 10, 3) for k, v in pl.profile_collect().items()}
        print(json.dumps(line), flush=True)
