"""Diagnostic: how much of the per-frame latency chain under a tile partition can be hidden by keeping several
independent frames in flight.  Emulated with D separate contexts (own stream, own state) on one GPU, each rendering
part 0 of `parts` of different samples."""
import ctypes, os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200
from mcrt_b200 import camera, pt, rr, scene_library

quads, parts = int(sys.argv[1]), int(sys.argv[2])
scene = scene_library.traversal_bench_scene(quads)
W, H, depth = 3840, 2160, 5
cam = camera.make_camera(scene.camera, W, H, (0.0, 0.0))
lib = None
for D in (1, 2, 3, 4):
    ctxs = []
    for d in range(D):
        st = torch.cuda.Stream()
        api = rr.IntersectionApi.Create(0, cuda_stream=st.cuda_stream)
        pl = pt.make_pt_pipeline(api, scene, W, H, depth, acc_type="hlbvh")
        pl.ensure_size()
        pl.set_partition(0, parts)
        lib = pt._lib()
        pt._check(lib.ptb_set_camera(pl.scene._h, cam.ctypes.data_as(ctypes.c_void_p)))
        ctxs.append((st, api, pl))
    def run(n, first):
        for f in range(first, first + n):
            st, api, pl = ctxs[f % D]
            pt._check(lib.ptb_pt_render_sample(pl.scene._h, f, depth))
    run(2 * D, 0)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    torch.cuda.synchronize()
    import time
    t0 = time.perf_counter()
    n = 24
    run(n, 2 * D)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"parts": parts, "frames_in_flight": D, "ms_per_frame": ms / n}), flush=True)
    for st, api, pl in ctxs:
        pl.scene.destroy()
        rr.IntersectionApi.Delete(api)
