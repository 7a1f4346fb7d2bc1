"""One JSON line per BASELINE.json config on ONE GPU (device-resident, CUDA events, 1 sample per pixel per step):

  C1  Cornell box, regular PT 512x512 depth 5
  C2  dragon_class (871,200-triangle glossy dielectric knot), regular PT 1920x1080 depth 8
  C3  sponza_class: that is bench.py itself; repeated here only with --all
  C4  BDPT, Cornell box with glass + mirror spheres, 1024x1024 depth 5
  C5  10 M-triangle knot, regular PT 3840x2160 depth 5 (one GPU's view of the workload; `--tiles R N` renders
      only the 64x64 tiles of rank R of N, which is what each of N GPUs does under the tile partition)

The numbers are supporting evidence for DESIGN.md section 6; the headline metric is bench.py's.
usage: python tools/bench_configs.py [C1 C2 C4 C5] [--steps K] [--tiles R N]
"""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200  # noqa: E402,F401
from mcrt_b200 import camera, pt, rr, scene_library  # noqa: E402


def run(name, scene, w, h, depth, steps, bdpt=False, tiles=None):
    stream = torch.cuda.Stream()
    api = rr.IntersectionApi.Create(0, cuda_stream=stream.cuda_stream)
    acc_type = os.environ.get("MCRT_ACC_TYPE", "fatbvh")      # hlbvh = the device builder (lbvh_build.cu)
    t0 = time.perf_counter()
    make = pt.make_bdpt_pipeline if bdpt else pt.make_pt_pipeline
    pl = make(api, scene, w, h, depth, acc_type=acc_type)
    setup = time.perf_counter() - t0
    pl.ensure_size()
    fif = int(os.environ.get("MCRT_FRAMES_IN_FLIGHT", "1"))
    if fif > 1 and not bdpt:
        pl.set_frames_in_flight(fif)
    if tiles:
        pl.set_partition(tiles[0], tiles[1])
    lib, hnd = pt._lib(), pl.scene._h
    render = lib.ptb_bdpt_render_sample if bdpt else lib.ptb_pt_render_sample
    fprops = pl.GI.filter_props()
    cams = []
    for f in range(steps + 3):
        off = camera.taa_pixel_offset(f, pl.GI.filterRadius, True)
        cams.append((camera.make_camera(scene.camera, w, h, off), off))

    def step(f):
        cam, off = cams[f]
        fprops.pixel_offset[0], fprops.pixel_offset[1] = off
        pt._check(lib.ptb_set_camera(hnd, cam.ctypes.data_as(ctypes.c_void_p)))
        pt._check(render(hnd, f, depth))
        pt._check(lib.ptb_accumulate(hnd, f, ctypes.byref(fprops)))

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for f in range(3):
            step(f)
        torch.cuda.synchronize()
        pl.reset_counters()
        ev0.record(stream)
        for f in range(3, 3 + steps):
            step(f)
        ev1.record(stream)
        torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    cnt = pl.counters()
    st = api.GetStats()
    img = pl.read_buffer("NextFrameImageCL").reshape(-1, 4)
    line = {"config": name, "integrator": "bdpt" if bdpt else "pt", "triangles": int(st["num_triangles"]),
            "width": w, "height": h, "max_depth": depth, "steps": steps, "tiles": tiles, "frames_in_flight": fif,
            "ms_per_step": ms / steps, "paths_per_s": cnt["samples"] / (ms * 1e-3),
            "mrays_per_s": (cnt["closest_rays"] + cnt["shadow_rays"]) / (ms * 1e-3) / 1e6,
            "rays_per_path": (cnt["closest_rays"] + cnt["shadow_rays"]) / max(cnt["samples"], 1),
            "bvh_nodes": int(st["num_nodes"]), "bvh_build_s": st["build_seconds"], "setup_s": setup,
            "image_mean": float(img[:, :3].mean()), "finite": bool(np.isfinite(img).all())}
    print(json.dumps(line), flush=True)
    pl.scene.destroy()
    rr.IntersectionApi.Delete(api)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["C1", "C2", "C4", "C5"])
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--tiles", type=int, nargs=2, default=None)
    a = ap.parse_args()
    for c in a.configs:
        if c == "C1":
            run("C1 cornell 512x512 d5", scene_library.cornell_box(), 512, 512, 5, a.steps * 5)
        elif c == "C2":
            run("C2 dragon_class 871200 tris 1920x1080 d8", scene_library.dragon_class(), 1920, 1080, 8, a.steps)
        elif c == "C3":
            run("C3 sponza_class 262144 tris 1920x1080 d5", scene_library.sponza_class(1024), 1920, 1080, 5, a.steps)
        elif c == "C4":
            run("C4 bdpt cornell + glass/mirror spheres 1024x1024 d5", scene_library.cornell_box(with_spheres=True),
                1024, 1024, 5, a.steps, bdpt=True)
        elif c == "C5":
            run("C5 knot 10M tris 3840x2160 d5", scene_library.traversal_bench_scene(2236), 3840, 2160, 5,
                max(a.steps // 2, 5), tiles=a.tiles)


if __name__ == "__main__":
    main()
