"""One frame of the C5-class scene under tile partition 0 of N (argv[2]), for ncu."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200
from mcrt_b200 import camera, pt, rr, scene_library
scene = scene_library.traversal_bench_scene(int(sys.argv[1]))
W, H, depth = 3840, 2160, 5
api = rr.IntersectionApi.Create(0)
pl = pt.make_pt_pipeline(api, scene, W, H, depth, acc_type="hlbvh")
pl.ensure_size()
lib, h = pt._lib(), pl.scene._h
cam = camera.make_camera(scene.camera, W, H, (0.0, 0.0))
pt._check(lib.ptb_set_camera(h, cam.ctypes.data_as(ctypes.c_void_p)))
pl.set_partition(0, int(sys.argv[2]))
pl.set_overlap(1)
for f in range(2):
    pt._check(lib.ptb_pt_render_sample(h, f, depth))
torch.cuda.synchronize()
