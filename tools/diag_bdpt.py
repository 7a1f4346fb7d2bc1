"""Per-depth / per-frame agreement of the CUDA BDPT with the compiled reference (run on a GPU box)."""
import sys
import numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import mcrt_b200  # noqa
from mcrt_b200 import scene_library
import test_bdpt_gpu as T

which = sys.argv[1] if len(sys.argv) > 1 else "cornell"
scene = {"cornell": lambda: scene_library.cornell_box(), "spheres": lambda: scene_library.cornell_box(True),
         "knot": lambda: scene_library.dragon_class(40, 16)}[which]()
w, h = 96, 96
for depth in (1, 2, 3, 5):
    got, cnt = T.gpu_frames(scene, w, h, depth, 3)
    want, rb = T.ref_frames(scene, w, h, depth, 3)
    for f in range(3):
        e = T.rel_err(got[f][:, :3], want[f][:, :3])
        print("depth", depth, "frame", f, "frac ok", float((e <= 1e-4).mean()), "mean got", got[f][:, :3].mean(0),
              "want", want[f][:, :3].mean(0), "finite", bool(np.isfinite(got[f]).all()))
    print("  counters", cnt, "ref rays", rb.rays_traced, rb.connection_rays_traced)
