"""Diagnostic: per-depth agreement of the CUDA path tracer with the compiled reference."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import mcrt_b200
from mcrt_b200 import rr, pt, camera, scene_library
from oracle import ref
import test_pt_gpu as T

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "knot"
    scene = scene_library.dragon_class(60, 24) if name == "knot" else scene_library.cornell_box(name == "spheres")
    w, h = 160, 96
    for depth in (1, 2, 3, 4):
        api = rr.IntersectionApi.Create(0)
        got, _, cnt = T.render_frames(api, scene, w, h, depth, 1)
        rr.IntersectionApi.Delete(api)
        want, rp = T.ref_frames(scene, w, h, depth, 1)
        g, wv = got[0][:, :3], want[0][:, :3]
        e = T.rel_err(g, wv)
        nz = (wv.max(1) > 0) | (g.max(1) > 0)
        print("depth", depth, "nonzero px", int(nz.sum()), "frac ok (all)", float((e <= 1e-4).mean()),
              "frac ok (nonzero)", float((e[nz] <= 1e-4).mean()), "mean got", g.mean(0), "mean want", wv.mean(0),
              "nan got", int(np.isnan(g).sum()), "nan want", int(np.isnan(wv).sum()),
              "rays", cnt["closest_rays"], cnt["shadow_rays"], "ref rays", rp.rays_traced, rp.shadow_rays_traced)
        bad = np.nonzero(e > 1e-4)[0][:5]
        for i in bad:
            print("   px", i, "got", g[i], "want", wv[i])
main()
