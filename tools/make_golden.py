"""Generates the committed golden vectors under tests/golden/ by running the COMPILED REFERENCE
(oracle/_ref/libmcrt_ref.so: the reference's own .cl kernels + Radeon Rays' Bvh2) in this container.
Re-run after any change to the scene builders; tests/test_golden_cpu.py checks that the committed
files still match what this script produces."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mcrt_b200  # noqa
from mcrt_b200 import camera, scene_library
from oracle import ref
from raygen import camera_rays, random_rays

OUT = os.path.join(ROOT, "tests", "golden")


def cornell_pt(w=64, h=64, depth=5, frames=2):
    scene = scene_library.cornell_box()
    o = ref.RefOracle()
    nodes = o.build_scene_bvh(scene)
    rp = ref.RefPathTracer(o, scene, nodes, w, h)
    out = {}
    for f in range(frames):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        tr = {}
        out["radiance_f%d" % f] = rp.render_sample(cam, f, depth, tr).copy()
        if f == 0:
            out["primary_shapeid"] = tr["primary_hits"]["shapeid"].copy()
            out["primary_primid"] = tr["primary_hits"]["primid"].copy()
            out["primary_uvwt"] = tr["primary_hits"]["uvwt"].copy()
    return out


def cornell_spheres_bdpt(w=64, h=64, depth=5, frames=2):
    """C4 (Cornell + glass + mirror sphere), BDPT.  The reference sums strategies with float atomics in
    an unspecified order, so these vectors are reproducible to rounding only (compared with rtol 1e-5)."""
    scene = scene_library.cornell_box(True)
    o = ref.RefOracle()
    nodes = o.build_scene_bvh(scene)
    rb = ref.RefBDPT(o, scene, nodes, w, h, depth)
    out = {}
    for f in range(frames):
        cam = camera.make_camera(scene.camera, w, h, camera.taa_pixel_offset(f))
        out["radiance_f%d" % f] = rb.render_sample(cam, f).copy()
    return out


def cornell_isect(n=4096):
    scene = scene_library.cornell_box()
    o = ref.RefOracle()
    nodes = o.build_scene_bvh(scene)
    lo, hi = scene.world_bounds()
    rays = random_rays(n, lo - 1.0, hi + 1.0, seed=1234)
    hits = o.intersect(nodes, rays)
    occ = o.occluded(nodes, rays)
    return {"rays": rays.view(np.uint8), "shapeid": hits["shapeid"], "primid": hits["primid"], "uvwt": hits["uvwt"],
            "occluded": occ}


def main():
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "cornell_pt_64x64_d5.npz"), **cornell_pt())
    np.savez_compressed(os.path.join(OUT, "cornell_isect_4096.npz"), **cornell_isect())
    np.savez_compressed(os.path.join(OUT, "cornell_spheres_bdpt_64x64_d5.npz"), **cornell_spheres_bdpt())
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
