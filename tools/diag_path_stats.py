"""Diagnostic: where C3's paths end.  Runs the stage-wise (reference-layout) pipeline and reads the ray /
intersection buffers back after every stage: live rays, hits, emitter hits, shadow rays per bounce."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200
from mcrt_b200 import rr, pt, scene_library, layouts as L
from mcrt_b200.pt import _lib, _check


def main():
    tex = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    w, h, depth = 960, 540, 5
    scene = scene_library.sponza_class(tex)
    api = rr.IntersectionApi.Create(0)
    pl = pt.make_pt_pipeline(api, scene, w, h, depth, fused=False)
    pl.ensure_size()
    hs = pl.scene.host
    light_of_shape = np.asarray(hs.shapes["lightID"]).copy()
    # drive the stages by hand (RTPathTracingPass.update, fused=False) with read-backs in between
    pl.set_camera_for_frame()
    hd = pl.scene._h
    _check(_lib().ptb_generate_primary_rays(hd))
    rp, rn = pl.device_buffer("RayBuffer")
    ip, inb = pl.device_buffer("PrimaryIntersectionBufferCL")
    sp, sn = pl.device_buffer("ShadowRayBuffer")
    op, on = pl.device_buffer("ShadowRayOcclusionBuffer")
    rays, isect = api.CreateFromDevicePtr(rp, rn), api.CreateFromDevicePtr(ip, inb)
    shadow, occl = api.CreateFromDevicePtr(sp, sn), api.CreateFromDevicePtr(op, on)
    api.QueryIntersection(rays, w * h, isect).Wait()
    P = w * h
    for b in range(depth):
        r = pl.read_buffer("RayBuffer", np.uint8).view(L.RAY)
        it = pl.read_buffer("PrimaryIntersectionBufferCL", np.uint8).view(L.INTERSECTION)
        active = r["extra"][:, 1] != 0
        hit = active & (it["shapeid"] != -1)
        emit = np.zeros(P, bool)
        if light_of_shape is not None:
            sid = np.clip(it["shapeid"], 0, len(light_of_shape) - 1)
            emit = hit & (light_of_shape[sid] != -1)
        _check(_lib().ptb_pt_shade(hd, 0, depth, b))
        sr = pl.read_buffer("ShadowRayBuffer", np.uint8).view(L.RAY)
        tr = pl.read_buffer("TempRadianceBuffer").reshape(-1, 4)
        r2 = pl.read_buffer("RayBuffer", np.uint8).view(L.RAY)
        print("bounce %d: live rays %8d (%.3f of P)  hits %8d  misses %7d  emitter hits %7d  non-black NEE %8d  continues %8d" %
              (b, active.sum(), active.mean(), hit.sum(), (active & ~hit).sum(), emit.sum(),
               (hit & (tr[:, :3].max(1) > 0)).sum(), (r2["extra"][:, 1] != 0).sum()))
        api.QueryOcclusion(shadow, P, occl).Wait()
        _check(_lib().ptb_pt_shadow_resolve(hd, b))
        if b + 1 < depth:
            api.QueryIntersection(rays, P, isect).Wait()


main()
