"""Where along a slow ray does its traversal work lie?  C5's scene (10 M-triangle knot), 4K primary rays and their
first-bounce rays through the instrumented closest-hit query; for the costliest rays, the same ray is queried with
64 increasing tmax values: W(tmax) = nodes + triangles visited is the cumulative work up to that distance.
  gpurun -- python tools/straggler_profile.py [hlbvh|fatbvh]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200  # noqa: E402,F401
from mcrt_b200 import rr, scene_library, traversal_bench as tb, layouts as L  # noqa: E402


def counted(api, rays):
    n = rays.shape[0]
    rb, hb, cb = api.CreateBuffer(n * 48, rays), api.CreateBuffer(n * 32), api.CreateBuffer(n * 8)
    api.QueryIntersectionCounted(rb, n, hb, cb)
    cnt, ev = api.MapBuffer(cb, rr.MAP_READ, 0, n * 8, dtype=np.int32)
    ev.Wait()
    c = cnt.reshape(-1, 2).copy()
    api.UnmapBuffer(cb)
    h, ev = api.MapBuffer(hb, rr.MAP_READ, 0, n * 32, dtype=L.INTERSECTION)
    ev.Wait()
    hits = h.copy()
    api.UnmapBuffer(hb)
    for b in (rb, hb, cb):
        api.DeleteBuffer(b)
    return c, hits


def main():
    acc = sys.argv[1] if len(sys.argv) > 1 else "hlbvh"
    scene = scene_library.traversal_bench_scene(2236)
    api = rr.IntersectionApi.Create(0)
    api.SetOption("acc.type", acc)
    rr.attach_scene(api, scene)
    api.Commit()
    prim = tb.camera_rays(scene, 1920, 1080)
    c, hits = counted(api, prim)
    idx, p, gn = tb.hit_frames(scene, prim, hits)
    sets = {"primary": (prim, c)}
    b = tb.bounce_rays(p, gn)
    sets["bounce"] = (b, counted(api, b)[0])
    for name, (rays, cnt) in sets.items():
        w = cnt.sum(1)
        print("%s %s: rays %d  mean work %.1f  p99 %.0f  p99.99 %.0f  max %d (nodes %d, tris %d)" % (
            acc, name, rays.shape[0], w.mean(), np.percentile(w, 99), np.percentile(w, 99.99), w.max(),
            cnt[w.argmax(), 0], cnt[w.argmax(), 1]))
        for r in np.argsort(-w)[:4]:
            ray = rays[r:r + 1]
            full = int(w[r])
            # how far the ray is inside the scene at all
            lo, hi = (np.asarray(x, np.float64) for x in scene.world_bounds())
            o, d = ray["o"][0, :3].astype(np.float64), ray["d"][0, :3].astype(np.float64)
            with np.errstate(divide="ignore", invalid="ignore"):
                ta, tb_ = (lo - o) / d, (hi - o) / d
            t_in, t_out = max(0.0, np.nanmax(np.minimum(ta, tb_))), np.nanmin(np.maximum(ta, tb_))
            K = 64
            probe = np.repeat(ray, K)
            ts = t_in + (t_out - t_in) * (np.arange(1, K + 1) / K)
            probe["o"][:, 3] = ts.astype(np.float32)
            cw = counted(api, probe)[0].sum(1)
            dw = np.diff(np.concatenate([[0], cw]))
            top = np.argsort(-dw)[:3]
            print("   ray %d work %d, inside the scene for t in [%.3f, %.3f]; work per 1/64 of that stretch: max %d, "
                  "top-3 slices %s hold %.0f %% ; slices with > 5 %% of the work: %d ; W(t) at 1/4, 1/2, 3/4: %d %d %d" % (
                      r, full, t_in, t_out, dw.max(), sorted(top.tolist()), 100.0 * dw[top].sum() / max(cw[-1], 1),
                      int((dw > 0.05 * cw[-1]).sum()), cw[15], cw[31], cw[47]))
    rr.IntersectionApi.Delete(api)


if __name__ == "__main__":
    main()
