"""Traversal micro-benchmark T: Grays/s of `QueryIntersection` / `QueryOcclusion` through the
IntersectionApi mirror on three ray sets of res x res rays each (the survey's 2^24 = 4096 x 4096):

  primary     coherent camera rays
  incoherent  real first-bounce rays: origin = primary hit point + gn * 1e-5 (gn turned against the incoming ray),
              direction = cosine-hemisphere sample around gn from a wang_hash-seeded xorshift stream, tmax = 1000
  shadow      the same origins towards uniformly sampled points on the scene's area light, tmax = distance

Ray buffers are produced once on the host and reused; a query is timed with CUDA events on the API's stream
(median of `reps` launches after one warm-up).  Per-ray visited-node / tested-triangle counts come from the
instrumented closest-hit kernel (QueryIntersectionCounted).  Product-side measurement code: no oracle involved.
"""
import numpy as np

from . import layouts as L
from . import rr
from . import scenes


def camera_rays(scene, width, height):
    """Pixel-corner primary rays of `scene.camera` (PathTracing.cl:13-35), as RadeonRays::ray records."""
    from . import camera
    cam = camera.make_camera(scene.camera, width, height, (0.0, 0.0))
    xs = (np.arange(width, dtype=np.float32) / np.float32(width))[None, :, None]
    ys = (np.arange(height, dtype=np.float32) / np.float32(height))[:, None, None]
    r00, r10, r11, r01 = (cam[k][0][:3].astype(np.float32) for k in ("r00", "r10", "r11", "r01"))
    top = r00 + (r10 - r00) * xs
    bot = r01 + (r11 - r01) * xs
    d = top + (bot - top) * ys
    d = (d / np.linalg.norm(d, axis=2, keepdims=True)).astype(np.float32).reshape(-1, 3)
    rays = L.make_rays(width * height)
    rays["o"][:, :3] = cam["pos"][0][:3]
    rays["o"][:, 3] = 1000.0
    rays["d"][:, :3] = d
    return rays


def hit_frames(scene, rays, hits):
    """Hit point and geometric normal (turned against the ray) of every ray that hit."""
    hit = hits["shapeid"] >= 0
    idx = np.nonzero(hit)[0]
    gn = np.zeros((idx.size, 3), np.float32)
    sid, pid = hits["shapeid"][idx], hits["primid"][idx]
    for s in np.unique(sid):
        m = scene.meshes[int(s)]
        sel = sid == s
        tri = m.indices[pid[sel]]
        pos = m.positions
        if m.transform is not None:
            pos = (pos @ m.transform[:3, :3].T + m.transform[:3, 3]).astype(np.float32)
        p0, p1, p2 = pos[tri[:, 0]], pos[tri[:, 1]], pos[tri[:, 2]]
        n = np.cross(p1 - p0, p2 - p0)
        gn[sel] = n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
    d = rays["d"][idx, :3]
    gn[(gn * d).sum(1) > 0] *= -1.0
    p = rays["o"][idx, :3] + hits["uvwt"][idx, 3:4] * d + gn * np.float32(1e-5)
    return idx, p.astype(np.float32), gn


def bounce_rays(p, gn, seed=1):
    n = p.shape[0]
    s = scenes.wang_hash(np.arange(n, dtype=np.uint32) + np.uint32(seed * 7919))
    s, u1 = scenes.rand_float(s)
    s, u2 = scenes.rand_float(s)
    r = np.sqrt(u1).astype(np.float32)
    phi = (2.0 * np.pi * u2).astype(np.float32)
    lx, lz = r * np.cos(phi), r * np.sin(phi)
    ly = np.sqrt(np.maximum(0.0, 1.0 - u1)).astype(np.float32)
    a = np.where(np.abs(gn[:, 0:1]) > 0.9, np.float32([[0, 1, 0]]), np.float32([[1, 0, 0]]))
    t = np.cross(a, gn)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(gn, t)
    d = lx[:, None] * t + ly[:, None] * gn + lz[:, None] * b
    out = L.make_rays(n)
    out["o"][:, :3] = p
    out["o"][:, 3] = 1000.0
    out["d"][:, :3] = d / np.linalg.norm(d, axis=1, keepdims=True)
    return out


def light_points(scene, n, seed=2):
    """Uniform points on the scene's first area light (disk or triangle mesh)."""
    s = scenes.wang_hash(np.arange(n, dtype=np.uint32) + np.uint32(seed * 104729))
    s, u1 = scenes.rand_float(s)
    s, u2 = scenes.rand_float(s)
    for lt in scene.lights:
        if int(lt["type"]) == L.RT_DISK_AREA_LIGHT:
            r = float(lt["radius"]) * np.sqrt(u1)
            phi = 2.0 * np.pi * u2
            nrm = lt["d"][:3].astype(np.float64)
            a = np.array([0.0, 1.0, 0.0]) if abs(nrm[0]) > 0.9 else np.array([1.0, 0.0, 0.0])
            t = np.cross(a, nrm)
            t /= np.linalg.norm(t)
            b = np.cross(nrm, t)
            return (lt["p"][:3] + (r * np.cos(phi))[:, None] * t + (r * np.sin(phi))[:, None] * b).astype(np.float32)
        if int(lt["type"]) == L.RT_TRIANGLE_MESH_AREA_LIGHT:
            m = scene.meshes[int(lt["shapeId"])]
            s, u3 = scenes.rand_float(s)
            tri = m.indices[np.minimum((u3 * m.indices.shape[0]).astype(np.int64), m.indices.shape[0] - 1)]
            pos = m.positions
            if m.transform is not None:
                pos = (pos @ m.transform[:3, :3].T + m.transform[:3, 3]).astype(np.float32)
            su = np.sqrt(u1)[:, None]
            b0, b1 = 1.0 - su, u2[:, None] * su
            return (b0 * pos[tri[:, 0]] + b1 * pos[tri[:, 1]] + (1.0 - b0 - b1) * pos[tri[:, 2]]).astype(np.float32)
    lo, hi = scene.world_bounds()
    return np.broadcast_to(((lo + hi) * 0.5 + np.array([0, (hi - lo)[1], 0])).astype(np.float32), (n, 3)).copy()


def shadow_rays(scene, p):
    n = p.shape[0]
    tgt = light_points(scene, n)
    d = tgt - p
    dist = np.linalg.norm(d, axis=1)
    out = L.make_rays(n)
    out["o"][:, :3] = p
    out["o"][:, 3] = dist
    out["d"][:, :3] = d / np.maximum(dist, 1e-20)[:, None]
    return out


def _time(fn, stream, reps):
    import torch
    fn().Wait()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ev = fn()
        e1.record(stream)
        ev.Wait()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    return float(np.median(ts))


def run(api, scene, stream, res=4096, reps=3):
    """api: a committed rr.IntersectionApi created on `stream` (a torch.cuda.Stream).  Returns
    {set: {rays, closest_grays_s, anyhit_grays_s, nodes_per_ray, tris_per_ray, hit_frac}}."""
    prim = camera_rays(scene, res, res)
    hits = api.QueryIntersectionHost(prim)
    idx, p, gn = hit_frames(scene, prim, hits)
    sets = {"primary": prim, "incoherent": bounce_rays(p, gn), "shadow": shadow_rays(scene, p)}
    out = {}
    for name, rays in sets.items():
        n = rays.shape[0]
        rb, hb, ob, cb = api.CreateBuffer(n * 48, rays), api.CreateBuffer(n * 32), api.CreateBuffer(n * 4), api.CreateBuffer(n * 8)
        t_c = _time(lambda: api.QueryIntersection(rb, n, hb), stream, reps)
        t_a = _time(lambda: api.QueryOcclusion(rb, n, ob), stream, reps)
        api.QueryIntersectionCounted(rb, n, hb, cb)
        cnt, ev = api.MapBuffer(cb, rr.MAP_READ, 0, n * 8, dtype=np.int32)
        ev.Wait()
        c = cnt.reshape(-1, 2).astype(np.float64).mean(0)
        api.UnmapBuffer(cb)
        h, ev = api.MapBuffer(hb, rr.MAP_READ, 0, n * 32, dtype=L.INTERSECTION)
        ev.Wait()
        hit_frac = float((h["shapeid"] >= 0).mean())
        api.UnmapBuffer(hb)
        for b in (rb, hb, ob, cb):
            api.DeleteBuffer(b)
        out[name] = {"rays": int(n), "closest_grays_s": n / t_c / 1e9, "anyhit_grays_s": n / t_a / 1e9,
                     "nodes_per_ray": float(c[0]), "tris_per_ray": float(c[1]),
                     "l2_gather_bytes_per_ray": float(c[0] * 80 + c[1] * 48), "hit_frac": hit_frac}
    return out
