"""Traversal micro-benchmark T: Mrays/s of QueryIntersection / QueryOcclusion on the
~1 M-triangle knot for primary, incoherent (first-bounce cosine) and shadow ray sets, with
per-ray visited-node / tested-triangle counts from the instrumented kernel."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mcrt_b200  # noqa
from mcrt_b200 import rr, scene_library, layouts as L, scenes
from raygen import camera_rays


def cosine_bounce_rays(rays, hits, scene_tris_normals=None, seed=1):
    """Incoherent set: origin = primary hit point, direction = cosine hemisphere around the (approximate)
    geometric normal facing the viewer, from a wang_hash-seeded xorshift stream."""
    n = rays.shape[0]
    hit = hits["shapeid"] >= 0
    t = hits["uvwt"][:, 3:4]
    p = rays["o"][:, :3] + t * rays["d"][:, :3]
    s = scenes.wang_hash(np.arange(n, dtype=np.uint32) + np.uint32(seed * 7919))
    s, u1 = scenes.rand_float(s)
    s, u2 = scenes.rand_float(s)
    # uniform direction on the sphere, flipped against the incoming ray (cheap stand-in for the normal)
    z = 1.0 - 2.0 * u1
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    phi = 2 * np.pi * u2
    d = np.stack([r * np.cos(phi), z, r * np.sin(phi)], 1).astype(np.float32)
    flip = (d * rays["d"][:, :3]).sum(1) > 0
    d[flip] *= -1
    out = L.make_rays(n)
    out["o"][:, :3] = p - rays["d"][:, :3] * 1e-3
    out["o"][:, 3] = 1000.0
    out["d"][:, :3] = d
    out["extra"][:, 1] = hit.astype(np.int32)
    return out[hit]


def time_query(api, fn, rb, n, ob, reps):
    ev = fn(rb, n, ob)
    ev.Wait()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn(rb, n, ob).Wait()
        ts.append(time.perf_counter() - t0)
    return min(ts), float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quads", type=int, default=707)
    ap.add_argument("--res", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--refit", action="store_true", help="also time a device refit against a full rebuild")
    ap.add_argument("--builder", default="sah", choices=["sah", "median", "lbvh"],
                    help="lbvh = the device builder (what acc.type = hlbvh selects)")
    a = ap.parse_args()
    scene = scene_library.traversal_bench_scene(a.quads)
    api = rr.IntersectionApi.Create(0)
    shapes = rr.attach_scene(api, scene)
    api.SetOption("bvh.builder", a.builder)
    t0 = time.time()
    api.Commit()
    st = api.GetStats()
    rebuilds = []
    for k in range(3):           # warm rebuilds (the first device build also pays CUDA module load)
        shapes[0].SetTransform(np.eye(4, dtype=np.float32))
        t1 = time.perf_counter()
        api.Commit()
        rebuilds.append(round(time.perf_counter() - t1, 4))
    print(json.dumps({"builder": a.builder, "on_device": st["last_build_on_device"], "levels": st["num_levels"],
                      "tris": st["num_triangles"], "nodes": st["num_nodes"], "build_s": st["build_seconds"],
                      "rebuild_s": rebuilds,
                      "node_MB": st["node_bytes"] / 1e6, "tri_MB": st["triangle_bytes"] / 1e6}))
    prim, _ = camera_rays(scene, a.res, a.res)
    hits = api.QueryIntersectionHost(prim)
    sets = {"primary": prim, "incoherent": cosine_bounce_rays(prim, hits)}
    sh = sets["incoherent"].copy()
    sh["o"][:, 3] = 6.0
    sets["shadow"] = sh
    for name, rays in sets.items():
        n = rays.shape[0]
        rb = api.CreateBuffer(n * 48, rays)
        hb = api.CreateBuffer(n * 32)
        ob = api.CreateBuffer(n * 4)
        cb = api.CreateBuffer(n * 8)
        tmin, tmed = time_query(api, api.QueryIntersection, rb, n, hb, a.reps)
        omin, omed = time_query(api, api.QueryOcclusion, rb, n, ob, a.reps)
        api.QueryIntersectionCounted(rb, n, hb, cb)
        cnt, ev = api.MapBuffer(cb, rr.MAP_READ, 0, n * 8, dtype=np.int32)
        ev.Wait()
        c = cnt.reshape(-1, 2).astype(np.float64).mean(0)
        api.UnmapBuffer(cb)
        h, ev = api.MapBuffer(hb, rr.MAP_READ, 0, n * 32, dtype=L.INTERSECTION)
        ev.Wait()
        hitfrac = float((h["shapeid"] >= 0).mean())
        api.UnmapBuffer(hb)
        print(json.dumps({"set": name, "rays": n, "closest_Mrays_s": n / tmin / 1e6, "closest_med_Mrays_s": n / tmed / 1e6,
                          "anyhit_Mrays_s": n / omin / 1e6, "nodes_per_ray": c[0], "tris_per_ray": c[1],
                          "bytes_per_ray": c[0] * 80 + c[1] * 48, "hit_frac": hitfrac}))
        for b in (rb, hb, ob, cb):
            api.DeleteBuffer(b)
    if a.refit:
        refit_section(scene, sets["incoherent"], a.reps)


def refit_section(scene, rays, reps):
    """Commit after a transform change: full host rebuild vs device refit (option bvh.refit), and what the kept
    topology costs in traversal speed after every shape was rotated by 30 degrees."""
    def rot_y(deg):
        c, s_ = np.cos(np.deg2rad(deg)), np.sin(np.deg2rad(deg))
        m = np.eye(4, dtype=np.float32)
        m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s_, -s_, c
        return m

    out = {}
    for mode in ("rebuild", "refit"):
        api = rr.IntersectionApi.Create(0)
        shapes = rr.attach_scene(api, scene)
        if mode == "refit":
            api.SetOption("bvh.refit", 1.0)
        api.Commit()
        ts = []
        for k in range(3):
            for sh in shapes:
                sh.SetTransform(rot_y(10.0 * (k + 1)))
            t0 = time.perf_counter()
            api.Commit()
            ts.append(time.perf_counter() - t0)
        n = rays.shape[0]
        rb, hb = api.CreateBuffer(n * 48, rays), api.CreateBuffer(n * 32)
        tmin, _ = time_query(api, api.QueryIntersection, rb, n, hb, reps)
        out[mode] = {"commit_ms": [round(t * 1e3, 3) for t in ts], "incoherent_closest_Mrays_s": n / tmin / 1e6,
                     "last_commit_refit": api.GetStats()["last_commit_refit"]}
        rr.IntersectionApi.Delete(api)
    print(json.dumps({"refit_vs_rebuild": out}))


if __name__ == "__main__":
    main()
