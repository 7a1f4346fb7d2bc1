"""Diagnostic: distribution of per-ray traversal work (visited nodes, tested triangles) -- looks for the few rays
that keep a persistent traversal kernel alive after all the others are done."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import mcrt_b200
from mcrt_b200 import rr, scene_library, layouts as L
from raygen import camera_rays

quads = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
acc = sys.argv[2] if len(sys.argv) > 2 else "hlbvh"
scene = scene_library.traversal_bench_scene(quads)
api = rr.IntersectionApi.Create(0)
rr.attach_scene(api, scene)
api.SetOption("acc.type", acc)
api.Commit()
rays, _ = camera_rays(scene, 1920, 1080)
n = rays.shape[0]
rb, hb, cb = api.CreateBuffer(n * 48, rays), api.CreateBuffer(n * 32), api.CreateBuffer(n * 8)
api.QueryIntersectionCounted(rb, n, hb, cb)
cnt, ev = api.MapBuffer(cb, rr.MAP_READ, 0, n * 8, dtype=np.int32)
ev.Wait()
c = cnt.reshape(-1, 2).copy()
api.UnmapBuffer(cb)
for k, name in enumerate(("nodes", "tris")):
    v = np.sort(c[:, k])
    print(name, "mean %.2f" % v.mean(), "p99", v[int(0.99 * n)], "p99.99", v[int(0.9999 * n)], "max", v[-1],
          "rays > 200:", int((v > 200).sum()), "> 1000:", int((v > 1000).sum()))
worst = np.argsort(c[:, 0])[-5:]
for i in worst:
    print("ray", i, "px", (i % 1920, i // 1920), "nodes", c[i, 0], "tris", c[i, 1], "d", rays["d"][i, :3], "o", rays["o"][i])
print(json.dumps(api.GetStats()))
