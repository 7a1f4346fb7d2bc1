"""Host-side tree-quality measurement (development tool): builds the benchmark scene's 8-wide tree with the HOST builder
(no GPU), walks two sets of incoherent rays through it with tools/bvh_quality.cpp (the control flow of
csrc/traverse.cuh) and prints node visits / triangle tests / fetched bytes per ray.

  python tools/bvh_quality.py [--scene sponza|knot|cornell] [--rays 200000]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from mcrt_b200 import rr, scene_library  # noqa: E402


def scene_soup(scene):
    verts, idx, ids, base = [], [], [], 0
    for i, m in enumerate(scene.meshes):
        verts.append(m.positions)
        idx.append(m.indices + base)
        ids.append(np.full(m.indices.shape[0], i, np.int32))
        base += m.positions.shape[0]
    return np.concatenate(verts).astype(np.float32), np.concatenate(idx).astype(np.int32), np.concatenate(ids)


def surface_rays(v, idx, n, rng):
    """Cosine-distributed rays leaving random surface points (area-weighted) on the side that faces the scene centre."""
    a, b, c = v[idx[:, 0]].astype(np.float64), v[idx[:, 1]].astype(np.float64), v[idx[:, 2]].astype(np.float64)
    cr = np.cross(b - a, c - a)
    area = 0.5 * np.linalg.norm(cr, axis=1)
    pick = rng.choice(idx.shape[0], n, p=area / area.sum())
    u, w = rng.rand(n), rng.rand(n)
    flip = u + w > 1
    u[flip], w[flip] = 1 - u[flip], 1 - w[flip]
    p = a[pick] + u[:, None] * (b[pick] - a[pick]) + w[:, None] * (c[pick] - a[pick])
    nrm = cr[pick] / np.maximum(np.linalg.norm(cr[pick], axis=1, keepdims=True), 1e-30)
    centre = 0.5 * (v.min(0) + v.max(0)).astype(np.float64)
    nrm[np.einsum("ij,ij->i", nrm, centre - p) < 0] *= -1
    r1, r2 = rng.rand(n), rng.rand(n)
    phi, st, ct = 2 * np.pi * r1, np.sqrt(r2), np.sqrt(1 - r2)
    t = np.where(np.abs(nrm[:, :1]) > 0.9, np.array([[0, 1.0, 0]]), np.array([[1.0, 0, 0]]))
    t = np.cross(nrm, t)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    bt = np.cross(nrm, t)
    d = st[:, None] * (np.cos(phi)[:, None] * t + np.sin(phi)[:, None] * bt) + ct[:, None] * nrm
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = p + 1e-3 * nrm
    rays[:, 3] = 1e30
    rays[:, 4:7] = d
    return rays


def volume_rays(v, n, rng):
    lo, hi = v.min(0), v.max(0)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = lo + rng.rand(n, 3) * (hi - lo)
    d = rng.randn(n, 3)
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 3] = 1e30
    return rays


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="sponza")
    ap.add_argument("--rays", type=int, default=200000)
    ap.add_argument("--order", default="", help="ideal = nearest first with re-test on pop (a bound); sorted = nearest first only")
    args = ap.parse_args()
    if args.scene == "sponza":
        scene = scene_library.sponza_class(16)
    elif args.scene == "knot":
        scene = scene_library.traversal_bench_scene(707)
    else:
        scene = scene_library.cornell_box(True)
    v, idx, ids = scene_soup(scene)
    exe = os.path.join(tempfile.gettempdir(), "bvh_quality")
    subprocess.check_call(["g++", "-O2", "-o", exe, os.path.join(ROOT, "tools", "bvh_quality.cpp")])
    import time
    t0 = time.time()
    nodes, tris = rr.host_bvh(v, idx, ids)
    build_s = time.time() - t0
    rng = np.random.RandomState(7)
    with tempfile.TemporaryDirectory() as td:
        nodes.tofile(os.path.join(td, "n"))
        tris.tofile(os.path.join(td, "t"))
        for name, rays in (("surface", surface_rays(v, idx, args.rays, rng)), ("volume", volume_rays(v, args.rays, rng))):
            rays.tofile(os.path.join(td, "r"))
            out = subprocess.check_output([exe, os.path.join(td, "n"), os.path.join(td, "t"), os.path.join(td, "r")] +
                                          ([args.order] if args.order else []))
            d = json.loads(out)
            d["set"], d["build_s"] = name, round(build_s, 3)
            print(json.dumps(d))


if __name__ == "__main__":
    main()
