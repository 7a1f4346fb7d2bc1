"""The BASELINE scene configs C1..C5 and the traversal micro-benchmark scene T.

C1/C4 geometry is this repo's own statement of the classic Cornell-box measurements (public
domain data, Cornell Program of Computer Graphics) in the form the reference's importer hands to
RTScene: z negated and winding flipped (assimp MakeLeftHanded | FlipWindingOrder,
engine/resource/AssetImporter.cpp:40), scaled by 5 (PathTracingApp.cpp:371), one RTShape per
group, flat per-face normals.  The real Dragon / Sponza / San Miguel meshes are not in the
reference tree (.MISSING_LARGE_BLOBS), so C2/C3/C5 are procedural stand-ins of the named
triangle counts, as BASELINE.json prescribes.
"""
import numpy as np

from . import layouts as L
from .scenes import (CameraSpec, Mesh, Scene, directional_light, disk_light, lambert, make_material, mesh_area,
                     point_light, tri_mesh_light, value_noise, hash01)

f4 = np.float32

# Cornell box quads (right-handed OBJ space, metres); each row = 4 corners, counter-clockwise seen
# from inside the box.
_CORNELL = {
    "floor": [[(-1.01, 0.00, 0.99), (1.00, 0.00, 0.99), (1.00, 0.00, -1.04), (-0.99, 0.00, -1.04)]],
    "ceiling": [[(-1.02, 1.99, 0.99), (-1.02, 1.99, -1.04), (1.00, 1.99, -1.04), (1.00, 1.99, 0.99)]],
    "backWall": [[(-0.99, 0.00, -1.04), (1.00, 0.00, -1.04), (1.00, 1.99, -1.04), (-1.02, 1.99, -1.04)]],
    "rightWall": [[(1.00, 0.00, -1.04), (1.00, 0.00, 0.99), (1.00, 1.99, 0.99), (1.00, 1.99, -1.04)]],
    "leftWall": [[(-1.01, 0.00, 0.99), (-0.99, 0.00, -1.04), (-1.02, 1.99, -1.04), (-1.02, 1.99, 0.99)]],
    "shortBox": [
        [(0.53, 0.60, 0.75), (0.70, 0.60, 0.17), (0.13, 0.60, 0.00), (-0.05, 0.60, 0.57)],
        [(-0.05, 0.00, 0.57), (-0.05, 0.60, 0.57), (0.13, 0.60, 0.00), (0.13, 0.00, 0.00)],
        [(0.53, 0.00, 0.75), (0.53, 0.60, 0.75), (-0.05, 0.60, 0.57), (-0.05, 0.00, 0.57)],
        [(0.70, 0.00, 0.17), (0.70, 0.60, 0.17), (0.53, 0.60, 0.75), (0.53, 0.00, 0.75)],
        [(0.13, 0.00, 0.00), (0.13, 0.60, 0.00), (0.70, 0.60, 0.17), (0.70, 0.00, 0.17)],
        [(0.53, 0.00, 0.75), (0.70, 0.00, 0.17), (0.13, 0.00, 0.00), (-0.05, 0.00, 0.57)],
    ],
    "tallBox": [
        [(-0.53, 1.20, 0.09), (0.04, 1.20, -0.09), (-0.14, 1.20, -0.67), (-0.71, 1.20, -0.49)],
        [(-0.53, 0.00, 0.09), (-0.53, 1.20, 0.09), (-0.71, 1.20, -0.49), (-0.71, 0.00, -0.49)],
        [(-0.71, 0.00, -0.49), (-0.71, 1.20, -0.49), (-0.14, 1.20, -0.67), (-0.14, 0.00, -0.67)],
        [(-0.14, 0.00, -0.67), (-0.14, 1.20, -0.67), (0.04, 1.20, -0.09), (0.04, 0.00, -0.09)],
        [(0.04, 0.00, -0.09), (0.04, 1.20, -0.09), (-0.53, 1.20, 0.09), (-0.53, 0.00, 0.09)],
        [(-0.53, 0.00, 0.09), (0.04, 0.00, -0.09), (-0.14, 0.00, -0.67), (-0.71, 0.00, -0.49)],
    ],
    "light": [[(-0.24, 1.98, 0.16), (-0.24, 1.98, -0.22), (0.23, 1.98, -0.22), (0.23, 1.98, 0.16)]],
}
_CORNELL_KD = {
    "floor": (0.725, 0.71, 0.68), "ceiling": (0.725, 0.71, 0.68), "backWall": (0.725, 0.71, 0.68),
    "rightWall": (0.14, 0.45, 0.091), "leftWall": (0.63, 0.065, 0.05), "shortBox": (0.725, 0.71, 0.68),
    "tallBox": (0.725, 0.71, 0.68), "light": (0.78, 0.78, 0.78),
}
CORNELL_CAMERA = CameraSpec(position=(-0.324, 4.428, -13.0), euler_deg=(0.0, 0.0, 0.0))


def quads_to_mesh(name, quads, material, scale=5.0) -> Mesh:
    """Quads -> 2 triangles each, left-handed flip (z negated, winding reversed), flat normals."""
    pos, nrm, uvs, idx = [], [], [], []
    for q in quads:
        p = np.asarray(q, np.float64) * scale
        p[:, 2] = -p[:, 2]
        n = np.cross(p[2] - p[0], p[1] - p[0])          # winding flipped -> normal of reversed order
        n = n / np.linalg.norm(n)
        base = len(pos)
        pos.extend(p)
        nrm.extend([n] * 4)
        uvs.extend([(0, 0), (1, 0), (1, 1), (0, 1)])
        idx.append((base + 0, base + 2, base + 1))
        idx.append((base + 0, base + 3, base + 2))
    return Mesh(name, np.asarray(pos, f4), np.asarray(nrm, f4), np.asarray(uvs, f4), np.asarray(idx, np.int32),
                material)


def uv_sphere(name, center, radius, n_lon, n_lat, material) -> Mesh:
    """Closed UV sphere, smooth normals; 2*n_lon*(n_lat-1) triangles."""
    pos, nrm, uvs, idx = [], [], [], []
    for j in range(n_lat + 1):
        th = np.pi * j / n_lat
        for i in range(n_lon + 1):
            ph = 2 * np.pi * i / n_lon
            d = np.array([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)])
            pos.append(np.asarray(center) + radius * d)
            nrm.append(d)
            uvs.append((i / n_lon, j / n_lat))
    w = n_lon + 1
    for j in range(n_lat):
        for i in range(n_lon):
            a, b, c, d = j * w + i, j * w + i + 1, (j + 1) * w + i, (j + 1) * w + i + 1
            if j != 0:
                idx.append((a, b, c))
            if j != n_lat - 1:
                idx.append((b, d, c))
    return Mesh(name, np.asarray(pos, f4), np.asarray(nrm, f4), np.asarray(uvs, f4), np.asarray(idx, np.int32),
                material)


def cornell_box(with_spheres: bool = False) -> Scene:
    """C1 (Original, 36 triangles) or C4 (walls + mirror sphere + glass sphere, BDPT config)."""
    groups = ["floor", "ceiling", "backWall", "rightWall", "leftWall"]
    if not with_spheres:
        groups += ["shortBox", "tallBox"]
    groups.append("light")
    meshes, mats = [], []
    for g in groups:
        mats.append(lambert(_CORNELL_KD[g]))
        # roughness = clamp(sqrt(2/(Ns+2)),1e-5,1) with Ns=10 (RTScene.cpp:840); unused while ks == 0
        mats[-1]["uber_roughness"] = np.sqrt(2.0 / 12.0)
        meshes.append(quads_to_mesh(g, _CORNELL[g], len(mats) - 1))
    if with_spheres:
        # left: perfect mirror (kr = 1); right: perfect specular transmission (kt.w = 0 -> delta), eta 1.5
        mats.append(make_material(kr=(1, 1, 1)))
        meshes.insert(-1, uv_sphere("leftSphere", (-2.1, 1.65, -1.6), 1.65, 36, 16, len(mats) - 1))
        mats.append(make_material(kt=(1, 1, 1, 0), eta=1.5))
        meshes.insert(-1, uv_sphere("rightSphere", (2.2, 1.65, 1.3), 1.65, 36, 16, len(mats) - 1))
    light_shape = len(meshes) - 1
    lights = tri_mesh_light(light_shape, (17.0, 12.0, 4.0), mesh_area(meshes[light_shape]))
    sc = Scene("cornell_spheres" if with_spheres else "cornell", meshes, np.concatenate(mats), lights, CORNELL_CAMERA)
    return sc.finalize()


# ----------------------------------------------------------------------------- C2 / T: torus knot
def grid_mesh(name, P, N, UV, material, closed_u=False, closed_v=False) -> Mesh:
    """Triangulate a (nu+1) x (nv+1) lattice of points P[i,j] (2 triangles per quad)."""
    nu, nv = P.shape[0] - 1, P.shape[1] - 1
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    w = nv + 1
    a = (i * w + j).reshape(-1)
    b = ((i + 1) * w + j).reshape(-1)
    c = (i * w + j + 1).reshape(-1)
    d = ((i + 1) * w + j + 1).reshape(-1)
    idx = np.empty((2 * a.size, 3), np.int32)
    idx[0::2] = np.stack([a, b, c], 1)
    idx[1::2] = np.stack([b, d, c], 1)
    return Mesh(name, P.reshape(-1, 3).astype(f4), N.reshape(-1, 3).astype(f4), UV.reshape(-1, 2).astype(f4), idx,
                material)


def torus_knot_mesh(nu, nv, seed=2, p=2, q=3, tube=0.45, material=0, fit=10.0) -> Mesh:
    """(p,q) torus knot tube, nu x nv quads -> 2*nu*nv triangles, smooth normals, radius perturbed
    by 1 + 0.02*(hash-0.5), scaled to fit a `fit`-unit box resting on y = 0."""
    u = np.linspace(0.0, 2 * np.pi, nu + 1)
    v = np.linspace(0.0, 2 * np.pi, nv + 1)
    r = 2.0 + np.cos(q * u)
    c = np.stack([r * np.cos(p * u), -np.sin(q * u), r * np.sin(p * u)], 1)
    dc = np.gradient(c, u, axis=0)
    t = dc / np.linalg.norm(dc, axis=1, keepdims=True)
    up = np.array([0.0, 1.0, 0.0])
    n1 = np.cross(t, up)
    n1 /= np.linalg.norm(n1, axis=1, keepdims=True)
    n2 = np.cross(t, n1)
    ii, jj = np.meshgrid(np.arange(nu + 1) % nu, np.arange(nv + 1) % nv, indexing="ij")
    pert = 1.0 + 0.02 * (hash01(ii, jj, seed) - 0.5)
    nrm = n1[:, None, :] * np.cos(v)[None, :, None] + n2[:, None, :] * np.sin(v)[None, :, None]
    P = c[:, None, :] + nrm * (tube * pert)[:, :, None]
    lo, hi = P.reshape(-1, 3).min(0), P.reshape(-1, 3).max(0)
    s = fit / (hi - lo).max()
    P = (P - np.array([0.5 * (lo[0] + hi[0]), lo[1], 0.5 * (lo[2] + hi[2])])) * s
    UV = np.stack(np.meshgrid(np.arange(nu + 1) / nu, np.arange(nv + 1) / nv, indexing="ij"), -1)
    m = grid_mesh("knot", P, nrm, UV, material)
    m.indices = m.indices[:, [0, 2, 1]].copy()      # geometric normal (reference convention) = outward
    return m


def floor_mesh(name, half, y, material, n=1) -> Mesh:
    g = np.linspace(-half, half, n + 1)
    X, Z = np.meshgrid(g, g, indexing="ij")
    P = np.stack([X, np.full_like(X, y), Z], -1)
    N = np.zeros_like(P)
    N[..., 1] = 1.0
    UV = np.stack([(X + half) / (2 * half), (Z + half) / (2 * half)], -1)
    m = grid_mesh(name, P, N, UV, material)
    m.indices = m.indices[:, [0, 2, 1]].copy()      # face up (+y) with the reference's gn convention
    return m


def dragon_class(nu=660, nv=660) -> Scene:
    """C2: 871,200-triangle glossy-dielectric knot (Trowbridge-Reitz BRDF + glossy BTDF) on a
    2-triangle floor, one directional + one disk light."""
    mats = [make_material(kd=(0.05, 0.05, 0.05), ks=(0.9, 0.9, 0.9), kt=(0.9, 0.95, 0.9, 1.0), roughness=(0.1, 0.1),
                          eta=1.5), lambert((0.5, 0.5, 0.5))]
    meshes = [torus_knot_mesh(nu, nv, material=0), floor_mesh("floor", 12.0, 0.0, 1)]
    sc = Scene("dragon_class", meshes, np.concatenate(mats), np.zeros(0, dtype=L.LIGHT),
               CameraSpec(position=(-11.0, 9.5, -11.0), euler_deg=(27.0, 45.0, 0.0)))
    lo, hi = sc.world_bounds()
    sc.lights = np.concatenate([directional_light((72.0, 0.0, 0.0), (40.0, 40.0, 40.0), lo, hi),
                                disk_light((0.0, 14.0, 0.0), (0.0, -1.0, 0.0), 2.0, (20.0, 20.0, 20.0))])
    return sc.finalize()


def traversal_bench_scene(nq=707) -> Scene:
    """T: the C2 knot re-tessellated to ~1 M triangles (707 x 707 quads x 2) + a 2-triangle floor."""
    sc = dragon_class(nq, nq)
    sc.name = "traversal_1M"
    return sc
