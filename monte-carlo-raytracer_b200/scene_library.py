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
        # the file's "bottom face" is `f -12 -11 -10 -9`, which addresses the RIGHT face's vertices again
        # (CornellBox-Original.obj:107): the imported box has that face twice and no bottom
        [(0.70, 0.00, 0.17), (0.70, 0.60, 0.17), (0.53, 0.60, 0.75), (0.53, 0.00, 0.75)],
    ],
    "tallBox": [
        [(-0.53, 1.20, 0.09), (0.04, 1.20, -0.09), (-0.14, 1.20, -0.67), (-0.71, 1.20, -0.49)],
        [(-0.53, 0.00, 0.09), (-0.53, 1.20, 0.09), (-0.71, 1.20, -0.49), (-0.71, 0.00, -0.49)],
        [(-0.71, 0.00, -0.49), (-0.71, 1.20, -0.49), (-0.14, 1.20, -0.67), (-0.14, 0.00, -0.67)],
        [(-0.14, 0.00, -0.67), (-0.14, 1.20, -0.67), (0.04, 1.20, -0.09), (0.04, 0.00, -0.09)],
        [(0.04, 0.00, -0.09), (0.04, 1.20, -0.09), (-0.53, 1.20, 0.09), (-0.53, 0.00, 0.09)],
        # likewise `f -8 -7 -6 -5` (CornellBox-Original.obj:155) repeats the FRONT face
        [(0.04, 0.00, -0.09), (0.04, 1.20, -0.09), (-0.53, 1.20, 0.09), (-0.53, 0.00, 0.09)],
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


# ----------------------------------------------------------------------------- C3: Sponza-class atrium
def lattice_noise(size, seed, octaves=4, base=8):
    """Tileable value noise in [0,1] on a size x size image: per octave a (base*2^o)^2 lattice of
    wang-hashed values, bilinearly interpolated with wrap-around (so REPEAT wrapping is seamless)."""
    out = np.zeros((size, size), np.float64)
    amp, tot = 1.0, 0.0
    ys, xs = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    for o in range(octaves):
        n = base << o
        gy, gx = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
        lat = hash01(gx, gy, seed * 131 + o)
        fx, fy = xs * (n / size), ys * (n / size)
        ix, iy = np.floor(fx).astype(np.int64), np.floor(fy).astype(np.int64)
        tx, ty = fx - ix, fy - iy
        tx, ty = tx * tx * (3 - 2 * tx), ty * ty * (3 - 2 * ty)
        ix1, iy1 = (ix + 1) % n, (iy + 1) % n
        out += amp * ((lat[iy, ix] * (1 - tx) + lat[iy, ix1] * tx) * (1 - ty) +
                      (lat[iy1, ix] * (1 - tx) + lat[iy1, ix1] * tx) * ty)
        tot += amp
        amp *= 0.5
    return out / tot


def make_texture_set(tex_id_base, size, tint, with_opacity=False):
    """Diffuse (RGBA8, alpha 255), specular and normal-map textures of one material, plus an optional
    alpha-mask texture; returns a list of (size,size,4) uint8 arrays (level 0 only: the kernels never
    sample another level, textures.cl:204-209)."""
    h = lattice_noise(size, 3000 + tex_id_base, 5, 4)
    g = lattice_noise(size, 3001 + tex_id_base, 3, 16)
    tint = np.asarray(tint, np.float64)
    diff = np.empty((size, size, 4), np.uint8)
    diff[..., :3] = np.clip((0.35 + 0.65 * h[..., None]) * tint * (0.85 + 0.15 * g[..., None]) * 255.0, 0, 255)
    diff[..., 3] = 255
    spec = np.empty((size, size, 4), np.uint8)
    spec[..., :3] = np.clip((0.2 + 0.8 * g[..., None]) * 255.0, 0, 255)
    spec[..., 3] = 255
    # normal map from the height field by central differences (wrap), encoded 0.5 * n + 0.5
    dx = (np.roll(h, -1, 1) - np.roll(h, 1, 1)) * (size / 64.0)
    dy = (np.roll(h, -1, 0) - np.roll(h, 1, 0)) * (size / 64.0)
    n = np.stack([-dx, -dy, np.ones_like(h)], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    nm = np.empty((size, size, 4), np.uint8)
    nm[..., :3] = np.clip((0.5 * n + 0.5) * 255.0 + 0.5, 0, 255)
    nm[..., 3] = 255
    out = [diff, spec, nm]
    if with_opacity:
        m = lattice_noise(size, 3002 + tex_id_base, 2, 8)
        op = np.empty((size, size, 4), np.uint8)
        op[..., :3] = np.where(m[..., None] > 0.35, 255, 0)
        op[..., 3] = 255
        out.append(op)
    return out


def cylinder_mesh(name, cx, cz, radius, y0, y1, nseg, nring, material, flutes=12) -> Mesh:
    th = np.linspace(0.0, 2 * np.pi, nseg + 1)
    ys = np.linspace(y0, y1, nring + 1)
    T, Y = np.meshgrid(th, ys, indexing="ij")
    r = radius * (1.0 + 0.04 * np.cos(flutes * T)) * (1.0 + 0.08 * np.exp(-((Y - y0) / 0.4) ** 2) +
                                                      0.08 * np.exp(-((y1 - Y) / 0.4) ** 2))
    P = np.stack([cx + r * np.cos(T), Y, cz + r * np.sin(T)], -1)
    N = np.stack([np.cos(T), np.zeros_like(T), np.sin(T)], -1)
    UV = np.stack([T / (2 * np.pi) * 2.0, (Y - y0) / (y1 - y0) * 4.0], -1)
    m = grid_mesh(name, P, N, UV, material)
    m.indices = m.indices[:, [0, 2, 1]].copy()     # geometric normal outward, like the shading normal
    return m


def plane_grid(name, origin, eu, ev, nu, nv, material, uv_scale, normal, height_seed=None, amp=0.0) -> Mesh:
    """Planar nu x nv-quad grid spanned by eu, ev from origin; optional noise displacement along normal."""
    a = np.linspace(0.0, 1.0, nu + 1)
    b = np.linspace(0.0, 1.0, nv + 1)
    A, B = np.meshgrid(a, b, indexing="ij")
    origin, eu, ev, normal = (np.asarray(v, np.float64) for v in (origin, eu, ev, normal))
    P = origin + A[..., None] * eu + B[..., None] * ev
    if height_seed is not None:
        hgt = value_noise(A * 12.0, B * 8.0, height_seed, 4) * amp
        P = P + hgt[..., None] * normal
    N = np.broadcast_to(normal, P.shape).copy()
    UV = np.stack([A * uv_scale[0], B * uv_scale[1]], -1)
    m = grid_mesh(name, P, N, UV, material)
    p, i = m.positions, m.indices
    gn = np.cross(p[i[0, 0]] - p[i[0, 2]], p[i[0, 1]] - p[i[0, 2]])
    if np.dot(gn, normal) < 0:
        m.indices = m.indices[:, [0, 2, 1]].copy()
    return m


def merge_meshes(name, meshes, material) -> Mesh:
    pos, nrm, uvs, idx, base = [], [], [], [], 0
    for m in meshes:
        pos.append(m.positions); nrm.append(m.normals); uvs.append(m.uvs); idx.append(m.indices + base)
        base += m.positions.shape[0]
    return Mesh(name, np.concatenate(pos), np.concatenate(nrm), np.concatenate(uvs), np.concatenate(idx), material)


def sponza_class(tex_size=1024) -> Scene:
    """C3: 262,144 triangles (2^18) in 24 shapes / 24 materials, every material with 1024^2 RGBA8
    diffuse + specular + normal-map textures (2 also alpha-masked), point + directional + triangle-mesh
    area light.  Hall 24 x 8 x 12 with an open central nave, two rows of 8 fluted columns."""
    X0, X1, Z0, Z1, H = -12.0, 12.0, -6.0, 6.0, 8.0
    tints = [(0.80, 0.72, 0.60), (0.70, 0.70, 0.72), (0.75, 0.62, 0.50), (0.62, 0.68, 0.60)]
    meshes = []
    # 0 floor, 1 aisle ceilings (two strips, nave open to the sky), 2-3 long walls, 4-5 end walls, 6 relief
    meshes.append(plane_grid("floor", (X0, 0, Z0), (X1 - X0, 0, 0), (0, 0, Z1 - Z0), 128, 128, 0, (12, 6), (0, 1, 0)))
    c0 = plane_grid("c0", (X0, H, Z0), (X1 - X0, 0, 0), (0, 0, 2.5), 128, 64, 1, (12, 2), (0, -1, 0))
    c1 = plane_grid("c1", (X0, H, Z1 - 2.5), (X1 - X0, 0, 0), (0, 0, 2.5), 128, 64, 1, (12, 2), (0, -1, 0))
    meshes.append(merge_meshes("ceiling", [c0, c1], 1))
    meshes.append(plane_grid("wallN", (X0, 0, Z1), (X1 - X0, 0, 0), (0, H, 0), 128, 64, 2, (12, 4), (0, 0, -1)))
    meshes.append(plane_grid("wallS", (X0, 0, Z0), (X1 - X0, 0, 0), (0, H, 0), 128, 64, 3, (12, 4), (0, 0, 1)))
    meshes.append(plane_grid("wallW", (X0, 0, Z0), (0, 0, Z1 - Z0), (0, H, 0), 128, 64, 4, (6, 4), (1, 0, 0)))
    meshes.append(plane_grid("wallE", (X1, 0, Z0), (0, 0, Z1 - Z0), (0, H, 0), 128, 64, 5, (6, 4), (-1, 0, 0)))
    meshes.append(plane_grid("relief", (X0 + 2, 0.5, Z1 - 0.35), (X1 - X0 - 4, 0, 0), (0, 5.0, 0), 217, 151, 6, (8, 3),
                             (0, 0, -1), height_seed=33, amp=0.3))
    k = 7
    for row, cz in enumerate((-3.5, 3.5)):
        for j in range(8):
            cx = X0 + 1.5 + j * 3.0
            meshes.append(cylinder_mesh("column%d_%d" % (row, j), cx, cz, 0.42, 0.0, H, 64, 32, k))
            k += 1
    # 23: emissive ceiling panel under the north aisle ceiling (2 triangles), facing down
    meshes.append(plane_grid("panel", (-2.0, H - 0.05, Z1 - 2.2), (4.0, 0, 0), (0, 0, 1.6), 1, 1, 23, (1, 1), (0, -1, 0)))
    assert len(meshes) == 24

    mats, tex_images = [], []
    for m in range(24):
        mat = make_material(kd=(1, 1, 1), ks=(0.25, 0.25, 0.25), roughness=(np.sqrt(2.0 / 22.0),) * 2)
        masked = m in (6, 9)
        imgs = make_texture_set(10 * m, tex_size, tints[m % 4], with_opacity=masked)
        base = len(tex_images)
        tex_images += imgs
        mat["uber_diffuseTexId"], mat["uber_glossyTexId"], mat["uber_normalMapId"] = base, base + 1, base + 2
        if masked:
            mat["uber_opacityTexId"] = base + 3
        mats.append(mat)
    textures = np.zeros(len(tex_images), dtype=L.TEXDESC)
    off = 0
    for i, im in enumerate(tex_images):
        textures[i]["width"], textures[i]["height"] = im.shape[1], im.shape[0]
        textures[i]["numMipLevels"], textures[i]["format"], textures[i]["wrap"] = 1, L.TEX_FORMAT_RGBA8, L.TEX_WRAP_REPEAT
        textures[i]["memOffset"] = off
        off += im.size
    texdata = np.concatenate([im.reshape(-1) for im in tex_images])

    sc = Scene("sponza_class", meshes, np.concatenate(mats), np.zeros(0, dtype=L.LIGHT),
               CameraSpec(position=(-10.6, 2.6, 0.4), euler_deg=(-4.0, 84.0, 0.0)), textures=textures, texdata=texdata)
    lo, hi = sc.world_bounds()
    sc.lights = np.concatenate([point_light((0.0, 5.0, 0.0), (30.0, 30.0, 30.0)),
                                directional_light((72.0, 0.0, 0.0), (40.0, 40.0, 40.0), lo, hi),
                                tri_mesh_light(23, (15.0, 15.0, 15.0), mesh_area(meshes[23]))])
    return sc.finalize()
