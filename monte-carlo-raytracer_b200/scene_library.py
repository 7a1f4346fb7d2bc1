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
