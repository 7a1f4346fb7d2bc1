"""Headless scene builders producing the reference's host scene arrays (RTHostScene,
scene/RTScene.h:44-61) for the BASELINE configs.

Every builder is a deterministic function of integer seeds built on the reference's own
wang_hash + xorshift (assets/kernels/rng.cl:48-66) so the arrays can be regenerated
bit-identically anywhere.  The flattening follows RTScene::attachMesh
(scene/RTScene.cpp:564-678): one RTShape per sub-mesh, shared index / vertex pools,
`startIdx` / `startVertex` offsets, indices relative to `startVertex`, shape id == index into
the RTShape array, one Radeon Rays mesh per shape.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import layouts as L

f4 = np.float32


# ----------------------------------------------------------------------------- hashing
def wang_hash(seed):
    """rng.cl:58-66, vectorised on uint32."""
    s = np.asarray(seed, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        s = (s ^ np.uint32(61)) ^ (s >> np.uint32(16))
        s = s * np.uint32(9)
        s = s ^ (s >> np.uint32(4))
        s = s * np.uint32(0x27D4EB2D)
        s = s ^ (s >> np.uint32(15))
    return s


def xorshift(state):
    """rng.cl:48-56: returns (new_state); new_state is also the random number."""
    s = np.asarray(state, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        s = s + np.uint32(2463534242)
        s = s ^ (s << np.uint32(13))
        s = s ^ (s >> np.uint32(17))
        s = s ^ (s << np.uint32(5))
    return s


def rand_float(state):
    """rng.cl:104-107: (float)u / 0xffffffff with the divisor rounded to 2^32 in fp32."""
    s = xorshift(state)
    return s, (s.astype(np.float32) / np.float32(4294967296.0)).astype(np.float32)


def hash01(ix, iy, seed):
    """Uniform [0,1) from two integer lattice coordinates and a seed."""
    h = wang_hash(np.asarray(ix, np.uint32) * np.uint32(73856093) ^ np.asarray(iy, np.uint32) * np.uint32(19349663)
                  ^ np.uint32((int(seed) * 83492791) & 0xFFFFFFFF))
    return (h >> np.uint32(8)).astype(np.float64) / float(1 << 24)


def value_noise(x, y, seed, octaves=1):
    """Bilinear value noise on the integer lattice, summed over octaves (persistence 0.5)."""
    x = np.asarray(x, np.float64)
    y = np.asarray(y, np.float64)
    out = np.zeros(np.broadcast(x, y).shape)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        fx, fy = x * (1 << o), y * (1 << o)
        ix, iy = np.floor(fx).astype(np.int64), np.floor(fy).astype(np.int64)
        tx, ty = fx - ix, fy - iy
        tx, ty = tx * tx * (3 - 2 * tx), ty * ty * (3 - 2 * ty)
        s = seed + 7919 * o
        v00 = hash01(ix & 0xFFFF, iy & 0xFFFF, s)
        v10 = hash01((ix + 1) & 0xFFFF, iy & 0xFFFF, s)
        v01 = hash01(ix & 0xFFFF, (iy + 1) & 0xFFFF, s)
        v11 = hash01((ix + 1) & 0xFFFF, (iy + 1) & 0xFFFF, s)
        out += amp * ((v00 * (1 - tx) + v10 * tx) * (1 - ty) + (v01 * (1 - tx) + v11 * tx) * ty)
        tot += amp
        amp *= 0.5
    return out / tot


# ----------------------------------------------------------------------------- containers
@dataclass
class Mesh:
    """One sub-mesh = one RTShape = one Radeon Rays mesh (RTScene.cpp:600-670)."""
    name: str
    positions: np.ndarray            # (V,3) float32, object space
    normals: np.ndarray              # (V,3) float32
    uvs: np.ndarray                  # (V,2) float32
    indices: np.ndarray              # (T,3) int32, relative to this mesh's first vertex
    material: int = L.RT_INVALID_ID
    transform: Optional[np.ndarray] = None   # 4x4 row-major float32; None = identity
    # Second and later uses of a mesh (RTScene::attachMesh, RTScene.cpp:575-598): index of the Mesh in Scene.meshes
    # whose geometry this shape shares.  Its own RTShape has the template's startIdx / startVertex / numTriangles and
    # its own transform; Radeon Rays gets CreateInstance(template) instead of a second CreateMesh.  The arrays above
    # then ARE the template's (same objects), so code that flattens shapes keeps working.
    instance_of: Optional[int] = None


@dataclass
class CameraSpec:
    position: tuple
    euler_deg: tuple
    fov_y: float = 45.0              # passed to glm::perspective as-is, i.e. RADIANS
    near: float = 0.3
    far: float = 30.0


@dataclass
class Scene:
    name: str
    meshes: List[Mesh]
    materials: np.ndarray
    lights: np.ndarray
    camera: CameraSpec
    textures: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=L.TEXDESC))
    texdata: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.uint8))
    # flattened pools (filled by finalize)
    shapes: np.ndarray = None
    indices: np.ndarray = None
    positions: np.ndarray = None
    uvs: np.ndarray = None
    normals: np.ndarray = None

    @property
    def num_triangles(self) -> int:
        return int(sum(m.indices.shape[0] for m in self.meshes))

    def finalize(self) -> "Scene":
        """RTScene::attachMesh flattening + light <-> shape links (RTScene.cpp:478-553,564-678)."""
        nshape = len(self.meshes)
        shapes = L.default_shape(nshape)
        idx_chunks, pos_chunks, uv_chunks, nrm_chunks = [], [], [], []
        start_idx = start_vtx = 0
        for i, m in enumerate(self.meshes):
            T = np.eye(4, dtype=f4) if m.transform is None else np.asarray(m.transform, f4)
            shapes[i]["toWorldTransform"] = T
            shapes[i]["toWorldInverseTranspose"] = np.linalg.inv(T.astype(np.float64)).T.astype(f4)
            shapes[i]["materialId"] = m.material
            if m.instance_of is not None:       # shares the template's index / vertex ranges (RTScene.cpp:583-590)
                t = int(m.instance_of)
                assert t < i and self.meshes[t].instance_of is None, "an instance must follow its template mesh"
                for k in ("startIdx", "startVertex", "numTriangles"):
                    shapes[i][k] = shapes[t][k]
                continue
            shapes[i]["startIdx"] = start_idx
            shapes[i]["startVertex"] = start_vtx
            shapes[i]["numTriangles"] = m.indices.shape[0]
            idx_chunks.append(np.ascontiguousarray(m.indices, np.uint32).reshape(-1))
            pos_chunks.append(m.positions.astype(f4))
            uv_chunks.append(m.uvs.astype(f4))
            nrm_chunks.append(m.normals.astype(f4))
            start_idx += 3 * m.indices.shape[0]
            start_vtx += m.positions.shape[0]
        for li in range(len(self.lights)):
            sid = int(self.lights[li]["shapeId"])
            if self.lights[li]["type"] in (L.RT_DISK_AREA_LIGHT, L.RT_TRIANGLE_MESH_AREA_LIGHT) and sid >= 0:
                shapes[sid]["lightID"] = li
        n = len(self.lights)
        if n:
            self.lights["choicePdf"] = f4(1.0) / f4(n)          # RTScene.cpp:811-819
        self.shapes = shapes
        self.indices = np.concatenate(idx_chunks)
        pos = np.concatenate(pos_chunks)
        nrm = np.concatenate(nrm_chunks)
        self.positions = np.zeros((pos.shape[0], 4), f4)
        self.positions[:, :3] = pos
        self.normals = np.zeros((nrm.shape[0], 4), f4)
        self.normals[:, :3] = nrm
        self.uvs = np.ascontiguousarray(np.concatenate(uv_chunks), f4)
        return self

    def world_bounds(self):
        lo = np.full(3, np.inf)
        hi = np.full(3, -np.inf)
        for m in self.meshes:
            p = m.positions.astype(np.float64)
            if m.transform is not None:
                T = np.asarray(m.transform, np.float64)
                p = p @ T[:3, :3].T + T[:3, 3]
            lo = np.minimum(lo, p.min(0))
            hi = np.maximum(hi, p.max(0))
        return lo, hi


def mesh_area(m: Mesh) -> float:
    """MeshRenderer::computeSurfaceArea: sum of world-space triangle areas."""
    p = m.positions.astype(np.float64)
    if m.transform is not None:
        T = np.asarray(m.transform, np.float64)
        p = p @ T[:3, :3].T + T[:3, 3]
    a, b, c = p[m.indices[:, 0]], p[m.indices[:, 1]], p[m.indices[:, 2]]
    return float(0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).sum())


def lambert(kd):
    m = L.default_material(1)
    m["uber_kd"][0, :3] = kd
    m["uber_ks"][0, :3] = 0.0
    return m


def make_material(kd=(0, 0, 0), ks=(0, 0, 0), kr=(0, 0, 0), kt=(0, 0, 0, 0), roughness=(0.1, 0.1), eta=1.5,
                  opacity=(1, 1, 1)):
    m = L.default_material(1)
    m["uber_kd"][0, :3] = kd
    m["uber_ks"][0, :3] = ks
    m["uber_kr"][0, :3] = kr
    m["uber_kt"][0, :4] = kt
    m["uber_roughness"][0] = roughness
    m["uber_eta"] = eta
    m["uber_opacity"][0, :3] = opacity
    return m


def tri_mesh_light(shape_id: int, intensity, area: float, position=(0, 0, 0)):
    l = np.zeros(1, dtype=L.LIGHT)
    l["type"] = L.RT_TRIANGLE_MESH_AREA_LIGHT
    l["flags"] = L.RT_LIGHT_FLAG_AREA
    l["p"][0, :3] = position
    l["shapeId"] = shape_id
    l["area"] = area
    l["intensity"][0, :3] = intensity
    return l


def euler_forward(euler_deg):
    from .camera import euler_yxz_matrix
    return euler_yxz_matrix(euler_deg)[:, 2]


def directional_light(euler_deg, intensity, scene_lo, scene_hi):
    """RTScene::setLight RT_DIRECTIONAL_LIGHT (RTScene.cpp:484-497)."""
    l = np.zeros(1, dtype=L.LIGHT)
    fwd = euler_forward(euler_deg)
    radius = 0.5 * float(np.linalg.norm(scene_hi - scene_lo))
    l["type"] = L.RT_DIRECTIONAL_LIGHT
    l["flags"] = L.RT_LIGHT_FLAG_DELTA_DIRECTION
    l["d"][0, :3] = fwd
    l["intensity"][0, :3] = intensity
    l["radius"] = radius
    l["p"][0, :3] = 0.5 * (scene_lo + scene_hi) - fwd * radius
    l["area"] = np.pi * radius * radius
    l["shapeId"] = -1
    return l


def point_light(position, intensity):
    l = np.zeros(1, dtype=L.LIGHT)
    l["type"] = L.RT_POINT_LIGHT
    l["flags"] = L.RT_LIGHT_FLAG_DELTA_POSITION
    l["p"][0, :3] = position
    l["intensity"][0, :3] = intensity
    l["shapeId"] = -1
    return l


def disk_light(position, normal, radius, intensity):
    l = np.zeros(1, dtype=L.LIGHT)
    n = np.asarray(normal, np.float64)
    l["type"] = L.RT_DISK_AREA_LIGHT
    l["flags"] = L.RT_LIGHT_FLAG_AREA
    l["p"][0, :3] = position
    l["d"][0, :3] = n / np.linalg.norm(n)
    l["radius"] = radius
    l["area"] = np.pi * radius * radius
    l["intensity"][0, :3] = intensity
    l["shapeId"] = -1
    return l
