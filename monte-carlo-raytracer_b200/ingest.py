"""Headless scene ingestion in the reference's file formats: Wavefront OBJ / MTL and
image files -> the RTHostScene arrays the device path consumes.

What the reference does with an OBJ, restated for a host without assimp / OpenGL:
  * import flags  aiProcessPreset_TargetRealtime_Fast | MakeLeftHanded | FlipWindingOrder
    (engine/resource/AssetImporter.cpp:40): polygons are fan-triangulated (v0, v_i, v_i+1), z is negated,
    triangle winding is reversed, uvs are NOT flipped, missing normals are generated per face (flat);
  * one sub-mesh per (group, material) in file order; every sub-mesh becomes one RTShape with its own
    slice of the shared index / vertex pools and one Radeon Rays mesh whose id is the RTShape index
    (scene/RTScene.cpp:600-670);
  * material:  Kd -> uber_kd, Ks -> uber_ks, roughness = clamp(sqrt(2 / (Ns + 2)), 1e-5, 1), kr = kt = 0,
    opacity 1, eta 1.5 (scene/RTScene.cpp:821-853, kernel_data.h:89-95); texture slots map_Kd -> diffuse,
    map_Bump / bump / norm -> normal map, map_d -> opacity, map_Ks -> glossy (RTScene.cpp:855-881);
  * textures are packed as level-0 RGBA8 rows in file order into one byte buffer with a TextureDesc2D
    table, wrap = REPEAT (RTScene.cpp:680-766);
  * a group whose material has a non-zero Ke becomes a triangle-mesh area light of that intensity.
"""
import os

import numpy as np

from . import layouts as L
from .scenes import CameraSpec, Mesh, Scene, make_material, mesh_area, tri_mesh_light

f4 = np.float32


def parse_mtl(path):
    mats, cur = {}, None
    if not os.path.exists(path):
        return mats
    with open(path, "r", encoding="latin-1") as f:
        for line in f:
            line = line.split("#", 1)[0].split()
            if not line:
                continue
            key, args = line[0], line[1:]
            if key == "newmtl":
                cur = {"Kd": (0.6, 0.6, 0.6), "Ks": (0.0, 0.0, 0.0), "Ke": (0.0, 0.0, 0.0), "Ns": 0.0, "maps": {}}
                mats[" ".join(args)] = cur
            elif cur is None:
                continue
            elif key in ("Kd", "Ks", "Ke"):
                v = [float(a) for a in args[:3]]
                cur[key] = tuple(v + [v[-1]] * (3 - len(v)))
            elif key == "Ns":
                cur["Ns"] = float(args[0])
            elif key.lower() in ("map_kd", "map_bump", "bump", "norm", "map_d", "map_ks"):
                slot = {"map_kd": "diffuse", "map_bump": "normal", "bump": "normal", "norm": "normal", "map_d": "opacity",
                        "map_ks": "glossy"}[key.lower()]
                cur["maps"][slot] = args[-1]          # options such as -bm precede the file name
    return mats


def load_image_rgba8(path):
    """(H, W, 4) uint8, rows in file order.  TGA / PNG / JPEG through Pillow (what soil2 covers in the engine)."""
    from PIL import Image
    with Image.open(path) as im:
        return np.asarray(im.convert("RGBA"), np.uint8).copy()


class TexturePool:
    def __init__(self):
        self.images, self.ids = [], {}

    def get(self, path):
        key = os.path.normpath(path)
        if key not in self.ids:
            self.ids[key] = len(self.images)
            self.images.append(load_image_rgba8(path))
        return self.ids[key]

    def pack(self):
        textures = np.zeros(len(self.images), dtype=L.TEXDESC)
        off = 0
        for i, im in enumerate(self.images):
            textures[i]["width"], textures[i]["height"] = im.shape[1], im.shape[0]
            textures[i]["numMipLevels"], textures[i]["format"], textures[i]["wrap"] = 1, L.TEX_FORMAT_RGBA8, L.TEX_WRAP_REPEAT
            textures[i]["memOffset"] = off
            off += im.size
        data = np.concatenate([im.reshape(-1) for im in self.images]) if self.images else np.zeros(0, np.uint8)
        return textures, data


def _resolve(n, count):
    n = int(n)
    return n - 1 if n > 0 else count + n


def load_obj(path, scale=1.0, camera=None, extra_lights=(), name=None) -> Scene:
    """OBJ + MTL (+ textures next to it) -> finalized Scene."""
    base = os.path.dirname(os.path.abspath(path))
    V, VT, VN = [], [], []
    mtl = {}
    submeshes, order = {}, []           # (group, material) -> list of faces [(vi, ti, ni), ...]
    group, material = "default", None
    with open(path, "r", encoding="latin-1") as f:
        for line in f:
            line = line.split("#", 1)[0].split()
            if not line:
                continue
            key, args = line[0], line[1:]
            if key == "v":
                V.append([float(a) for a in args[:3]])
            elif key == "vt":
                VT.append([float(a) for a in (args + ["0"])[:2]])
            elif key == "vn":
                VN.append([float(a) for a in args[:3]])
            elif key in ("g", "o"):
                group = " ".join(args) if args else "default"
            elif key == "usemtl":
                material = " ".join(args)
            elif key == "mtllib":
                mtl.update(parse_mtl(os.path.join(base, " ".join(args))))
            elif key == "f":
                corners = []
                for a in args:
                    p = (a.split("/") + ["", ""])[:3]
                    corners.append((_resolve(p[0], len(V)), _resolve(p[1], len(VT)) if p[1] else -1,
                                    _resolve(p[2], len(VN)) if p[2] else -1))
                k = (group, material)
                if k not in submeshes:
                    submeshes[k] = []
                    order.append(k)
                for i in range(1, len(corners) - 1):          # fan triangulation
                    submeshes[k].append((corners[0], corners[i], corners[i + 1]))
    V = np.asarray(V, np.float64).reshape(-1, 3)
    VT = np.asarray(VT, np.float64).reshape(-1, 2)
    VN = np.asarray(VN, np.float64).reshape(-1, 3)

    pool = TexturePool()
    mat_ids, mats, meshes, light_specs = {}, [], [], []
    for (g, mname) in order:
        faces = submeshes[(g, mname)]
        m = mtl.get(mname, {"Kd": (0.6, 0.6, 0.6), "Ks": (0, 0, 0), "Ke": (0, 0, 0), "Ns": 0.0, "maps": {}})
        if mname not in mat_ids:
            rough = float(np.clip(np.sqrt(2.0 / (m["Ns"] + 2.0)), 1e-5, 1.0))
            rm = make_material(kd=m["Kd"], ks=m["Ks"], roughness=(rough, rough))
            for slot, field in (("diffuse", "uber_diffuseTexId"), ("normal", "uber_normalMapId"),
                                ("opacity", "uber_opacityTexId"), ("glossy", "uber_glossyTexId")):
                if slot in m["maps"]:
                    rm[field] = pool.get(os.path.join(base, m["maps"][slot]))
            mat_ids[mname] = len(mats)
            mats.append(rm)
        # un-indexed corners -> vertices (deduplicated per sub-mesh), left-handed flip, winding reversed
        vert_index, pos, nrm, uvs, idx = {}, [], [], [], []
        for tri in faces:
            p = [V[c[0]] for c in tri]
            flat = np.cross(p[1] - p[0], p[2] - p[0])
            ln = np.linalg.norm(flat)
            flat = flat / ln if ln > 0 else np.array([0.0, 1.0, 0.0])
            out = []
            for c in tri:
                n = VN[c[2]] if c[2] >= 0 else flat
                t = VT[c[1]] if c[1] >= 0 else np.zeros(2)
                key = (c[0], c[1], c[2]) if c[2] >= 0 else (c[0], c[1], -1, tuple(np.round(flat, 9)))
                if key not in vert_index:
                    vert_index[key] = len(pos)
                    pos.append([V[c[0]][0] * scale, V[c[0]][1] * scale, -V[c[0]][2] * scale])
                    nrm.append([n[0], n[1], -n[2]])
                    uvs.append([t[0], t[1]])
                out.append(vert_index[key])
            idx.append((out[0], out[2], out[1]))
        mesh = Mesh(g, np.asarray(pos, f4), np.asarray(nrm, f4), np.asarray(uvs, f4), np.asarray(idx, np.int32),
                    mat_ids[mname])
        if any(c > 0 for c in m["Ke"]):
            light_specs.append((len(meshes), m["Ke"]))
        meshes.append(mesh)

    lights = [tri_mesh_light(si, ke, mesh_area(meshes[si])) for si, ke in light_specs] + list(extra_lights)
    textures, texdata = pool.pack()
    sc = Scene(name or os.path.splitext(os.path.basename(path))[0], meshes,
               np.concatenate(mats) if mats else np.zeros(0, dtype=L.MATERIAL),
               np.concatenate(lights) if lights else np.zeros(0, dtype=L.LIGHT),
               camera or CameraSpec(position=(0.0, 0.0, -5.0), euler_deg=(0.0, 0.0, 0.0)),
               textures=textures, texdata=texdata)
    return sc.finalize()


def load_entities(entities, camera=None, extra_lights=(), name="entities") -> Scene:
    """A scene of placed assets: `entities` is a list of (obj_path, transform 4x4 row-major or None[, scale]) -- the
    engine's MeshRenderer entities.  The first use of an asset imports it (load_obj's rules); every later use of the
    SAME asset shares its geometry, as RTScene::attachMesh does (RTScene.cpp:570-598): one more RTShape per sub-mesh
    with the template's startIdx / startVertex / numTriangles and material, its own transform, and
    IntersectionApi::CreateInstance instead of CreateMesh -- which makes Radeon Rays (and this library) commit a
    two-level structure.  Materials and textures are pooled over all assets."""
    meshes, mats, lights = [], [], []
    tex_chunks, texdata_chunks, tex_bytes, tex_count = [], [], 0, 0
    first_use = {}
    for ent in entities:
        path, xf = ent[0], ent[1]
        scale = ent[2] if len(ent) > 2 else 1.0
        T = None if xf is None else np.ascontiguousarray(xf, f4).reshape(4, 4)
        key = (os.path.abspath(path), float(scale))
        if key in first_use:
            for t in first_use[key]:
                tm = meshes[t]
                meshes.append(Mesh(tm.name, tm.positions, tm.normals, tm.uvs, tm.indices, tm.material, T, instance_of=t))
            continue
        part = load_obj(path, scale=scale)
        mat_base, tex_base, shape_base = len(mats), tex_count, len(meshes)
        m2 = part.materials.copy()
        for f in ("uber_normalMapId", "uber_diffuseTexId", "uber_glossyTexId", "uber_specReflectionTexId",
                  "uber_transmissionTexId", "uber_opacityTexId", "uber_roughnessTexId", "uber_iorTexId"):
            m2[f] = np.where(m2[f] >= 0, m2[f] + tex_base, m2[f])
        mats.extend(m2[i:i + 1] for i in range(len(m2)))
        if len(part.textures):
            t2 = part.textures.copy()
            t2["memOffset"] += tex_bytes
            tex_chunks.append(t2)
            texdata_chunks.append(part.texdata)
            tex_bytes += part.texdata.size
            tex_count += len(t2)
        first_use[key] = []
        for m in part.meshes:
            first_use[key].append(len(meshes))
            meshes.append(Mesh(m.name, m.positions, m.normals, m.uvs, m.indices,
                               m.material + mat_base if m.material >= 0 else m.material, T))
        for lt in part.lights:
            l2 = lt.copy()
            if int(l2["shapeId"]) >= 0:
                l2["shapeId"] = int(l2["shapeId"]) + shape_base
            lights.append(np.array([l2], dtype=L.LIGHT))
    lights += list(extra_lights)
    sc = Scene(name, meshes, np.concatenate(mats) if mats else np.zeros(0, dtype=L.MATERIAL),
               np.concatenate(lights) if lights else np.zeros(0, dtype=L.LIGHT),
               camera or CameraSpec(position=(0.0, 0.0, -5.0), euler_deg=(0.0, 0.0, 0.0)),
               textures=np.concatenate(tex_chunks) if tex_chunks else np.zeros(0, dtype=L.TEXDESC),
               texdata=np.concatenate(texdata_chunks) if texdata_chunks else np.zeros(0, dtype=np.uint8))
    return sc.finalize()
