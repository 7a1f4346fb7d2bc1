"""oracle/ref.py -- TEST INFRASTRUCTURE.  Python driver of oracle/_ref/libmcrt_ref.so.

libmcrt_ref.so is the reference's OWN source (assets/kernels/*.cl, Radeon Rays' Bvh2 builder and
intersect_bvh2_lds.cl) compiled for the host by oracle/ref_build/Makefile.  This module restates
only the reference's host loops around those kernels:

  * RTPrimaryRaysPass::update            (renderPasses/RTPrimaryRaysPass.cpp:32-66)
  * RTPathTracingPass::update            (renderPasses/RTPathTracingPass.cpp:40-114)
  * IntersectorLDS::Process / Intersect  (RadeonRays/src/intersector/intersector_lds.cpp:163-326)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It never runs on the product path.
"""
import ctypes
import os

import numpy as np

import mcrt_b200  # noqa: F401  (registers the package)
from mcrt_b200 import layouts as L

_HERE = os.path.dirname(os.path.abspath(__file__))


class RefSceneDesc(ctypes.Structure):
    _fields_ = [("shapes", ctypes.c_void_p), ("indices", ctypes.c_void_p), ("positions", ctypes.c_void_p),
                ("uvs", ctypes.c_void_p), ("normals", ctypes.c_void_p), ("textures", ctypes.c_void_p),
                ("texdata", ctypes.c_void_p), ("sobol", ctypes.c_void_p), ("lights", ctypes.c_void_p),
                ("materials", ctypes.c_void_p), ("camera", ctypes.c_void_p),
                ("num_shapes", ctypes.c_int32), ("num_textures", ctypes.c_int32), ("num_lights", ctypes.c_int32),
                ("num_materials", ctypes.c_int32), ("num_indices", ctypes.c_int64), ("num_vertices", ctypes.c_int64),
                ("texdata_bytes", ctypes.c_int64)]


def lib_path(fast: bool = False, sobol: bool = False) -> str:
    name = "libmcrt_ref_sobol.so" if sobol else ("libmcrt_ref_fast.so" if fast else "libmcrt_ref.so")
    return os.path.join(_HERE, "_ref", name)


def available(fast: bool = False, sobol: bool = False) -> bool:
    return os.path.exists(lib_path(fast, sobol))


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(ctypes.c_void_p)


class RefOracle:
    """The compiled reference, one NDRange per call."""

    def __init__(self, fast: bool = False, sobol_table=None):
        """sobol_table: uint32 generator matrices -> the Sobol-sampler build of the kernels is used."""
        self.sobol_table = None if sobol_table is None else np.ascontiguousarray(sobol_table, np.uint32)
        self.lib = ctypes.CDLL(lib_path(fast, self.sobol_table is not None))
        lib = self.lib
        lib.ref_set_threads.argtypes = [ctypes.c_int]
        lib.ref_get_threads.restype = ctypes.c_int
        lib.ref_rr_world_create.restype = ctypes.c_void_p
        lib.ref_rr_world_destroy.argtypes = [ctypes.c_void_p]
        lib.ref_rr_world_add_mesh.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_void_p]
        lib.ref_rr_world_add_instance.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                  ctypes.c_void_p]
        lib.ref_rr_world_build.argtypes = [ctypes.c_void_p, ctypes.c_float, ctypes.c_int, ctypes.c_int]
        lib.ref_rr_world_build.restype = ctypes.c_longlong
        lib.ref_rr_world_nodes.argtypes = [ctypes.c_void_p]
        lib.ref_rr_world_nodes.restype = ctypes.c_void_p
        lib.ref_rr_intersect.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
        lib.ref_rr_occluded.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
        lib.ref_pt_generate_rays.argtypes = [ctypes.c_void_p] * 3
        lib.ref_pt_shade.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                     ctypes.c_void_p]
        lib.ref_pt_shadow_pass.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]

    def set_threads(self, n=None):
        """Run the kernels' OpenMP loops on `n` threads (default: every core this process may use), whatever
        OMP_NUM_THREADS says; returns the team size a parallel region then really gets."""
        if n is None:
            n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.lib.ref_set_threads(int(n))
        return int(self.lib.ref_get_threads())

    # ------------------------------------------------------------------ Radeon Rays side
    def build_bvh(self, meshes, use_sah=True, traversal_cost=10.0, num_bins=64):
        """meshes: iterable of (positions(V,3) f32, indices(T,3) i32, shape_id, transform or None).
        Returns the Bvh2 node array as raw bytes (n,64) -- engine defaults: FatBVH + SAH, cost 10,
        64 bins (GUI/PathTracingSettings.h:241-249)."""
        w = self.lib.ref_rr_world_create()
        keep = []
        try:
            for pos, idx, sid, xf in meshes:
                pos = np.ascontiguousarray(pos, np.float32)
                idx = np.ascontiguousarray(idx, np.int32)
                m = None if xf is None else np.ascontiguousarray(xf, np.float32)
                minv = None if xf is None else np.ascontiguousarray(np.linalg.inv(m.astype(np.float64)), np.float32)
                keep += [pos, idx, m, minv]
                self.lib.ref_rr_world_add_mesh(w, _ptr(pos), pos.shape[0], 12, _ptr(idx), 12, idx.shape[0], int(sid),
                                               _ptr(m), _ptr(minv))
            n = self.lib.ref_rr_world_build(w, traversal_cost, num_bins, 1 if use_sah else 0)
            src = self.lib.ref_rr_world_nodes(w)
            nodes = np.frombuffer(ctypes.string_at(src, n * 64), dtype=np.uint8).reshape(n, 64).copy()
        finally:
            self.lib.ref_rr_world_destroy(w)
        return nodes

    def build_scene_bvh(self, scene, **kw):
        return self.build_bvh([(m.positions, m.indices, i, m.transform) for i, m in enumerate(scene.meshes)], **kw)

    def intersect(self, nodes, rays, hits=None):
        """intersect_main; `hits` (prefilled) is updated in place like the device buffer would be."""
        if hits is None:
            hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
            hits["shapeid"] = -1
            hits["primid"] = -1
        self.lib.ref_rr_intersect(_ptr(nodes), _ptr(rays), rays.shape[0], _ptr(hits))
        return hits

    def occluded(self, nodes, rays, out=None):
        if out is None:
            out = np.zeros(rays.shape[0], dtype=np.int32)
        self.lib.ref_rr_occluded(_ptr(nodes), _ptr(rays), rays.shape[0], _ptr(out))
        return out

    # ------------------------------------------------------------------ path tracer side
    def scene_desc(self, scene, camera):
        d = RefSceneDesc()
        d.shapes, d.indices, d.positions = _ptr(scene.shapes), _ptr(scene.indices), _ptr(scene.positions)
        d.uvs, d.normals = _ptr(scene.uvs), _ptr(scene.normals)
        d.textures, d.texdata, d.sobol = _ptr(scene.textures), _ptr(scene.texdata), _ptr(self.sobol_table)
        d.lights, d.materials, d.camera = _ptr(scene.lights), _ptr(scene.materials), _ptr(camera)
        d.num_shapes, d.num_textures = len(scene.shapes), len(scene.textures)
        d.num_lights, d.num_materials = len(scene.lights), len(scene.materials)
        d.num_indices, d.num_vertices = scene.indices.size, scene.positions.shape[0]
        d.texdata_bytes = scene.texdata.size
        return d

    def generate_rays(self, camera):
        w, h = int(camera["width"][0]), int(camera["height"][0])
        rays = np.zeros(w * h, dtype=L.RAY)
        diffs = np.zeros(w * h, dtype=L.RAYDIFF)
        self.lib.ref_pt_generate_rays(_ptr(rays), _ptr(diffs), _ptr(camera))
        return rays


class RefPathTracer:
    """RTPrimaryRaysPass + RTPathTracingPass state for one image size: the buffers live across
    frames exactly like the reference's device buffers (zero-filled at creation)."""

    def __init__(self, oracle: RefOracle, scene, nodes, width, height):
        self.o, self.scene, self.nodes, self.w, self.h = oracle, scene, nodes, width, height
        n = width * height
        self.rays = np.zeros(n, dtype=L.RAY)
        self.diffs = np.zeros(n, dtype=L.RAYDIFF)
        self.isect = np.zeros(n, dtype=L.INTERSECTION)
        self.shadow_rays = np.zeros(n, dtype=L.RAY)
        self.occlusion = np.zeros(n, dtype=np.int32)
        self.temp_radiance = np.zeros((n, 4), np.float32)
        self.radiance = np.zeros((n, 4), np.float32)
        self.throughput = np.zeros(n, dtype=L.THROUGHPUT)
        self.rays_traced = 0
        self.shadow_rays_traced = 0

    def render_sample(self, camera, frame, max_depth, trace=None, pixels=None):
        """One frame = one sample per pixel.  `trace`, if a dict, receives copies of the primary
        hits, of every bounce's temp radiance (for per-stage parity checks) and the per-bounce path record
        ("path_trace", the layout of the CUDA side's "PathTraceBuffer").  `pixels`: global indices of the
        only pixels that carry a path (tile partition); the others stay black."""
        lib, o = self.o.lib, self.o
        desc = o.scene_desc(self.scene, camera)
        lib.ref_pt_generate_rays(_ptr(self.rays), _ptr(self.diffs), _ptr(camera))
        if pixels is not None:
            # only these pixels carry a path (a rank's tile partition): every other ray is switched off, which is
            # what the kernels' `extra.y != 0` tests key on; indices and W*H stay global (E14)
            keep = np.zeros(self.w * self.h, bool)
            keep[np.asarray(pixels, np.int64)] = True
            self.rays["extra"][~keep, 1] = 0
            self.isect["shapeid"][~keep] = -1
        o.intersect(self.nodes, self.rays, self.isect)
        self.rays_traced += int((self.rays["extra"][:, 1] != 0).sum())
        if trace is not None:
            trace["primary_rays"] = self.rays.copy()
            trace["primary_hits"] = self.isect.copy()
            trace["bounce_radiance"] = []
        if trace is not None:
            trace["path_trace"] = np.zeros((max_depth, self.w * self.h, 4), np.int32)
        for b in range(max_depth):
            if trace is not None:
                if trace.get("keep_rays"):      # the ray whose hit bounce b shades (RTRay records, before the kernel rewrites them)
                    trace.setdefault("closest_rays", []).append(self.rays.copy())
                # the record ptb_set_path_trace keeps on the CUDA side (include/ptb.h), from the reference's buffers
                rec = trace["path_trace"][b]
                shaded = (self.rays["extra"][:, 1] != 0) & (self.isect["shapeid"] != -1) & (self.isect["primid"] != -1)
                rec[:, 0] = np.where(shaded, self.isect["shapeid"], 0)
                rec[:, 1] = np.where(shaded, self.isect["primid"], 0)
            lib.ref_pt_shade(ctypes.byref(desc), self.w, self.h, _ptr(self.shadow_rays), _ptr(self.rays),
                             _ptr(self.isect), frame, max_depth, b, _ptr(self.temp_radiance), _ptr(self.throughput))
            if trace is not None:
                emitted = shaded & (self.throughput["ignoreOcclusion"] != 0)
                shadow = shaded & ~emitted & (self.temp_radiance[:, :3] != 0).any(axis=1)
                spawned = shaded & ~emitted & (self.rays["extra"][:, 1] != 0) & (b + 1 < max_depth)
                rec[:, 2] = (shaded * 1 + shadow * 2 + emitted * 4 + spawned * 8 +
                             np.where(shaded, self.throughput["prevBsdfFlags"] << 8, 0))
            o.occluded(self.nodes, self.shadow_rays, self.occlusion)
            self.shadow_rays_traced += int((self.shadow_rays["extra"][:, 1] != 0).sum())
            if trace is not None and trace.get("keep_rays"):
                trace.setdefault("shadow_rays", []).append(self.shadow_rays.copy())
            if trace is not None:
                rec[:, 3] = np.where(shadow, (self.occlusion != -1) | (self.shadow_rays["extra"][:, 1] == 0), 0)
            lib.ref_pt_shadow_pass(_ptr(self.shadow_rays), _ptr(self.occlusion), self.w, self.h, b,
                                   _ptr(self.throughput), _ptr(self.temp_radiance), _ptr(self.radiance))
            if trace is not None:
                trace["bounce_radiance"].append(self.radiance.copy())
            if b + 1 < max_depth:
                o.intersect(self.nodes, self.rays, self.isect)
                self.rays_traced += int((self.rays["extra"][:, 1] != 0).sum())
        return self.radiance


class RefBDPT:
    """RTBDPTPass state and host loop (renderPasses/RTBDPTPass.cpp:67-128,130-448,451-478): buffers are
    zero-filled at creation and live across frames (the kernels read some of them before writing,
    SURVEY App. A E12), so frames must be rendered in order from frame 0."""

    def __init__(self, oracle: RefOracle, scene, nodes, width, height, max_depth):
        self.o, self.scene, self.nodes, self.w, self.h, self.D = oracle, scene, nodes, width, height, max_depth
        lib = oracle.lib
        lib.ref_bdpt_vertex_size.restype = ctypes.c_int
        vp, i, fp, ip = ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p
        lib.ref_bdpt_generate_start_vertices.argtypes = [vp, i, i, i, i] + [vp] * 11
        lib.ref_bdpt_generate_secondary_vertices.argtypes = [vp, i, i, i, i, i, vp, vp, vp, vp, fp, ip, i]
        lib.ref_bdpt_prepare_connections.argtypes = [vp, i, i, i, i] + [vp] * 8
        lib.ref_bdpt_connect_vertices.argtypes = [vp, i, i, i, i] + [vp] * 10
        lib.ref_bdpt_copy_buffer.argtypes = [i, i, vp, vp]
        self.vsize = lib.ref_bdpt_vertex_size()
        n, D = width * height, max_depth
        self.max_conn = (D + 2) * (D + 3) // 2 - 2              # getMaxPossibleConnectionsCount (:397-401)
        z = lambda count, dt: np.zeros(count, dtype=dt)
        self.radiance = np.zeros((n, 4), np.float32)
        self.final = z(n * 3, np.float32)
        self.cam_vertices = z(n * (D + 2) * self.vsize, np.uint8)
        self.light_vertices = z(n * (D + 1) * self.vsize, np.uint8)
        self.cam_rays, self.light_rays = z(n, L.RAY), z(n, L.RAY)
        self.cam_isect, self.light_isect = z(n, L.INTERSECTION), z(n, L.INTERSECTION)
        self.cam_beta, self.light_beta = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32)
        self.cam_pdf, self.light_pdf = z(n, np.float32), z(n, np.float32)
        self.cam_counts, self.light_counts = z(n, np.int32), z(n, np.int32)
        self.conn_rays = z(n * self.max_conn, L.RAY)
        self.conn_vis = z(n * self.max_conn, np.int32)
        self.sampled_cam = z(n * D * self.vsize, np.uint8)
        self.sampled_light = z(n * D * self.vsize, np.uint8)
        self.temp_radiance = np.zeros((n * self.max_conn, 4), np.float32)
        self.rays_traced = 0
        self.connection_rays_traced = 0

    def _trace(self, rays, isect):
        self.o.intersect(self.nodes, rays, isect)
        self.rays_traced += int((rays["extra"][:, 1] != 0).sum())

    def render_sample(self, camera, frame, trace=None):
        lib, o, D, w, h = self.o.lib, self.o, self.D, self.w, self.h
        desc = o.scene_desc(self.scene, camera)
        d = ctypes.byref(desc)
        # generateStartVertices (:130-221)
        lib.ref_bdpt_generate_start_vertices(d, w, h, frame, D, _ptr(self.cam_vertices), _ptr(self.cam_rays),
                                             _ptr(self.cam_beta), _ptr(self.cam_pdf), _ptr(self.cam_counts),
                                             _ptr(self.light_vertices), _ptr(self.light_rays), _ptr(self.light_beta),
                                             _ptr(self.light_pdf), _ptr(self.light_counts), _ptr(self.final))
        self._trace(self.cam_rays, self.cam_isect)
        self._trace(self.light_rays, self.light_isect)
        # generateSecondaryVertices (:223-307)
        for depth in range(1, D + 2):
            lib.ref_bdpt_generate_secondary_vertices(d, w, h, frame, D, 1, _ptr(self.cam_vertices), _ptr(self.cam_rays),
                                                     _ptr(self.cam_isect), _ptr(self.cam_beta), _ptr(self.cam_pdf),
                                                     _ptr(self.cam_counts), depth)
            if depth <= D:
                lib.ref_bdpt_generate_secondary_vertices(d, w, h, frame, D, 0, _ptr(self.light_vertices),
                                                         _ptr(self.light_rays), _ptr(self.light_isect),
                                                         _ptr(self.light_beta), _ptr(self.light_pdf),
                                                         _ptr(self.light_counts), depth)
            self._trace(self.cam_rays, self.cam_isect)
            if depth <= D:
                self._trace(self.light_rays, self.light_isect)
        # prepareVertexConnections (:309-360)
        lib.ref_bdpt_prepare_connections(d, w, h, frame, D, _ptr(self.cam_vertices), _ptr(self.light_vertices),
                                         _ptr(self.sampled_cam), _ptr(self.sampled_light), _ptr(self.conn_rays),
                                         _ptr(self.cam_counts), _ptr(self.light_counts), _ptr(self.temp_radiance))
        o.occluded(self.nodes, self.conn_rays, self.conn_vis)
        self.connection_rays_traced += int((self.conn_rays["extra"][:, 1] != 0).sum())
        if trace is not None:
            trace["cam_counts"] = self.cam_counts.copy()
            trace["light_counts"] = self.light_counts.copy()
            trace["unweighted"] = self.temp_radiance.copy()
            trace["conn_active"] = (self.conn_rays["extra"][:, 1] != 0).copy()
            trace["conn_vis"] = self.conn_vis.copy()
        # makeConnections (:362-395), copyRadianceBuffer (:403-430)
        lib.ref_bdpt_connect_vertices(d, w, h, frame, D, _ptr(self.cam_vertices), _ptr(self.light_vertices),
                                      _ptr(self.sampled_cam), _ptr(self.sampled_light), _ptr(self.conn_rays),
                                      _ptr(self.conn_vis), _ptr(self.cam_counts), _ptr(self.light_counts),
                                      _ptr(self.temp_radiance), _ptr(self.final))
        if trace is not None:      # ConnectVertices multiplied by the visibility and filled in the s = 0 slots (:722-737)
            trace["connected"] = self.temp_radiance.copy()
        lib.ref_bdpt_copy_buffer(w, h, _ptr(self.final), _ptr(self.radiance))
        return self.radiance
