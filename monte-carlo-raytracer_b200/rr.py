"""Host-side mirror of RadeonRays::IntersectionApi over the C ABI (include/rrb.h).

Same method names, argument meaning and error behaviour as
third_party/RadeonRays/RadeonRays/include/radeon_rays.h:170-305 so that the parity tests read
like the reference's own (UnitTest/radeon_rays_apitest_cl.h).  Errors surface as
`RadeonRaysError`, the counterpart of RadeonRays::Exception (radeon_rays.h:132-137).

ctypes is plumbing only: every query runs in the CUDA library; there is no CPU fallback and the
import fails loudly when libmcrt_b200.so is missing.
"""
import ctypes
import os

import numpy as np

from . import layouts as L

NOALLOC_MAX_RAYS = 1 << 20
K_NULL_ID = -1
MAP_READ, MAP_WRITE = 1, 2

_lib = None


class RadeonRaysError(RuntimeError):
    pass


class RrbStats(ctypes.Structure):
    _fields_ = [("num_triangles", ctypes.c_int64), ("num_nodes", ctypes.c_int64), ("node_bytes", ctypes.c_int64),
                ("triangle_bytes", ctypes.c_int64), ("build_seconds", ctypes.c_double), ("launches", ctypes.c_int64),
                ("refits", ctypes.c_int64), ("last_commit_refit", ctypes.c_int64),
                ("last_build_on_device", ctypes.c_int64), ("num_levels", ctypes.c_int64),
                ("two_level", ctypes.c_int64), ("num_instances", ctypes.c_int64), ("two_level_rebuilds", ctypes.c_int64),
                ("top_level_updates", ctypes.c_int64)]


def lib_path() -> str:
    """libmcrt_b200.so next to this file.  MCRT_LIB_VARIANT=<name> selects build/variants/libmcrt_<name>.so, an
    experiment build of the same sources with other -D knobs (tools/build_variants.py); never a fallback."""
    here = os.path.dirname(os.path.abspath(__file__))
    variant = os.environ.get("MCRT_LIB_VARIANT")
    if variant:
        return os.path.join(here, "build", "variants", "libmcrt_%s.so" % variant)
    return os.path.join(here, "libmcrt_b200.so")


def load_library():
    """Load libmcrt_b200.so (built in-tree by build.py / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RadeonRaysError("%s is missing: run `python __graft_entry__.py` (build) first; there is no fallback" % path)
    lib = ctypes.CDLL(path)
    vp, i, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    pp = ctypes.POINTER(ctypes.c_void_p)
    sig = {
        "rrb_create": [i, vp, pp], "rrb_destroy": [vp],
        "rrb_create_mesh": [vp, vp, i, i, vp, i, vp, i, pp], "rrb_create_instance": [vp, vp, pp],
        "rrb_delete_shape": [vp, vp], "rrb_attach": [vp, vp], "rrb_detach": [vp, vp], "rrb_detach_all": [vp],
        "rrb_commit": [vp], "rrb_reset_id_counter": [vp], "rrb_is_world_empty": [vp, ctypes.POINTER(i)],
        "rrb_shape_set_transform": [vp, vp, vp], "rrb_shape_get_transform": [vp, vp, vp],
        "rrb_shape_set_id": [vp, i], "rrb_shape_get_id": [vp, ctypes.POINTER(i)],
        "rrb_shape_set_linear_velocity": [vp, vp], "rrb_shape_set_angular_velocity": [vp, vp],
        "rrb_set_option_str": [vp, ctypes.c_char_p, ctypes.c_char_p],
        "rrb_set_option_f32": [vp, ctypes.c_char_p, ctypes.c_float],
        "rrb_create_buffer": [vp, sz, vp, pp], "rrb_wrap_device_ptr": [vp, vp, sz, pp],
        "rrb_delete_buffer": [vp, vp], "rrb_buffer_device_ptr": [vp, pp, ctypes.POINTER(sz)],
        "rrb_map": [vp, vp, i, sz, sz, pp, pp], "rrb_unmap": [vp, vp, vp, pp],
        "rrb_event_complete": [vp, ctypes.POINTER(i)], "rrb_event_wait": [vp], "rrb_event_delete": [vp, vp],
        "rrb_query_intersection": [vp, vp, i, vp, vp, pp], "rrb_query_occlusion": [vp, vp, i, vp, vp, pp],
        "rrb_query_intersection_indirect": [vp, vp, vp, i, vp, vp, pp],
        "rrb_query_occlusion_indirect": [vp, vp, vp, i, vp, vp, pp],
        "rrb_query_intersection_host": [vp, vp, i, vp], "rrb_query_occlusion_host": [vp, vp, i, vp],
        "rrb_query_intersection_noalloc": [vp, vp, i, vp],
        "rrb_get_stats": [vp, ctypes.POINTER(RrbStats)], "rrb_query_intersection_counted": [vp, vp, i, vp, vp],
        "rrb_get_bvh": [vp, pp, pp, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)],
        "rrb_get_stream": [vp, ctypes.POINTER(i), pp],
        "rrb_host_bvh_build": [vp, i, vp, i, vp, i, pp],
        "rrb_host_bvh_get": [vp, pp, ctypes.POINTER(ctypes.c_int64), pp, ctypes.POINTER(ctypes.c_int64)],
        "rrb_host_bvh_get_refit_info": [vp, pp, pp, ctypes.POINTER(i)],
        "rrb_host_bvh_free": [vp],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = ctypes.c_int
    lib.rrb_last_error.restype = ctypes.c_char_p
    lib.rrb_device_count.restype = ctypes.c_int
    lib.rrb_device_info.argtypes = [i, ctypes.c_char_p, sz, ctypes.c_char_p, sz, ctypes.POINTER(i)]
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise RadeonRaysError(load_library().rrb_last_error().decode())


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Shape:
    """RadeonRays::Shape (radeon_rays.h:91-110)."""

    def __init__(self, handle):
        self._h = handle

    def SetTransform(self, m, minv=None):
        m = np.ascontiguousarray(m, np.float32).reshape(16)
        if minv is None:
            minv = np.linalg.inv(m.reshape(4, 4).astype(np.float64)).astype(np.float32)
        minv = np.ascontiguousarray(minv, np.float32).reshape(16)
        _check(load_library().rrb_shape_set_transform(self._h, _np_ptr(m), _np_ptr(minv)))

    def GetTransform(self):
        m, minv = np.zeros(16, np.float32), np.zeros(16, np.float32)
        _check(load_library().rrb_shape_get_transform(self._h, _np_ptr(m), _np_ptr(minv)))
        return m.reshape(4, 4), minv.reshape(4, 4)

    def SetId(self, id_):
        _check(load_library().rrb_shape_set_id(self._h, int(id_)))

    def GetId(self):
        out = ctypes.c_int()
        _check(load_library().rrb_shape_get_id(self._h, ctypes.byref(out)))
        return out.value

    def SetLinearVelocity(self, v):
        v = np.ascontiguousarray(v, np.float32)
        _check(load_library().rrb_shape_set_linear_velocity(self._h, _np_ptr(v)))

    def SetAngularVelocity(self, q):
        q = np.ascontiguousarray(q, np.float32)
        _check(load_library().rrb_shape_set_angular_velocity(self._h, _np_ptr(q)))


class Buffer:
    """RadeonRays::Buffer: a chunk of device memory hosted inside the API or wrapped (interop)."""

    def __init__(self, handle, size):
        self._h = handle
        self.size = size

    @property
    def device_ptr(self):
        p = ctypes.c_void_p()
        _check(load_library().rrb_buffer_device_ptr(self._h, ctypes.byref(p), None))
        return p.value


class Event:
    """RadeonRays::Event (radeon_rays.h:120-129)."""

    def __init__(self, handle):
        self._h = handle

    def Complete(self):
        out = ctypes.c_int()
        _check(load_library().rrb_event_complete(self._h, ctypes.byref(out)))
        return bool(out.value)

    def Wait(self):
        _check(load_library().rrb_event_wait(self._h))


class IntersectionApi:
    """RadeonRays::IntersectionApi (radeon_rays.h:170-305) on one CUDA device."""

    @staticmethod
    def GetDeviceCount():
        return load_library().rrb_device_count()

    @staticmethod
    def GetDeviceInfo(devidx):
        name, vendor, t = ctypes.create_string_buffer(256), ctypes.create_string_buffer(64), ctypes.c_int()
        _check(load_library().rrb_device_info(devidx, name, 256, vendor, 64, ctypes.byref(t)))
        return {"name": name.value.decode(), "vendor": vendor.value.decode(), "type": t.value}

    @classmethod
    def Create(cls, devidx=0, cuda_stream=None):
        h = ctypes.c_void_p()
        _check(load_library().rrb_create(int(devidx), cuda_stream, ctypes.byref(h)))
        return cls(h)

    @staticmethod
    def Delete(api):
        if api is not None and api._h:
            load_library().rrb_destroy(api._h)
            api._h = None

    def __init__(self, handle):
        self._h = handle
        self._lib = load_library()

    # ---- geometry
    def CreateMesh(self, vertices, vnum, vstride, indices, istride, numfacevertices, numfaces):
        vertices = np.ascontiguousarray(vertices, np.float32)
        indices = np.ascontiguousarray(indices, np.int32)
        nfv = None if numfacevertices is None else np.ascontiguousarray(numfacevertices, np.int32)
        h = ctypes.c_void_p()
        _check(self._lib.rrb_create_mesh(self._h, _np_ptr(vertices), int(vnum), int(vstride), _np_ptr(indices),
                                         int(istride), _np_ptr(nfv), int(numfaces), ctypes.byref(h)))
        return Shape(h)

    def CreateInstance(self, shape):
        h = ctypes.c_void_p()
        _check(self._lib.rrb_create_instance(self._h, shape._h, ctypes.byref(h)))
        return Shape(h)

    def DeleteShape(self, shape):
        _check(self._lib.rrb_delete_shape(self._h, shape._h))
        shape._h = None

    def AttachShape(self, shape):
        _check(self._lib.rrb_attach(self._h, shape._h))

    def DetachShape(self, shape):
        _check(self._lib.rrb_detach(self._h, shape._h))

    def DetachAll(self):
        _check(self._lib.rrb_detach_all(self._h))

    def Commit(self):
        _check(self._lib.rrb_commit(self._h))

    def ResetIdCounter(self):
        _check(self._lib.rrb_reset_id_counter(self._h))

    def IsWorldEmpty(self):
        out = ctypes.c_int()
        _check(self._lib.rrb_is_world_empty(self._h, ctypes.byref(out)))
        return bool(out.value)

    def SetOption(self, name, value):
        if isinstance(value, str):
            _check(self._lib.rrb_set_option_str(self._h, name.encode(), value.encode()))
        else:
            _check(self._lib.rrb_set_option_f32(self._h, name.encode(), float(value)))

    # ---- memory
    def CreateBuffer(self, size, initdata=None):
        h = ctypes.c_void_p()
        init = None
        if initdata is not None:
            init = np.ascontiguousarray(initdata)
            assert init.nbytes >= size
        _check(self._lib.rrb_create_buffer(self._h, int(size), _np_ptr(init), ctypes.byref(h)))
        return Buffer(h, int(size))

    def CreateFromDevicePtr(self, device_ptr, size):
        """Counterpart of CreateFromOpenClBuffer (radeon_rays_cl.h:41): non-owning wrap."""
        h = ctypes.c_void_p()
        _check(self._lib.rrb_wrap_device_ptr(self._h, ctypes.c_void_p(device_ptr), int(size), ctypes.byref(h)))
        return Buffer(h, int(size))

    def DeleteBuffer(self, buf):
        _check(self._lib.rrb_delete_buffer(self._h, buf._h))
        buf._h = None

    def MapBuffer(self, buf, map_type, offset, size, dtype=np.uint8):
        """Returns (host ndarray view, Event).  The view is valid until UnmapBuffer."""
        data, ev = ctypes.c_void_p(), ctypes.c_void_p()
        _check(self._lib.rrb_map(self._h, buf._h, int(map_type), int(offset), int(size), ctypes.byref(data),
                                 ctypes.byref(ev)))
        raw = (ctypes.c_uint8 * int(size)).from_address(data.value)
        arr = np.frombuffer(raw, dtype=dtype)
        return arr, Event(ev)

    def UnmapBuffer(self, buf, arr=None):
        ev = ctypes.c_void_p()
        _check(self._lib.rrb_unmap(self._h, buf._h, None, ctypes.byref(ev)))
        return Event(ev)

    def DeleteEvent(self, ev):
        _check(self._lib.rrb_event_delete(self._h, ev._h))
        ev._h = None

    # ---- ray casting
    def QueryIntersection(self, rays, numrays, hits, waitevent=None, want_event=True):
        ev = ctypes.c_void_p()
        w = None if waitevent is None else waitevent._h
        if isinstance(numrays, Buffer):
            raise TypeError("use QueryIntersectionIndirect for a device-side ray count")
        _check(self._lib.rrb_query_intersection(self._h, rays._h, int(numrays), hits._h, w,
                                                ctypes.byref(ev) if want_event else None))
        return Event(ev) if want_event else None

    def QueryOcclusion(self, rays, numrays, hitresults, waitevent=None, want_event=True):
        ev = ctypes.c_void_p()
        w = None if waitevent is None else waitevent._h
        _check(self._lib.rrb_query_occlusion(self._h, rays._h, int(numrays), hitresults._h, w,
                                             ctypes.byref(ev) if want_event else None))
        return Event(ev) if want_event else None

    def QueryIntersectionIndirect(self, rays, numrays_buf, maxrays, hits, waitevent=None):
        ev = ctypes.c_void_p()
        w = None if waitevent is None else waitevent._h
        _check(self._lib.rrb_query_intersection_indirect(self._h, rays._h, numrays_buf._h, int(maxrays), hits._h, w,
                                                         ctypes.byref(ev)))
        return Event(ev)

    def QueryOcclusionIndirect(self, rays, numrays_buf, maxrays, hitresults, waitevent=None):
        ev = ctypes.c_void_p()
        w = None if waitevent is None else waitevent._h
        _check(self._lib.rrb_query_occlusion_indirect(self._h, rays._h, numrays_buf._h, int(maxrays), hitresults._h,
                                                      w, ctypes.byref(ev)))
        return Event(ev)

    # ---- host-buffer path (numpy in, numpy out; copies inside)
    def QueryIntersectionHost(self, rays: np.ndarray, hits: np.ndarray = None) -> np.ndarray:
        assert rays.dtype == L.RAY and rays.flags.c_contiguous
        if hits is None:
            hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
            hits["shapeid"] = K_NULL_ID
            hits["primid"] = K_NULL_ID
        _check(self._lib.rrb_query_intersection_host(self._h, _np_ptr(rays), rays.shape[0], _np_ptr(hits)))
        return hits

    def QueryOcclusionHost(self, rays: np.ndarray, results: np.ndarray = None) -> np.ndarray:
        assert rays.dtype == L.RAY and rays.flags.c_contiguous
        if results is None:
            results = np.zeros(rays.shape[0], dtype=np.int32)
        _check(self._lib.rrb_query_occlusion_host(self._h, _np_ptr(rays), rays.shape[0], _np_ptr(results)))
        return results

    def QueryIntersectionNoAlloc(self, rays, numrays, hits):
        """rrb_query_intersection_noalloc on the device pointers of two Buffers (or raw ints): no allocation inside,
        asynchronous on the API's stream, at most RRB_NOALLOC_MAX_RAYS rays (include/rrb.h)."""
        rp = rays.device_ptr if isinstance(rays, Buffer) else int(rays)
        hp = hits.device_ptr if isinstance(hits, Buffer) else int(hits)
        _check(self._lib.rrb_query_intersection_noalloc(self._h, rp, int(numrays), hp))

    # ---- measurement
    def GetStats(self):
        s = RrbStats()
        _check(self._lib.rrb_get_stats(self._h, ctypes.byref(s)))
        return {k: getattr(s, k) for k, _ in RrbStats._fields_}

    def GetBvh(self):
        """(nodes device pointer, triangles device pointer, num_nodes, num_tris) of the committed tree."""
        pn, pt = ctypes.c_void_p(), ctypes.c_void_p()
        nn, nt = ctypes.c_int64(), ctypes.c_int64()
        _check(self._lib.rrb_get_bvh(self._h, ctypes.byref(pn), ctypes.byref(pt), ctypes.byref(nn), ctypes.byref(nt)))
        return pn.value, pt.value, nn.value, nt.value

    def QueryIntersectionCounted(self, rays, numrays, hits, counts):
        _check(self._lib.rrb_query_intersection_counted(self._h, rays._h, int(numrays), hits._h, counts._h))


def attach_scene(api: IntersectionApi, scene):
    """RTScene::attachMesh's Radeon Rays calls (scene/RTScene.cpp:659-670): one mesh per RTShape,
    id = index into the RTShape array, world transform set on the shape."""
    shapes = []
    for i, m in enumerate(scene.meshes):
        if getattr(m, "instance_of", None) is not None:      # RTScene.cpp:579: CreateInstance(instanceTemplateShape)
            s = api.CreateInstance(shapes[int(m.instance_of)])
        else:
            s = api.CreateMesh(m.positions, m.positions.shape[0], 12, m.indices, 0, None, m.indices.shape[0])
        s.SetId(i)
        if m.transform is not None:
            s.SetTransform(m.transform)
        api.AttachShape(s)
        shapes.append(s)
    return shapes


def host_bvh(vertices, indices, face_shape_ids=None, use_sah=True, refit_info=False):
    """Build the compressed wide BVH on the host only (no CUDA): returns (nodes u32 (N,20), tris f32 (T,12));
    with refit_info also (tri_mesh i32 (T,), level_offsets i32 (levels + 1,))."""
    lib = load_library()
    v = np.ascontiguousarray(vertices, np.float32)
    ix = np.ascontiguousarray(indices, np.int32)
    ids = None if face_shape_ids is None else np.ascontiguousarray(face_shape_ids, np.int32)
    h = ctypes.c_void_p()
    _check(lib.rrb_host_bvh_build(_np_ptr(v), v.shape[0], _np_ptr(ix), ix.shape[0], _np_ptr(ids), int(use_sah),
                                  ctypes.byref(h)))
    try:
        pn, pt = ctypes.c_void_p(), ctypes.c_void_p()
        nn, nt = ctypes.c_int64(), ctypes.c_int64()
        _check(lib.rrb_host_bvh_get(h, ctypes.byref(pn), ctypes.byref(nn), ctypes.byref(pt), ctypes.byref(nt)))
        nodes = np.frombuffer(ctypes.string_at(pn.value, nn.value * 80), np.uint32).reshape(-1, 20).copy()
        tris = np.frombuffer(ctypes.string_at(pt.value, nt.value * 48), np.float32).reshape(-1, 12).copy()
        extra = ()
        if refit_info:
            pm, pl, nl = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int()
            _check(lib.rrb_host_bvh_get_refit_info(h, ctypes.byref(pm), ctypes.byref(pl), ctypes.byref(nl)))
            extra = (np.frombuffer(ctypes.string_at(pm.value, nt.value * 4), np.int32).copy(),
                     np.frombuffer(ctypes.string_at(pl.value, (nl.value + 1) * 4), np.int32).copy())
    finally:
        lib.rrb_host_bvh_free(h)
    return (nodes, tris) + extra
