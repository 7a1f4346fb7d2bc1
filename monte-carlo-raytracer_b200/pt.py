"""Host-side mirror of the engine's path-tracer pass interface over the C ABI (include/ptb.h).

The reference dispatches the hot path through `RenderPass::update()` objects registered with a
`RenderPipeline` that doubles as a string-keyed blackboard
(source/engine/rendering/architecture/RenderPass.h:6-32, RenderPipeline.h:49-71,
RenderPipeline.cpp:36-67).  The classes here keep those names, the blackboard keys and the
`update()` call shape; their bodies call the CUDA library instead of CLW:

    RTPrimaryRaysPass   renderPasses/RTPrimaryRaysPass.cpp:32-134
    RTPathTracingPass   renderPasses/RTPathTracingPass.cpp:40-218
    RTBDPTPass          renderPasses/RTBDPTPass.cpp:67-511
    RTReconstructionPass renderPasses/RTReconstructionPass.cpp:41-123

ctypes is plumbing only; there is no CPU fallback.
"""
import ctypes

import numpy as np

from . import camera as camera_mod
from . import layouts as L
from . import rr


class PtbSceneDesc(ctypes.Structure):
    _fields_ = [("shapes", ctypes.c_void_p), ("indices", ctypes.c_void_p), ("positions", ctypes.c_void_p),
                ("uvs", ctypes.c_void_p), ("normals", ctypes.c_void_p), ("textures", ctypes.c_void_p),
                ("texdata", ctypes.c_void_p), ("sobol", ctypes.c_void_p), ("lights", ctypes.c_void_p),
                ("materials", ctypes.c_void_p), ("camera", ctypes.c_void_p),
                ("num_shapes", ctypes.c_int32), ("num_textures", ctypes.c_int32), ("num_lights", ctypes.c_int32),
                ("num_materials", ctypes.c_int32), ("num_indices", ctypes.c_int64), ("num_vertices", ctypes.c_int64),
                ("texdata_bytes", ctypes.c_int64)]


class PtbFilterProps(ctypes.Structure):
    _fields_ = [("filter_type", ctypes.c_int32), ("radius", ctypes.c_float * 2), ("mitchell_b", ctypes.c_float),
                ("mitchell_c", ctypes.c_float), ("lanczos_sinc_tau", ctypes.c_float),
                ("gaussian_alpha", ctypes.c_float), ("gaussian_exp_x", ctypes.c_float),
                ("gaussian_exp_y", ctypes.c_float), ("pixel_offset", ctypes.c_float * 2)]


class PtbCounters(ctypes.Structure):
    _fields_ = [("closest_rays", ctypes.c_int64), ("shadow_rays", ctypes.c_int64), ("shaded_vertices", ctypes.c_int64),
                ("samples", ctypes.c_int64), ("kernel_launches", ctypes.c_int64)]


_bound = False


def _lib():
    global _bound
    lib = rr.load_library()
    if not _bound:
        vp, i, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
        pp = ctypes.POINTER(ctypes.c_void_p)
        sig = {
            "ptb_create": [vp, pp], "ptb_destroy": [vp], "ptb_upload_scene": [vp, ctypes.POINTER(PtbSceneDesc)],
            "ptb_update_scene_array": [vp, ctypes.c_char_p, vp, ctypes.c_size_t],
            "ptb_resize": [vp, i, i], "ptb_set_camera": [vp, vp], "ptb_set_partition": [vp, i, i],
            "ptb_generate_primary_rays": [vp], "ptb_pt_shade": [vp, i, i, i], "ptb_pt_shadow_resolve": [vp, i],
            "ptb_pt_render_sample": [vp, i, i], "ptb_bdpt_render_sample": [vp, i, i],
            "ptb_accumulate": [vp, i, ctypes.POINTER(PtbFilterProps)], "ptb_reset_accumulation": [vp],
            "ptb_get_buffer": [vp, ctypes.c_char_p, pp, ctypes.POINTER(sz)],
            "ptb_read_buffer": [vp, ctypes.c_char_p, vp, sz], "ptb_synchronize": [vp],
            "ptb_get_counters": [vp, ctypes.POINTER(PtbCounters)], "ptb_reset_counters": [vp],
            "ptb_set_sampler": [vp, i, vp, i],
            "ptb_denoise": [vp, ctypes.c_char_p, i, ctypes.c_float, ctypes.c_float],
            "ptb_tonemap": [vp, ctypes.c_char_p, ctypes.c_float],
            "ptb_profile_enable": [vp, ctypes.c_int], "ptb_set_overlap": [vp, ctypes.c_int],
            "ptb_set_coherence": [vp, ctypes.c_int], "ptb_set_estimator": [vp, ctypes.c_int],
            "ptb_set_shade_split": [vp, ctypes.c_int],
            "ptb_set_cooperative_tail": [vp, ctypes.c_int],
            "ptb_set_frames_in_flight": [vp, ctypes.c_int],
            "ptb_selftest_fdiv": [vp, vp, vp, vp, vp, ctypes.c_int],
            "ptb_read_buffer_async": [vp, ctypes.c_char_p, vp, sz], "ptb_read_wait": [vp],
            "ptb_profile_collect": [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)],
            "ptb_profile_collect_kernels": [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)],
            "ptb_selftest_libm": [vp, i, vp, vp, i],
            "ptb_set_path_trace": [vp, i], "ptb_write_buffer": [vp, ctypes.c_char_p, vp, sz],
            "ptb_bdpt_reserve": [vp, i], "ptb_merge_accumulation": [vp, vp, i],
            "ptb_nccl_unique_id": [vp], "ptb_nccl_comm_create": [vp, vp, i, i, pp], "ptb_nccl_comm_destroy": [vp],
        }
        for name, args in sig.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = ctypes.c_int
        lib.ptb_last_error.restype = ctypes.c_char_p
        _bound = True
    return lib


class PathTracerError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise PathTracerError(_lib().ptb_last_error().decode())


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(ctypes.c_void_p)


# PathTracerSettings::GI defaults (GUI/PathTracingSettings.h:52-81)
class GISettings:
    def __init__(self):
        self.maxDepth = 2
        self.imageResolution = (1280, 720)
        self.useTAA = True
        self.filterType = L.RT_TRIANGLE_FILTER
        self.filterRadius = (2.0, 2.0)
        self.mitchellB = 1.0 / 3.0
        self.mitchellC = 1.0 / 3.0
        self.lanczosSincTau = 3.0
        self.gaussianAlpha = 1.0
        self.curPixelOffset = (0.0, 0.0)
        # post-processing (PathTracingSettings.h:82-87)
        self.denoiseKernelRadius = 1
        self.bilateralDenoiseSigmaRange = 0.1
        self.bilateralDenoiseSigmaSpatial = 1.0
        self.useDenoise = False
        self.useTonemapping = False
        self.minLuminance = 2.0

    def filter_props(self) -> PtbFilterProps:
        f = PtbFilterProps()
        f.filter_type = self.filterType
        f.radius[0], f.radius[1] = self.filterRadius
        f.mitchell_b, f.mitchell_c = self.mitchellB, self.mitchellC
        f.lanczos_sinc_tau = self.lanczosSincTau
        f.gaussian_alpha = self.gaussianAlpha
        # gaussianExpX/Y = exp(-alpha * r^2) (pbrt GaussianFilter; RTReconstructionPass.cpp:95-97)
        f.gaussian_exp_x = float(np.exp(np.float32(-self.gaussianAlpha * self.filterRadius[0] ** 2)))
        f.gaussian_exp_y = float(np.exp(np.float32(-self.gaussianAlpha * self.filterRadius[1] ** 2)))
        f.pixel_offset[0], f.pixel_offset[1] = self.curPixelOffset
        return f


class RTScene:
    """Device scene + Radeon Rays world: RTScene's constructor, attachMesh and commit
    (scene/RTScene.cpp:28-56,199-267,564-678) for a headless host scene."""

    def __init__(self, api: rr.IntersectionApi, scene, refit=False, acc_type="fatbvh", builder="sah"):
        self.api = api
        self.host = scene
        self.shapes = rr.attach_scene(api, scene)
        if refit:       # extension: transform-only commits refit the tree on the device (include/rrb.h)
            api.SetOption("bvh.refit", 1.0)
        # RTScene::commit: acc.type / bvh.builder from the GUI defaults (PathTracingSettings.h:241-249)
        # (acc.type = "hlbvh", Radeon Rays' GPU-built hierarchy, selects the device builder here too)
        api.SetOption("acc.type", acc_type)
        api.SetOption("bvh.builder", builder)
        api.SetOption("bvh.sah.traversal_cost", 10.0)
        api.SetOption("bvh.sah.num_bins", 64.0)
        api.Commit()
        h = ctypes.c_void_p()
        _check(_lib().ptb_create(api._h, ctypes.byref(h)))
        self._h = h
        self.upload()

    def upload(self):
        s = self.host
        d = PtbSceneDesc()
        d.shapes, d.indices, d.positions = _ptr(s.shapes), _ptr(s.indices), _ptr(s.positions)
        d.uvs, d.normals = _ptr(s.uvs), _ptr(s.normals)
        d.textures, d.texdata, d.sobol = _ptr(s.textures), _ptr(s.texdata), None
        d.lights, d.materials, d.camera = _ptr(s.lights), _ptr(s.materials), None
        d.num_shapes, d.num_textures = len(s.shapes), len(s.textures)
        d.num_lights, d.num_materials = len(s.lights), len(s.materials)
        d.num_indices, d.num_vertices = s.indices.size, s.positions.shape[0]
        d.texdata_bytes = s.texdata.size
        _check(_lib().ptb_upload_scene(self._h, ctypes.byref(d)))

    def updateDynamicEntities(self, transforms=None, lights=None, materials=None):
        """RTScene::updateDynamicEntities (scene/RTScene.cpp:317-389): `transforms` maps an RTShape index to its
        new row-major local-to-world matrix -> Shape::SetTransform, RTShape::toWorldTransform /
        toWorldInverseTranspose, one write of the shapes array, Commit; `lights` / `materials` are whole
        replacement arrays (same length)."""
        s = self.host
        if transforms:
            for i, m in transforms.items():
                m = np.ascontiguousarray(m, np.float32)
                self.shapes[i].SetTransform(m)
                s.meshes[i].transform = m
                s.shapes[i]["toWorldTransform"] = m
                s.shapes[i]["toWorldInverseTranspose"] = np.linalg.inv(m.astype(np.float64)).T.astype(np.float32)
            _check(_lib().ptb_update_scene_array(self._h, b"shapes", _ptr(s.shapes), s.shapes.nbytes))
            self.api.Commit()
        if lights is not None:
            lights = np.ascontiguousarray(lights, L.LIGHT)
            assert lights.shape == s.lights.shape
            s.lights = lights
            _check(_lib().ptb_update_scene_array(self._h, b"lights", _ptr(s.lights), s.lights.nbytes))
        if materials is not None:
            materials = np.ascontiguousarray(materials, L.MATERIAL)
            assert materials.shape == s.materials.shape
            s.materials = materials
            _check(_lib().ptb_update_scene_array(self._h, b"materials", _ptr(s.materials), s.materials.nbytes))

    def destroy(self):
        if self._h:
            _lib().ptb_destroy(self._h)
            self._h = None


class RenderPipeline:
    """engine/rendering/architecture/RenderPipeline: ordered passes + string-keyed blackboard."""

    def __init__(self, rt_scene: RTScene, settings: GISettings = None):
        self.scene = rt_scene
        self.GI = settings or GISettings()
        self.passes = []
        self.board = {}
        self.frameIndex = 0
        self._size = None

    def addRenderPasses(self, *passes):
        for p in passes:
            p.m_renderPipeline = self
            self.passes.append(p)

    def put(self, key, value):
        self.board[key] = value

    def fetch(self, key):
        return self.board.get(key)

    putPtr, fetchPtr = put, fetch

    def ensure_size(self):
        if self._size != tuple(self.GI.imageResolution):
            w, h = self.GI.imageResolution
            _check(_lib().ptb_resize(self.scene._h, int(w), int(h)))
            self._size = (int(w), int(h))

    def set_camera_for_frame(self):
        """PathTracingApp::update TAA jitter (:208-215) + camera upload (RTPrimaryRaysPass.cpp:96-107)."""
        self.ensure_size()
        w, h = self._size
        if self.GI.useTAA:
            self.GI.curPixelOffset = camera_mod.taa_pixel_offset(self.frameIndex, self.GI.filterRadius, True)
        cam = camera_mod.make_camera(self.scene.host.camera, w, h, self.GI.curPixelOffset)
        _check(_lib().ptb_set_camera(self.scene._h, _ptr(cam)))
        return cam

    def update(self):
        for p in self.passes:
            p.update()

    # ---- buffer helpers
    def device_buffer(self, name):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(_lib().ptb_get_buffer(self.scene._h, name.encode(), ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def read_buffer(self, name, dtype=np.float32, out=None):
        _, n = self.device_buffer(name)
        if out is None:
            out = np.empty(n // np.dtype(dtype).itemsize, dtype=dtype)
        _check(_lib().ptb_read_buffer(self.scene._h, name.encode(), _ptr(out), out.nbytes))
        return out

    def read_buffer_async(self, name, out):
        """Snapshot + copy to the (pinned) array `out` behind the next frame; valid after read_wait()."""
        _check(_lib().ptb_read_buffer_async(self.scene._h, name.encode(), _ptr(out), out.nbytes))

    def read_wait(self):
        _check(_lib().ptb_read_wait(self.scene._h))

    def counters(self):
        c = PtbCounters()
        _check(_lib().ptb_get_counters(self.scene._h, ctypes.byref(c)))
        return {k: getattr(c, k) for k, _ in PtbCounters._fields_}

    def reset_counters(self):
        _check(_lib().ptb_reset_counters(self.scene._h))

    def synchronize(self):
        _check(_lib().ptb_synchronize(self.scene._h))

    def profile_enable(self, on=True):
        _check(_lib().ptb_profile_enable(self.scene._h, int(bool(on))))

    STAGES = ("k_trace_closest", "k_trace_mixed", "k_trace_shadow", "k_shade")     # ptb_stage

    def profile_collect(self):
        """{kernel: (summed device ms, launches)} of the wavefront's kernels since the last collect."""
        ms, n = (ctypes.c_double * 4)(), (ctypes.c_int64 * 4)()
        _check(_lib().ptb_profile_collect(self.scene._h, ms, n))
        return {k: (ms[i], n[i]) for i, k in enumerate(self.STAGES)}

    KERNELS = STAGES + ("k_shade_scatter",)     # ptb_profile_collect_kernels: [3] is the surface kernel when [4] ran

    def profile_collect_kernels(self):
        """Like profile_collect, per kernel: with the split shade stage "k_shade" is the surface kernel alone and
        "k_shade_scatter" the scatter kernel (zero launches for the single shade kernel)."""
        ms, n = (ctypes.c_double * 5)(), (ctypes.c_int64 * 5)()
        _check(_lib().ptb_profile_collect_kernels(self.scene._h, ms, n))
        return {k: (ms[i], n[i]) for i, k in enumerate(self.KERNELS)}

    def set_overlap(self, halves):
        _check(_lib().ptb_set_overlap(self.scene._h, int(halves)))

    def set_frames_in_flight(self, depth):
        """Render up to `depth` consecutive frames concurrently on side streams (include/ptb.h); results unchanged."""
        _check(_lib().ptb_set_frames_in_flight(self.scene._h, int(depth)))

    def set_coherence(self, mode):
        """Queue order of the fused wavefront: 0 push order, 1 bounces >= 1 bucketed by the sampler's light / lobe
        decisions, 2 bounce 0 too (include/ptb.h)."""
        _check(_lib().ptb_set_coherence(self.scene._h, int(mode)))

    def set_shade_split(self, on):
        """The shade stage as one kernel per bounce (False), as surface + scatter kernels always (True) or on large
        wavefronts only ("auto", the default); identical results
        (include/ptb.h: ptb_set_shade_split)."""
        mode = {False: 0, True: 2, "auto": 1}.get(on, on)       # True = always (tests), "auto" = the library's default
        _check(_lib().ptb_set_shade_split(self.scene._h, int(mode)))

    def set_cooperative_tail(self, min_iters):
        """Cooperative tail of the traversal kernels: split a ray walked `min_iters` rounds after the queue ran dry
        over the warp's idle lanes (include/ptb.h); 0 = off; identical results."""
        _check(_lib().ptb_set_cooperative_tail(self.scene._h, int(min_iters)))

    def set_estimator(self, russian_roulette=False, nee_mis=False):
        """Opt-in estimator switches (include/ptb.h PTB_EST_*); both off = the reference's estimator."""
        _check(_lib().ptb_set_estimator(self.scene._h, (1 if russian_roulette else 0) | (2 if nee_mis else 0)))

    def write_buffer(self, name, data):
        """Whole-buffer host -> device write (ptb_write_buffer): checkpoint restore."""
        data = np.ascontiguousarray(data)
        _check(_lib().ptb_write_buffer(self.scene._h, name.encode(), _ptr(data), data.nbytes))

    def set_path_trace(self, max_depth):
        """Diagnostic per-bounce record of the fused wavefront ("PathTraceBuffer", include/ptb.h); 0 = off."""
        self.ensure_size()
        _check(_lib().ptb_set_path_trace(self.scene._h, int(max_depth)))

    def path_trace(self, max_depth):
        """(max_depth, P, 4) int32 records of the frame rendered last."""
        w, h = self._size
        return self.read_buffer("PathTraceBuffer", np.int32).reshape(-1, w * h, 4)[:max_depth]

    def bdpt_connections(self):
        """Per-strategy records of the BDPT frame rendered last (include/ptb.h: "BDPTConnection*"): dict of arrays
        over the N surviving (pixel, t, s) strategies -- pixel, t, s, target, C (N, 3), visible, weight, ray_o, ray_d."""
        n = int(self.read_buffer("BDPTCounters", np.uint32)[6])

        def head(name, dtype, width):
            out = np.empty(n * width, dtype=dtype)
            if n:
                _check(_lib().ptb_read_buffer(self.scene._h, name.encode(), _ptr(out), out.nbytes))
            return out.reshape(n, width) if width > 1 else out
        C = head("BDPTConnectionContribution", np.float32, 4)
        d = head("BDPTConnectionRayD", np.float32, 4)
        o = head("BDPTConnectionRayO", np.float32, 4)
        ts = d[:, 3].copy().view(np.int32)
        return {"pixel": head("BDPTConnectionPixel", np.int32, 1), "t": ts & 0xFF, "s": ts >> 8,
                "target": C[:, 3].copy().view(np.int32), "C": C[:, :3].copy(),
                "visible": head("BDPTConnectionVisibility", np.int32, 1) != 0,
                "weight": head("BDPTConnectionWeight", np.float32, 1), "ray_o": o, "ray_d": d[:, :3].copy()}

    def merge_accumulation(self, comm=None, root=0):
        """The one collective of the multi-GPU path (ptb_merge_accumulation): NCCL reduce of the accumulation
        buffers to `root` (all-reduce when root < 0) + the division, on the context's stream."""
        _check(_lib().ptb_merge_accumulation(self.scene._h, comm, int(root)))

    # ---- checkpoint / resume of the accumulation state (the reference keeps it only on the device,
    # RTReconstructionPass.cpp:212-213)
    def _integrator(self):
        for p in self.passes:
            if isinstance(p, (RTPathTracingPass, RTBDPTPass)):
                return p
        raise PathTracerError("the pipeline has no integrator pass")

    def save_checkpoint(self, path):
        """(sum(w L), sum(w), frame index) -- and for BDPT the sampled-light records the next frame reads before
        it rewrites them (E12) -- as one .npz; a render resumed from it continues bit for bit."""
        integ = self._integrator()
        w, h = self._size
        state = {"width": w, "height": h, "frame_index": integ.m_frameIndex, "max_depth": int(self.GI.maxDepth),
                 "bdpt": isinstance(integ, RTBDPTPass),
                 "accumulation": self.read_buffer("AccumulationBuffer"),
                 "image": self.read_buffer("NextFrameImageCL")}
        if state["bdpt"] and integ.m_frameIndex > 0:
            state["bdpt_sampled_light"] = self.read_buffer("BDPTSampledLightBuffer")
        np.savez(path, **state)

    def load_checkpoint(self, path):
        z = np.load(path)
        self.ensure_size()
        integ = self._integrator()
        if (int(z["width"]), int(z["height"])) != self._size or bool(z["bdpt"]) != isinstance(integ, RTBDPTPass) \
                or int(z["max_depth"]) != int(self.GI.maxDepth):
            raise PathTracerError("checkpoint does not match this pipeline (size, integrator or depth)")
        self.write_buffer("AccumulationBuffer", z["accumulation"])
        self.write_buffer("NextFrameImageCL", z["image"])
        if "bdpt_sampled_light" in z.files:
            _check(_lib().ptb_bdpt_reserve(self.scene._h, int(self.GI.maxDepth)))
            self.write_buffer("BDPTSampledLightBuffer", z["bdpt_sampled_light"])
        integ.m_frameIndex = int(z["frame_index"])
        self.frameIndex = integ.m_frameIndex

    def set_partition(self, part, num_parts):
        _check(_lib().ptb_set_partition(self.scene._h, int(part), int(num_parts)))

    def set_sampler(self, kind, matrices=None):
        """samplers.cl:15-17: kind 0 = Sobol (needs the generator matrices, sobol.matrices()), 1 = random."""
        n = 0
        if matrices is not None:
            matrices = np.ascontiguousarray(matrices, np.uint32)
            n = matrices.size // 52
        _check(_lib().ptb_set_sampler(self.scene._h, int(kind), _ptr(matrices), n))


class RenderPass:
    m_renderPipeline: RenderPipeline = None

    def update(self):
        raise NotImplementedError


class RTPrimaryRaysPass(RenderPass):
    """Stage-wise primary rays + primary hits in the reference's layouts (RTPrimaryRaysPass.cpp:32-66)."""

    def update(self):
        pl = self.m_renderPipeline
        pl.set_camera_for_frame()
        h, api = pl.scene._h, pl.scene.api
        _check(_lib().ptb_generate_primary_rays(h))
        w, hgt = pl._size
        rp, rn = pl.device_buffer("RayBuffer")
        ip, inb = pl.device_buffer("PrimaryIntersectionBufferCL")
        rays = api.CreateFromDevicePtr(rp, rn)      # the engine wraps on every call (:133)
        isect = api.CreateFromDevicePtr(ip, inb)
        api.QueryIntersection(rays, w * hgt, isect).Wait()
        pl.putPtr("RayBuffer", rays)
        pl.putPtr("PrimaryIntersectionBufferCL", isect)


class RTPathTracingPass(RenderPass):
    """RTPathTracingPass::update (RTPathTracingPass.cpp:40-114).

    fused=True (default) runs the whole frame -- primary rays included -- as the queue-based
    wavefront (ptb_pt_render_sample); fused=False keeps the reference's stage-by-stage dispatch over
    the reference-layout buffers and needs an RTPrimaryRaysPass before it."""

    def __init__(self, fused=True):
        self.fused = fused
        self.m_frameIndex = 0

    def update(self):
        pl = self.m_renderPipeline
        h, api = pl.scene._h, pl.scene.api
        depth = int(pl.GI.maxDepth)
        if len(pl.scene.host.lights) == 0:
            return                                   # :42-43
        pl.frameIndex = self.m_frameIndex
        if self.fused:
            pl.set_camera_for_frame()
            _check(_lib().ptb_pt_render_sample(h, self.m_frameIndex, depth))
        else:
            w, hgt = pl._size
            rays, isect = pl.fetchPtr("RayBuffer"), pl.fetchPtr("PrimaryIntersectionBufferCL")
            sp, sn = pl.device_buffer("ShadowRayBuffer")
            op, on = pl.device_buffer("ShadowRayOcclusionBuffer")
            shadow, occl = api.CreateFromDevicePtr(sp, sn), api.CreateFromDevicePtr(op, on)
            for i in range(depth):
                _check(_lib().ptb_pt_shade(h, self.m_frameIndex, depth, i))            # applyShading
                api.QueryOcclusion(shadow, w * hgt, occl).Wait()                       # applyVisibilityTest
                _check(_lib().ptb_pt_shadow_resolve(h, i))
                if i + 1 < depth:
                    api.QueryIntersection(rays, w * hgt, isect).Wait()
        pl.put("RadianceBufferCL", pl.device_buffer("RadianceBufferCL")[0])
        self.m_frameIndex += 1
        pl.frameIndex = self.m_frameIndex


class RTBDPTPass(RenderPass):
    """RTBDPTPass::update (RTBDPTPass.cpp:67-128)."""

    def __init__(self):
        self.m_frameIndex = 0

    def update(self):
        pl = self.m_renderPipeline
        if len(pl.scene.host.lights) == 0:
            return
        pl.frameIndex = self.m_frameIndex
        pl.set_camera_for_frame()
        _check(_lib().ptb_bdpt_render_sample(pl.scene._h, self.m_frameIndex, int(pl.GI.maxDepth)))
        pl.put("RadianceBufferCL", pl.device_buffer("RadianceBufferCL")[0])
        self.m_frameIndex += 1
        pl.frameIndex = self.m_frameIndex


class RTReconstructionPass(RenderPass):
    """RTReconstructionPass::updateReconstruction (RTReconstructionPass.cpp:71-123): the frame that was
    just rendered is frameIndex - 1 (the integrator pass has already incremented it)."""

    def update(self):
        pl = self.m_renderPipeline
        if pl.fetch("RadianceBufferCL") is None:
            return
        f = pl.GI.filter_props()
        _check(_lib().ptb_accumulate(pl.scene._h, pl.frameIndex - 1, ctypes.byref(f)))
        pl.put("NextFrameImageCL", pl.device_buffer("NextFrameImageCL")[0])
        pl.put("NextFrameImageName", "NextFrameImageCL")


class RTDenoisePass(RenderPass):
    """RTDenoisePass::update (RTDenoisePass.cpp:33-75): bilateral filter of "NextFrameImageCL"; its output
    replaces that blackboard entry for the passes after it."""

    def update(self):
        pl = self.m_renderPipeline
        if not pl.GI.useDenoise or pl.fetch("NextFrameImageCL") is None:
            return
        src = pl.fetch("NextFrameImageName") or "NextFrameImageCL"
        _check(_lib().ptb_denoise(pl.scene._h, src.encode(), int(pl.GI.denoiseKernelRadius),
                                  float(pl.GI.bilateralDenoiseSigmaSpatial), float(pl.GI.bilateralDenoiseSigmaRange)))
        pl.put("NextFrameImageCL", pl.device_buffer("DenoisedImageCL")[0])
        pl.put("NextFrameImageName", "DenoisedImageCL")


class RTToneMappingPass(RenderPass):
    """RTToneMappingPass::applyReinhardToneMapping (RTToneMappingPass.cpp:38-87); GI.minLuminance is passed
    as the kernel's Lwhite (:60)."""

    def update(self):
        pl = self.m_renderPipeline
        if not pl.GI.useTonemapping or pl.fetch("NextFrameImageCL") is None:
            return
        src = pl.fetch("NextFrameImageName") or "NextFrameImageCL"
        _check(_lib().ptb_tonemap(pl.scene._h, src.encode(), float(pl.GI.minLuminance)))
        pl.put("NextFrameImageCL", pl.device_buffer("TonemappedImageCL")[0])
        pl.put("NextFrameImageName", "TonemappedImageCL")


def write_pfm(path, image, width, height):
    """Headless stand-in for RTDisplayPass: dump a float4[P] buffer as a little-endian RGB PFM."""
    rgb = np.asarray(image, np.float32).reshape(height, width, -1)[::-1, :, :3]
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (width, height))
        f.write(np.ascontiguousarray(rgb, "<f4").tobytes())


def write_png(path, image, width, height):
    """Engine::takeScreenshot (source/engine/Engine.cpp:180-204) without the GL framebuffer: the float4[P] image the
    display pass would draw (normally "TonemappedImageCL"), converted as an RGBA8 framebuffer converts it -- clamp to
    [0, 1], x 255, round to nearest; NaN -> 0 -- and saved as an 8-bit RGBA PNG, rows top to bottom as write_pfm
    orders them.  Alpha is written opaque (the reference reads back the window's alpha, which the GUI leaves at 1).
    Encoder: zlib + the three mandatory chunks, no image library needed."""
    import struct
    import zlib
    img = np.asarray(image, np.float32).reshape(height, width, -1)
    rgba = np.empty((height, width, 4), np.uint8)
    rgb = np.nan_to_num(img[:, :, :3], nan=0.0, posinf=1.0, neginf=0.0)
    rgba[:, :, :3] = np.rint(np.clip(rgb, 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
    rgba[:, :, 3] = 255
    raw = np.concatenate([np.zeros((height, 1), np.uint8), rgba.reshape(height, width * 4)], axis=1).tobytes()

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", width, height, 8, 6, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw, 6)))
        f.write(chunk(b"IEND", b""))


def make_pt_pipeline(api, scene, width, height, max_depth, fused=True, use_taa=True, refit=False, acc_type="fatbvh"):
    """PathTracingApp::startPathTracing's PT pipeline (:248-255) without denoise / tonemap / display."""
    rts = RTScene(api, scene, refit=refit, acc_type=acc_type)
    gi = GISettings()
    gi.imageResolution = (width, height)
    gi.maxDepth = max_depth
    gi.useTAA = use_taa
    pl = RenderPipeline(rts, gi)
    if fused:
        pl.addRenderPasses(RTPathTracingPass(True), RTReconstructionPass())
    else:
        pl.addRenderPasses(RTPrimaryRaysPass(), RTPathTracingPass(False), RTReconstructionPass())
    pl.addRenderPasses(RTDenoisePass(), RTToneMappingPass())      # both off by default, as in the GUI
    return pl


def smoke_check(api, scene, width=64, height=64, depth=3):
    """One small frame of the fused wavefront on the current device; finite, non-black, and the
    stage-wise dispatch over the reference layouts reproduces it bit for bit."""
    out = []
    for fused in (True, False):
        a = api if fused else rr.IntersectionApi.Create(0)
        pl = make_pt_pipeline(a, scene, width, height, depth, fused=fused)
        pl.update()
        out.append(pl.read_buffer("RadianceBufferCL").reshape(-1, 4).copy())
        pl.scene.destroy()
        if not fused:
            rr.IntersectionApi.Delete(a)
    assert np.isfinite(out[0]).all() and out[0][:, :3].sum() > 0.0
    assert np.array_equal(out[0].view(np.uint32)[:, :3], out[1].view(np.uint32)[:, :3])
    return out[0]


def make_bdpt_pipeline(api, scene, width, height, max_depth, use_taa=True, acc_type="fatbvh"):
    """PathTracingApp::startPathTracing's BDPT pipeline (:256-263): RTBDPTPass + RTReconstructionPass."""
    rts = RTScene(api, scene, acc_type=acc_type)
    gi = GISettings()
    gi.imageResolution = (width, height)
    gi.maxDepth = max_depth
    gi.useTAA = use_taa
    pl = RenderPipeline(rts, gi)
    pl.addRenderPasses(RTBDPTPass(), RTReconstructionPass(), RTDenoisePass(), RTToneMappingPass())
    return pl
