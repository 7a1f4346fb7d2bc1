#!/usr/bin/env python
"""bench.py -- paths/s and Mrays/s of the wavefront path tracer on the Sponza-class 1080p workload
(BASELINE.json configs[2]: 262,144 triangles, textured + normal-mapped uber material, point +
directional + area light, regular PT 1920x1080, depth 5), one step = one sample per pixel.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's kernels on the host cores

Own arm, per rank: the scene is resident in HBM; a step is raygen -> (closest-hit trace -> shade/NEE
-> shadow trace + radiance add) x depth -> accumulation, launched through the C ABI (include/ptb.h).
N > 1 partitions by SAMPLE INDEX (rank r renders sample step*N + r of every pixel; weak scaling) and
merges with one NCCL reduce of the float4 accumulation buffers to rank 0 per step, behind the C ABI
(ptb_merge_accumulation).
`value` is timed with CUDA events on the launching stream (max over ranks); `e2e` drives the same
frame through the RenderPipeline mirror (camera built and uploaded on the host every step) and reads
the merged image back into pinned host memory every step.

The reference arm (and the `cpu_baseline` leg of the own arm) executes oracle/_ref -- the
reference's own OpenCL-C kernels and Radeon Rays BVH2 builder compiled for the host -- and is the
only place this file touches oracle/.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = "sponza_class_262144tri_1920x1080_pt_depth5"
METRIC, UNIT = "paths_per_sec", "paths/s"
# SURVEY 8(d) algorithmic bytes at the reference's buffer layouts
BYTES_CLOSEST_PER_RAY = 80       # 48 B RTRay read + 32 B RTIntersection write
BYTES_SHADOW_PER_RAY = 52        # 48 B RTRay read + 4 B occlusion write
BYTES_SHADE_PER_VERTEX = 948     # PathTracing + ShadowPass stage traffic per active pixel-bounce


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--depth", type=int, default=5)
    ap.add_argument("--tex", type=int, default=1024, help="texture edge length (1024 = the named workload)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--halves", type=int, default=0, choices=[0, 1, 2],
                    help="how a frame is issued (ptb_set_overlap): 0 = 1 wavefront when frames are in flight, else 2 half-frame wavefronts")
    ap.add_argument("--no-extras", action="store_true", help="skip the sustained loop, the L2 measurement and traversal T")
    ap.add_argument("--sustain-seconds", type=float, default=2.0)
    ap.add_argument("--t-res", type=int, default=4096, help="traversal T: rays per set = t_res^2 (SURVEY 8d: 2^24)")
    ap.add_argument("--frames-in-flight", type=int, default=2,
                    help="consecutive samples rendered concurrently on side streams (ptb_set_frames_in_flight); results "
                         "are bit-identical.  Measured on this workload (profiles/r01_c5_tiles.md): 1 -> 3.87 ms per step "
                         "device-resident / 3.90 end to end, 2 -> 3.65 / 3.69, 3 -> 3.66 / 3.83")
    return ap.parse_args()


def config_of(args, n_gpus):
    return {"workload": WORKLOAD, "triangles": 262144, "width": args.width, "height": args.height,
            "max_depth": args.depth, "spp_per_step": 1, "textures": "74 x %d^2 RGBA8" % args.tex,
            "lights": "point+directional+area", "partition": "sample-index x%d, 1 ncclReduce to rank 0 per step (ptb_merge_accumulation)" % n_gpus if n_gpus > 1 else "none",
            "frames_in_flight": args.frames_in_flight,
            "l2": "no flush: per-step working set (pixel state + textures) exceeds the 126 MB L2"}


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        fd, self.path = tempfile.mkstemp(suffix=".csv")
        os.close(fd)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows, reasons = [], set()
        with open(self.path) as f:
            for line in f:
                p = [x.strip() for x in line.split(",")]
                if len(p) < 7:
                    continue
                try:
                    rows.append((float(p[0]), float(p[1]), float(p[2])))
                except ValueError:
                    continue
                for name, v in zip(self.NAMES, p[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        os.unlink(self.path)
        if rows:
            pmax = max(r[2] for r in rows)
            load = [r for r in rows if r[2] >= 0.6 * pmax] or rows      # samples taken while kernels ran
            out.update(sm_mhz=float(np.median([r[0] for r in load])), sm_max_mhz=float(max(r[1] for r in rows)),
                       reasons=sorted(reasons), samples=len(rows), samples_under_load=len(load), power_w_max=pmax)
        return out


# ----------------------------------------------------------------------------- reference on the host
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_sample_size(args):
    """The host arm renders the frame `config` states, whole (r01 sub-sampled it 3 x 3; a ninth of the pixels is
    friendlier to the CPU's caches than the workload the line names)."""
    return args.width, args.height


class HostReference:
    """The reference's kernels (oracle/_ref) on the host cores: BVH2 build + the pass loops."""

    def __init__(self, scene, args):
        from oracle import ref
        self.fast = ref.available(True)
        if not ref.available(self.fast):
            raise RuntimeError("oracle/_ref is not built (make -C oracle ref needs /root/reference)")
        from mcrt_b200 import camera
        self.camera_mod = camera
        self.scene, self.depth = scene, args.depth
        self.w, self.h = cpu_sample_size(args)
        self.oracle = ref.RefOracle(self.fast)
        # every core this process may run on, set explicitly: torchrun exports OMP_NUM_THREADS=1 to its ranks
        self.threads = self.oracle.set_threads(host_threads())
        t = time.perf_counter()
        self.nodes = self.oracle.build_scene_bvh(scene)
        self.build_s = time.perf_counter() - t
        self.tracer = ref.RefPathTracer(self.oracle, scene, self.nodes, self.w, self.h)
        self.frame = 0

    def step(self):
        cam = self.camera_mod.make_camera(self.scene.camera, self.w, self.h, self.camera_mod.taa_pixel_offset(self.frame))
        self.tracer.render_sample(cam, self.frame, self.depth)
        self.frame += 1

    def rays(self):
        return self.tracer.rays_traced + self.tracer.shadow_rays_traced

    def describe(self):
        return ("the whole %dx%d frame, 1 spp per step, reference .cl kernels + RR BVH2 compiled %s, OpenMP on %d threads "
                "(counted inside a parallel region)" % (self.w, self.h, "-O3 AVX2" if self.fast else "-O2 strict fp32",
                                                        self.threads))


def run_cpu_baseline(scene, args, seconds):
    hr = HostReference(scene, args)
    hr.step()                                   # warm caches / page in
    r0, frames, t0 = hr.rays(), 0, time.perf_counter()
    while frames < 2 or time.perf_counter() - t0 < seconds:
        hr.step()
        frames += 1
        if frames >= 64:
            break
    dt = time.perf_counter() - t0
    paths = frames * hr.w * hr.h
    return {"value": paths / dt, "unit": UNIT, "cores": hr.threads, "kind": "reference",
            "sample": hr.describe() + ", %d steps in %.1f s" % (frames, dt),
            "mrays_per_s": (hr.rays() - r0) / dt / 1e6, "bvh_build_s": hr.build_s}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0                                # the host cores are one pool: rank 0 alone measures
    os.environ["OMP_NUM_THREADS"] = str(host_threads())     # before libgomp loads; ref_set_threads repeats it
    import mcrt_b200  # noqa: F401
    from mcrt_b200 import scene_library
    scene = scene_library.sponza_class(args.tex)
    hr = HostReference(scene, args)
    for _ in range(max(args.warmup, 1)):
        hr.step()
    r0, t0 = hr.rays(), time.perf_counter()
    for _ in range(args.steps):
        hr.step()
    dt = time.perf_counter() - t0
    value = args.steps * hr.w * hr.h / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(args, args.gpus), "mrays_per_s": (hr.rays() - r0) / dt / 1e6,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": hr.threads, "kind": "reference",
                             "sample": hr.describe(), "bvh_build_s": hr.build_s},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


# ----------------------------------------------------------------------------- this repo's CUDA path
class DevView:
    """__cuda_array_interface__ over a raw device pointer, so torch can address a ptb buffer."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def own_arm(args):
    import torch
    import torch.distributed as dist
    import mcrt_b200  # noqa: F401
    from mcrt_b200 import rr, pt, camera, scene_library, multi_gpu

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    scene = scene_library.sponza_class(args.tex)
    W, H, depth, P = args.width, args.height, args.depth, args.width * args.height
    stream = torch.cuda.Stream(device=dev)
    api = rr.IntersectionApi.Create(local, cuda_stream=stream.cuda_stream)
    acc_type = os.environ.get("MCRT_ACC_TYPE", "fatbvh")      # hlbvh = the device builder (lbvh_build.cu)
    t0 = time.perf_counter()
    pl = pt.make_pt_pipeline(api, scene, W, H, depth, acc_type=acc_type)          # attaches shapes, Commit() builds the BVH, uploads
    setup_s = time.perf_counter() - t0
    stats = api.GetStats()
    pl.ensure_size()
    fif = max(1, min(4, args.frames_in_flight))
    pl.set_frames_in_flight(fif)
    # with frames in flight a frame is issued as ONE wavefront: whole frames overlap each other, and that fills the
    # kernel tails better than two half-frame wavefronts of the same frame do (profiles/r01_frames_in_flight.md)
    timed_halves = args.halves or (1 if fif > 1 else 2)
    pl.set_overlap(timed_halves)
    pt_pass = pl.passes[0]
    lib, h = pt._lib(), pl.scene._h
    fprops = pl.GI.filter_props()

    # N > 1: the one collective of the path, behind the C ABI (ptb_merge_accumulation): ncclReduce of the 20 B/pixel
    # accumulation buffer to rank 0 on the context's stream + the division there; only rank 0 reads the image back
    comm = multi_gpu.create_comm(pl, rank, world) if world > 1 else None
    host_imgs = [torch.empty(4 * P, dtype=torch.float32, pin_memory=True) for _ in range(2)]   # double-buffered
    host_nps = [t.numpy() for t in host_imgs]
    host_np = host_nps[0]

    def cam_for(frame):
        off = camera.taa_pixel_offset(frame, pl.GI.filterRadius, True)
        return camera.make_camera(scene.camera, W, H, off), off

    def merge():
        """One ncclReduce of the accumulation buffers to rank 0 per step (progressive display).  It runs on the
        context's stream behind this step's accumulation, while the next frames render on their side streams."""
        pt._check(lib.ptb_merge_accumulation(h, comm, 0))

    def merge_flush():
        pass

    def device_step(frame, cam, off):
        """C-ABI calls only; the camera struct was built before the timed region."""
        fprops.pixel_offset[0], fprops.pixel_offset[1] = off
        pt._check(lib.ptb_set_camera(h, cam.ctypes.data_as(ctypes.c_void_p)))
        pt._check(lib.ptb_pt_render_sample(h, frame, depth))
        pt._check(lib.ptb_accumulate(h, frame, ctypes.byref(fprops)))
        if world > 1:
            merge()

    def e2e_step(frame, slot=0):
        """The call a user makes: pipeline.update() (host camera build + upload) and the image of every frame on
        the host -- read back through the pipelined ptb_read_buffer_async (device snapshot, then a copy into pinned
        memory behind the next frame's kernels; the last copy is waited for inside the timed region)."""
        pt_pass.m_frameIndex = frame
        pl.update()
        if world > 1:
            merge()
        if rank == 0:                   # the merged image exists on the root only
            pl.read_buffer_async("NextFrameImageCL", host_nps[slot & 1])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    K, Wm = args.steps, max(args.warmup, 3)
    frames = multi_gpu.frames_of_rank(rank, world, Wm + K)
    cams = [cam_for(f) for f in frames]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    with torch.cuda.stream(stream):
        # ---------------- device-resident timing
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()                     # runs through warm-up, both timed loops; idle samples are dropped
        for s in range(Wm):
            device_step(frames[s], *cams[s])
        merge_flush()
        barrier()
        pl.reset_counters()
        barrier()
        ev0.record(stream)
        for s in range(Wm, Wm + K):
            device_step(frames[s], *cams[s])
        merge_flush()
        ev1.record(stream)
        barrier()
        ms = max_over_ranks(ev0.elapsed_time(ev1))
        cnt = pl.counters()

        # ---------------- per-kernel durations for the roofline: the same K steps once more with every frame issued
        # as ONE wavefront (ptb_set_overlap(1)), so that a kernel's events bracket that kernel alone; in the timed
        # region above the two half-frame wavefronts overlap and per-kernel intervals are not separable
        pl.set_overlap(1)
        pl.profile_enable(True)
        ev0.record(stream)
        for s in range(Wm, Wm + K):
            device_step(frames[s], *cams[s])
        merge_flush()
        ev1.record(stream)
        barrier()
        ms_single = max_over_ranks(ev0.elapsed_time(ev1))
        kern = pl.profile_collect_kernels()
        split = kern["k_shade_scatter"][1] > 0       # the shade stage ran as surface + scatter kernels
        prof = {k: kern[k] for k in pl.STAGES}
        if split:                                     # the stage as a whole: one "launch" = one bounce = two kernels
            prof["k_shade"] = (kern["k_shade"][0] + kern["k_shade_scatter"][0], kern["k_shade"][1])
        pl.profile_enable(False)
        pl.set_overlap(timed_halves)

        # ---------------- end to end through the pipeline mirror, image read back every step
        pt._check(lib.ptb_reset_accumulation(h))
        for s in range(Wm):
            e2e_step(frames[s], s)
        merge_flush()
        pl.read_wait()
        barrier()
        ev0.record(stream)
        for s in range(Wm, Wm + K):
            e2e_step(frames[s], s)
        merge_flush()
        pl.read_wait()                      # the last frame's image is on the host before the clock stops
        ev1.record(stream)
        barrier()
        ms_e2e = max_over_ranks(ev0.elapsed_time(ev1))
        clocks = sampler.stop() if rank == 0 else None
        img_mean = float(host_nps[(Wm + K - 1) & 1].reshape(-1, 4)[:, :3].mean())

    paths_total = sum_over_ranks(float(cnt["samples"]))
    closest_total = sum_over_ranks(float(cnt["closest_rays"]))
    shadow_total = sum_over_ranks(float(cnt["shadow_rays"]))
    launches_total = int(sum_over_ranks(float(cnt["kernel_launches"])))
    value = paths_total / (ms * 1e-3)
    e2e_value = paths_total / (ms_e2e * 1e-3)

    if rank == 0:
        peaks, peak_src = {}, "fallback (B200_PROFILING.md)"
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
            peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # Per kernel: closest-hit rays of bounce 0 have a launch of their own; those of bounce b share one with the
        # shadow rays of bounce b - 1 (k_trace_mixed); the last shadow set is traced alone.  Algorithmic bytes of
        # a traversal launch = 80 B per closest-hit ray + 52 B per shadow ray it served.
        first_closest = float(cnt["samples"])                   # one primary ray per sample
        trace_bytes = cnt["closest_rays"] * BYTES_CLOSEST_PER_RAY + cnt["shadow_rays"] * BYTES_SHADOW_PER_RAY
        t_closest, t_mixed, t_shadow = prof["k_trace_closest"][0], prof["k_trace_mixed"][0], prof["k_trace_shadow"][0]
        stages = {
            "k_trace_closest": (t_closest, first_closest * BYTES_CLOSEST_PER_RAY, prof["k_trace_closest"][1]),
            "k_trace_mixed": (t_mixed, None, prof["k_trace_mixed"][1]),
            "k_trace_shadow": (t_shadow, None, prof["k_trace_shadow"][1]),
            "k_shade": (prof["k_shade"][0], cnt["shaded_vertices"] * BYTES_SHADE_PER_VERTEX, prof["k_shade"][1]),
        }
        # the split of the remaining rays between the mixed launches and the last shadow launch is not counted on
        # the device; apportion their bytes by time (same kernel body, same ray mix to first order)
        rest = trace_bytes - first_closest * BYTES_CLOSEST_PER_RAY
        for k, t in (("k_trace_mixed", t_mixed), ("k_trace_shadow", t_shadow)):
            stages[k] = (t, rest * t / max(t_mixed + t_shadow, 1e-12), stages[k][2])
        top = max(stages, key=lambda k: stages[k][0])
        top_ms, top_bytes, nb = stages[top]
        nb = max(nb, 1)
        achieved = top_bytes / max(top_ms * 1e-3, 1e-12) / 1e9
        traffic = None      # measured per build: run_extras() fills it from profiles/kernel_counters.json
        shade_name = "k_shade_surface+k_shade_scatter" if split else "k_shade"
        roofline = {"bound": "hbm", "kernel": shade_name if top == "k_shade" else top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                    "launches": nb, "avg_launch_ms": top_ms / nb, "algorithmic_bytes_per_launch": top_bytes / nb,
                    "stage_ms_per_step": dict({k: v[0] / K for k, v in stages.items()},
                                              **({"k_shade_surface": kern["k_shade"][0] / K, "k_shade_scatter": kern["k_shade_scatter"][0] / K}
                                                 if split else {})),
                    "stage_share_of_step": {k: v[0] / max(ms_single, 1e-9) for k, v in stages.items()},
                    "stage_timing": "separate %d-step loop, one frame at a time, one wavefront per frame (%.3f ms/step); the timed region %s"
                                    % (K, ms_single / K, ("keeps %d frames in flight on side streams" % fif) if fif > 1 else
                                       "overlaps two half-frame wavefronts on two streams"),
                    "trace_rays_per_s": (cnt["closest_rays"] + cnt["shadow_rays"]) /
                                        max((t_closest + t_mixed + t_shadow) * 1e-3, 1e-12)}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = run_cpu_baseline(scene, args, args.cpu_seconds)
        # ---------------- outside `value`, N = 1 only: sustained loop, L2 / issue rooflines, traversal micro-benchmark T
        extras = {}
        if world == 1 and not args.no_extras:
            extras = run_extras(args, torch, stream, dev, local, device_step, frames, cams, Wm, K, pl, api, scene,
                                roofline, top, stages, clocks)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config_of(args, world),
                "mrays_per_s": (closest_total + shadow_total) / (ms * 1e-3) / 1e6,
                "rays_per_path": (closest_total + shadow_total) / max(paths_total, 1.0),
                "roofline": roofline, "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": 176,
                        "d2h_bytes_per_step": int(host_np.nbytes), "image_mean": img_mean},
                "gpu_launches": launches_total, "clocks": clocks, **extras,
                "bvh": {"build_and_upload_s": setup_s, "nodes": int(stats.get("num_nodes", 0)) if isinstance(stats, dict) else None,
                        "acc_type": acc_type, "on_device": int(stats.get("last_build_on_device", 0)) if isinstance(stats, dict) else None,
                        "commit_s": float(stats.get("build_seconds", 0.0)) if isinstance(stats, dict) else None}}
        emit(line)
    if comm is not None:
        pl.synchronize()
        multi_gpu.destroy_comm(comm)
    pl.scene.destroy()
    rr.IntersectionApi.Delete(api)
    if world > 1:
        dist.destroy_process_group()
    return 0


def build_fingerprint():
    from mcrt_b200 import build
    return build._fingerprint()[:16]


def load_sidecar():
    """Per-kernel hardware counters of ONE step of this workload, measured by tools/measure_counters.py under ncu
    (dram bytes, L2 bytes, warp instructions per launch).  They are tied to the build they were measured on: a
    sidecar whose fingerprint differs from the library's is refused (None), never silently reused."""
    path = os.path.join(ROOT, "profiles", "kernel_counters.json")
    try:
        with open(path) as f:
            d = json.load(f)
    except OSError:
        return None
    return d if d.get("build_fingerprint") == build_fingerprint() else None


def measure_l2_copy_gbs(torch, dev, stream):
    """Device-to-device copy of a buffer that fits the L2 twice over (2 x 24 MB), read + write bytes per second:
    the denominator of roofline.l2, obtained the way MEASURED_PEAKS.json's hbm_gbs was."""
    n = 24 * 1024 * 1024 // 4
    a = torch.empty(n, dtype=torch.float32, device=dev).normal_()
    b = torch.empty_like(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(5):
        for _ in range(3):
            b.copy_(a)
        e0.record(stream)
        for _ in range(40):
            b.copy_(a)
        e1.record(stream)
        e1.synchronize()
        best = max(best, 40 * 2 * n * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    return best


def run_extras(args, torch, stream, dev, local, device_step, frames, cams, Wm, K, pl, api, scene, roofline, top, stages,
               clocks):
    from mcrt_b200 import rr, scene_library, traversal_bench
    out = {}
    # (1) sustained: the same device-resident step for >= sustain_seconds, clocks sampled while it runs
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pl.reset_counters()
    n, t0 = 0, time.perf_counter()
    ev0.record(stream)
    while time.perf_counter() - t0 < args.sustain_seconds:
        for s in range(Wm, Wm + K):
            device_step(frames[s], *cams[s])
        n += K
        pl.synchronize()
    ev1.record(stream)
    ev1.synchronize()
    ms = ev0.elapsed_time(ev1)
    c = sampler.stop()
    out["sustained"] = {"seconds": ms * 1e-3, "steps": n, "value": pl.counters()["samples"] / (ms * 1e-3), "unit": UNIT,
                        "sm_mhz": c.get("sm_mhz"), "power_w_max": c.get("power_w_max"), "reasons": c.get("reasons")}
    # (2) the non-HBM rooflines of the dominant kernel: issue slots and L2 bytes, from the measured per-launch counters
    side = load_sidecar()
    l2_peak = measure_l2_copy_gbs(torch, dev, stream)
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    roofline["l2_copy_gbs_measured"] = l2_peak
    if side and top in side.get("kernels", {}):
        k = side["kernels"][top]
        t_launch = roofline["avg_launch_ms"] * 1e-3
        roofline["traffic"] = k["dram_bytes_per_launch"]
        roofline["traffic_source"] = "ncu, %s, build %s" % (side.get("when"), side.get("build_fingerprint"))
        issue_peak = sms * 4 * mhz * 1e6
        roofline["issue"] = {"warp_insts_per_launch": k["warp_insts_per_launch"], "achieved_per_s": k["warp_insts_per_launch"] / t_launch,
                             "peak_per_s": issue_peak, "frac": k["warp_insts_per_launch"] / t_launch / issue_peak,
                             "threads_per_inst": k.get("threads_per_inst"), "sm_mhz": mhz}
        roofline["l2"] = {"bytes_per_launch": k["l2_bytes_per_launch"], "achieved": k["l2_bytes_per_launch"] / t_launch / 1e9,
                          "peak": l2_peak, "unit": "GB/s", "frac": k["l2_bytes_per_launch"] / t_launch / 1e9 / l2_peak}
    else:
        roofline["traffic"] = None
        roofline["traffic_source"] = "no counters for this build (tools/measure_counters.py writes profiles/kernel_counters.json)"
    # (3) traversal micro-benchmark T: the 1 M-triangle knot, and this workload's closed atrium
    t_scene = scene_library.traversal_bench_scene(707)
    api_t = rr.IntersectionApi.Create(local, cuda_stream=stream.cuda_stream)
    rr.attach_scene(api_t, t_scene)
    api_t.Commit()
    out["traversal_T"] = {"rays_per_set": args.t_res * args.t_res,
                          "knot_1M": traversal_bench.run(api_t, t_scene, stream, args.t_res),
                          "sponza_class": traversal_bench.run(api, scene, stream, args.t_res),
                          "bytes_per_ray": {"closest": BYTES_CLOSEST_PER_RAY, "anyhit": BYTES_SHADOW_PER_RAY}}
    rr.IntersectionApi.Delete(api_t)
    return out


_RESULT_FD = None


def emit(line):
    """The ONE JSON line of the contract, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    # Libraries (NCCL prints its version banner, torchrun its OMP note) write to stdout; the contract is one JSON
    # line there.  Everything but that line goes to stderr.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return own_arm(args)


if __name__ == "__main__":
    sys.exit(main())
