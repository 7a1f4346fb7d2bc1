/* ptb.h -- C ABI of the B200 path-tracer stages (inside libmcrt_b200.so).
 *
 * Drop-in boundary, surface 2: what the engine's render passes call instead of compiling and
 * launching the assets/kernels .cl files through CLW (source/application/PathTracer/raytracing/renderPasses/
 * RTPrimaryRaysPass.cpp:32-134, RTPathTracingPass.cpp:40-218, RTBDPTPass.cpp:67-511,
 * RTReconstructionPass.cpp:71-123) and instead of RTScene::setSceneArgs / upload*
 * (scene/RTScene.cpp:178-197,680-809).  Pass classes mirroring the reference's are in
 * monte-carlo-raytracer_b200/host/passes.hpp; the binding a maintainer adds is in INTEGRATION.md.
 *
 * A ptb_context lives on the device and stream of the rrb_api it was created from and traces
 * against that API's committed world (rrb_commit must have run).  All functions return 0 or a
 * negative rrb_status; the message is ptb_last_error().  No CPU fallback exists.
 *
 * Buffer names (ptb_get_buffer) are the reference's blackboard keys and members:
 *   "RayBuffer"                    RTRay[P] 48 B            RTPrimaryRaysPass.cpp:64
 *   "PrimaryIntersectionBufferCL"  Intersection[P] 32 B     RTPrimaryRaysPass.cpp:65
 *   "ShadowRayBuffer"              RTRay[P]                 RTPathTracingPass.cpp:215
 *   "ShadowRayOcclusionBuffer"     int32[P]                 RTPathTracingPass.cpp:216
 *   "TempRadianceBuffer"           float4[P]                RTPathTracingPass.cpp:211
 *   "ThroughputBuffer"             RTThroughput[P] 32 B     RTPathTracingPass.cpp:212
 *   "RadianceBufferCL"             float4[P]                RTPathTracingPass.cpp:106 / RTBDPTPass.cpp:120
 *   "NextFrameImageCL"             float4[P] (was a GL-interop image2d_t, RTReconstructionPass.cpp:67)
 *   "WeightedRadianceBuffer", "FilterWeightsBuffer"  float4[P], float[P]  RTReconstructionPass.cpp:212-213
 *   "AccumulationBuffer"           the two above as the one allocation they are: float[4P] || float[P]
 *   "BDPTSampledLightBuffer"       BDPT state that persists across frames (exists after the first BDPT frame)
 *   "PathTraceBuffer"              diagnostic, see ptb_set_path_trace
 */
#ifndef PTB_H
#define PTB_H
#include <stddef.h>
#include <stdint.h>

#include "rrb.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ptb_context ptb_context;

/* Host pointers to the RTHostScene arrays (scene/RTScene.h:44-61) in the device layouts of
 * kernel_data.h:36-152,278-294.  float3 pools have a 16-byte stride.  tangents / binormals /
 * colors are bound by the reference but never read on this path and are not taken. */
typedef struct ptb_scene_desc {
    const void* shapes;          /* RTShape[num_shapes], 160 B */
    const uint32_t* indices;     /* 3 per triangle, relative to RTShape::startVertex */
    const float* positions;      /* float4 stride */
    const float* uvs;            /* float2 */
    const float* normals;        /* float4 stride */
    const void* textures;        /* TextureDesc2D[num_textures], 16 B */
    const unsigned char* texdata;
    const uint32_t* sobol;       /* 1024 x 52 direction numbers or NULL (random sampler is the default) */
    const void* lights;          /* RTLight[num_lights], 80 B */
    const void* materials;       /* RTMaterial[num_materials], 128 B */
    const void* camera;          /* unused here: see ptb_set_camera */
    int32_t num_shapes, num_textures, num_lights, num_materials;
    int64_t num_indices, num_vertices, texdata_bytes;
} ptb_scene_desc;

/* RTFilterProperties (kernel_data.h:63-80), packed for C. */
typedef struct ptb_filter_props {
    int32_t filter_type;         /* RT_BOX_FILTER..RT_LANCZOS_SINC_FILTER = 0..4 */
    float radius[2];
    float mitchell_b, mitchell_c;
    float lanczos_sinc_tau;
    float gaussian_alpha, gaussian_exp_x, gaussian_exp_y;
    float pixel_offset[2];
} ptb_filter_props;

typedef struct ptb_counters {
    int64_t closest_rays;        /* active rays given to closest-hit traversal */
    int64_t shadow_rays;         /* active rays given to any-hit traversal */
    int64_t shaded_vertices;     /* path vertices shaded */
    int64_t samples;             /* pixel-samples (paths) started */
    int64_t kernel_launches;     /* CUDA kernels launched by this context */
} ptb_counters;

const char* ptb_last_error(void);

int ptb_create(rrb_api* api, ptb_context** out);
int ptb_destroy(ptb_context* ctx);

/* RTScene::uploadShapes/uploadLights/uploadMaterials/uploadTextures + buffer creation. Blocking. */
int ptb_upload_scene(ptb_context* ctx, const ptb_scene_desc* scene);
/* PathTracerSettings::GI.imageResolution; (re)allocates every per-pixel buffer, zero-filled. */
/* Hot update of one scene array, in place and stream-ordered: `array` is "shapes", "lights" or "materials";
 * `bytes` must equal the size uploaded by ptb_upload_scene.  Replaces the whole-array WriteBuffer calls of
 * RTScene::updateDynamicEntities (scene/RTScene.cpp:356-389); the matching BVH update is rrb_commit
 * (a device refit when the rrb option "bvh.refit" is set). */
int ptb_update_scene_array(ptb_context* ctx, const char* array, const void* data, size_t bytes);
int ptb_resize(ptb_context* ctx, int width, int height);
/* Sampler variant (samplers.cl:15-17,75-84; a compile-time switch in the reference): kind 0 = Sobol,
 * 1 = random (default).  `matrices` (may be NULL to keep the table uploaded with the scene) holds
 * num_dimensions x 52 uint32 generator-matrix columns in the layout of sampling/sobol.h:31-34.  A frame
 * uses at most 5 dimensions per bounce in PT and 3 * maxDepth + 5 in BDPT. */
int ptb_set_sampler(ptb_context* ctx, int kind, const uint32_t* matrices, int num_dimensions);
/* WriteBuffer(camera) of RTPrimaryRaysPass.cpp:106-107 / RTBDPTPass.cpp:168-169: 176 bytes. */
int ptb_set_camera(ptb_context* ctx, const void* rt_pinhole_camera);
/* Multi-GPU image partition: this context renders only pixels whose 64x64 tile index t satisfies
 * t % num_parts == part (round-robin tiles); buffers stay full-frame so global pixel
 * indices seed the sampler.  Default (0, 1) = whole image. */
int ptb_set_partition(ptb_context* ctx, int part, int num_parts);
/* Host-only: the global pixel indices part `part` of `num_parts` owns (64x64 tiles, round-robin in
 * row-major tile order).  `out` may be NULL to query the count.  Needs no device. */
int ptb_partition_pixels(int width, int height, int part, int num_parts, uint32_t* out, int64_t capacity,
                         int64_t* count);

/* ---- stage-wise entry points in the reference's buffer layouts (one call = one kernel the reference launches) */
int ptb_generate_primary_rays(ptb_context* ctx);                                 /* GeneratePerspectiveRays */
int ptb_pt_shade(ptb_context* ctx, int frame, int max_depth, int bounce);       /* PathTracing */
int ptb_pt_shadow_resolve(ptb_context* ctx, int bounce);                         /* ShadowPass */

/* ---- fused wavefront: RTPrimaryRaysPass::update + RTPathTracingPass::update for one frame (one sample
 * per pixel, `max_depth` bounces) on SoA queues with stream compaction.  Result: "RadianceBufferCL". */
int ptb_pt_render_sample(ptb_context* ctx, int frame, int max_depth);
/* Same for RTBDPTPass::update. */
int ptb_bdpt_render_sample(ptb_context* ctx, int frame, int max_depth);
/* RTBDPTPass::createBuffers (RTBDPTPass.cpp:451-478) alone: allocates the zero-filled BDPT buffers for `max_depth`
 * (ptb_bdpt_render_sample does it on first use), so that a checkpoint can be restored before the first frame. */
int ptb_bdpt_reserve(ptb_context* ctx, int max_depth);
/* Per-strategy records of the BDPT frame rendered last (diagnostic; the reference keeps the same information in its
 * connection buffers: connectionRays / tempRadianceBuffer / connectionVisibilities, RTBDPTPass.cpp:451-478, written by
 * PrepareConnections, BDPT.cl:460-652).  One record per (pixel, t, s) strategy whose unweighted contribution is not
 * black -- compacted, in no particular order; N = element 6 of "BDPTCounters" (uint32[16]):
 *   "BDPTConnectionContribution" float4[N]  xyz = unweighted contribution C (the reference's tempRadianceBuffer entry
 *                                            before visibility and MIS weight), w = bits of the TARGET pixel index
 *                                            (t = 1 strategies splat to the pixel the light vertex projects to)
 *   "BDPTConnectionRayD"         float4[N]  xyz = connection ray direction, w = bits of  t | (s << 8)
 *   "BDPTConnectionRayO"         float4[N]  xyz = connection ray origin, w = its length (s = 0: no ray, zeros)
 *   "BDPTConnectionPixel"        int32[N]   the pixel whose subpaths are connected
 *   "BDPTConnectionVisibility"   int32[N]   1 = unoccluded or no ray needed (s = 0), 0 = occluded
 *   "BDPTConnectionWeight"       float[N]   the MIS weight the record was accumulated with (0 when occluded):
 *                                            radiance[target] += C * weight
 * Valid after ptb_bdpt_render_sample until the next BDPT frame or ptb_resize; capacity maxConnections * P records. */
/* How ptb_pt_render_sample issues a frame: 2 (default) = the pixel set is split in two halves whose wavefronts run
 * on two streams, so that each kernel's drain-out tail is filled by the other half's kernels;
 * 1 = one wavefront on the context's stream.  Per-sample results are identical. */
int ptb_set_overlap(ptb_context* ctx, int halves);
/* Frames in flight: with depth D > 1 (at most 4; default 1) consecutive ptb_pt_render_sample calls are issued on D
 * side streams, each with its own wavefront state and camera copy, so that the latency chain of one frame -- every
 * traversal launch ends with the few rays that graze the mesh -- is filled by the work of the following frames.
 * ptb_accumulate consumes the frames in the order they were rendered, on the context's stream; every other entry
 * point first waits for the frames in flight, so the call sequence of the render passes and every per-sample value
 * are unchanged.  "RadianceBufferCL" names the buffer of the frame rendered last.  Costs 304 B per pixel of device
 * memory per extra frame. */
int ptb_set_frames_in_flight(ptb_context* ctx, int depth);
/* Order of the live-path queues of ptb_pt_render_sample.  The reference's sampler is a pure function of (pixel, frame,
 * bounce) (samplers.cl:75-84), so which light a vertex will sample and which half of the lobe-selection variate it
 * will draw are known before its ray is traced: 1 = queues of bounces >= 1 are bucketed by that key (warps of the
 * shade kernel take the same light / lobe branches), 2 = bounce 0 as well, 3 = bucketed inside each thread block's
 * chunk of the queue only (keeps neighbouring pixels together: texture / geometry locality), 0 = push order,
 * 4 = the consumer sorts: every thread block of the shade kernel sorts its 2048-entry chunk of the queue in shared
 * memory by (light type, lobe half, material id), drops the rays that missed, and stages the scene's small tables in
 * shared memory with bulk asynchronous copies (csrc/pt_kernels.cu: shade_sorted; measured slower than 3 on the
 * benchmark scene, profiles/r02_shade.md).  Per-sample results are identical in all modes. */
#define PTB_DEFAULT_COHERENCE 0      /* push order: fastest with the split shade stage (r02q); 3 was r01's default for the single kernel */
/* Opt-in estimator switches of ptb_pt_render_sample (default 0 = the reference's estimator, PathTracing.cl:52-184,
 * which has neither): Russian roulette from the third vertex on, and power-heuristic MIS between next-event
 * estimation and BSDF sampling for triangle-mesh area lights.  Both keep the expectation of the reference's
 * estimator (MIS: for light meshes whose triangles have equal areas, lights.cl:106,122) and change its variance,
 * so they are excluded from per-sample parity; the stage-wise entry points ignore them. */
#define PTB_EST_RUSSIAN_ROULETTE 1
#define PTB_EST_NEE_MIS 2
int ptb_set_estimator(ptb_context* ctx, int flags);
int ptb_set_coherence(ptb_context* ctx, int mode);
/* The shade stage of ptb_pt_render_sample as ONE kernel per bounce (0) or as TWO (1: on wavefronts of at least
 * 393,216 entries, where it pays -- a 1080p half frame +4.4 %, a 512^2 half frame -12 %; 2: always): k_shade_surface -- surface
 * interaction, normal map, emission, material textures; drops the rays that missed; writes a 112-byte record per live
 * vertex -- and k_shade_scatter -- light sampling, BSDF evaluation, lobe sampling on dense records (csrc/pt_kernels.cu,
 * pt_shade.cuh).  Same expressions on the same values: per-sample results are identical.  Used for the reference's
 * estimator only, and only while no material binds a specular-reflection / transmission / ior / roughness texture
 * (the second kernel re-reads those properties from the material table); otherwise the single kernel runs. */
#define PTB_DEFAULT_SHADE_SPLIT 1
int ptb_set_shade_split(ptb_context* ctx, int on);
/* Cooperative tail of the traversal kernels of ptb_pt_render_sample (flat trees; no reference counterpart).  A ray that
 * is still being walked `min_iters` rounds after its launch's queue ran dry is put on a list of stragglers; every warp
 * that is out of work pulls rays off that list and walks each with all 32 lanes, lane k restricted to the k-th of 32
 * intervals of t (same origin and direction, so every t is the value the single walk computes); the warp's smallest
 * (t, shape id, prim id) is the single walk's answer, tie rule included (csrc/pt_wavefront.cuh).  It shortens the tail a
 * few slow rays put on every launch (profiles/r02_tail.md).  0 = off.  Per-sample results are identical. */
#define PTB_DEFAULT_COOPERATIVE_TAIL 0
int ptb_set_cooperative_tail(ptb_context* ctx, int min_iters);

/* Diagnostic record of ptb_pt_render_sample (no reference counterpart; parity evidence only).  With max_depth > 0 a
 * buffer "PathTraceBuffer" of int4[max_depth][P] is kept: entry [bounce * P + pixel] = {shape id, prim id, flags,
 * shadow ray occluded} of the vertex the pixel's path shaded at that bounce in the frame rendered last (all zero
 * = the path was not alive).  flags: bit 0 a hit was shaded, bit 1 a shadow ray was traced, bit 2 the path ended
 * on an emitter, bit 3 a next ray was spawned, bits 8.. the sampled BSDF flags (bxdfs.cl BSDF_*).  While it is on,
 * frames are rendered one at a time.  0 switches it off and frees the buffer; ptb_resize drops it. */
int ptb_set_path_trace(ptb_context* ctx, int max_depth);

/* Per-kernel device timing of the fused wavefront (no reference counterpart; measurement only).
 * While enabled, CUDA events are recorded on the context's stream around every traversal and shade launch;
 * collect() synchronises, returns the summed milliseconds and launch counts per kernel (indexed by ptb_stage)
 * and rearms the recorder. */
enum ptb_stage {
    PTB_STAGE_TRACE_CLOSEST = 0,   /* closest-hit rays of bounce 0 */
    PTB_STAGE_TRACE_MIXED = 1,     /* closest-hit rays of bounce b + shadow rays of bounce b - 1, one launch */
    PTB_STAGE_TRACE_SHADOW = 2,    /* shadow rays of the last bounce */
    PTB_STAGE_SHADE = 3,           /* the shade stage: k_shade, or k_shade_surface + k_shade_scatter (ptb_set_shade_split) */
    PTB_STAGE_SHADE_SCATTER = 4,   /* ptb_profile_collect_kernels only: the scatter kernel; [3] is then the surface kernel */
    PTB_NUM_STAGES = 5
};
int ptb_profile_enable(ptb_context* ctx, int on);
int ptb_profile_collect(ptb_context* ctx, double ms[4], int64_t launches[4]);
/* The same per KERNEL: with the split shade stage, entry PTB_STAGE_SHADE holds the surface kernel alone and entry
 * PTB_STAGE_SHADE_SCATTER the scatter kernel (0 for the single kernel). */
int ptb_profile_collect_kernels(ptb_context* ctx, double ms[PTB_NUM_STAGES], int64_t launches[PTB_NUM_STAGES]);

/* ReconstructionPass (reconstruction.cl:6-60): clamp, filter weight of the frame's pixel offset,
 * running weighted mean into "NextFrameImageCL"; frame 0 restarts the mean (reconstruction.cl:47-57). */
int ptb_accumulate(ptb_context* ctx, int frame, const ptb_filter_props* filter);
int ptb_reset_accumulation(ptb_context* ctx);

/* Post-processing on float4[P] buffers named as in ptb_get_buffer (the GL-interop images of the reference
 * become device buffers; the display pass itself is out of scope).
 *  ptb_denoise: BilateralDenoise (Denoise.cl:6-47, RTDenoisePass.cpp:33-75)   src -> "DenoisedImageCL";
 *               radius <= 0 or a sigma <= 0 writes nothing, like the kernel's early return.
 *  ptb_tonemap: ReinhardToneMapping (ToneMapping.cl:41-63, RTToneMappingPass.cpp:38-87; the host passes
 *               GI.minLuminance as Lwhite)                                     src -> "TonemappedImageCL". */
int ptb_denoise(ptb_context* ctx, const char* src_name, int radius, float sigma_spatial, float sigma_range);
int ptb_tonemap(ptb_context* ctx, const char* src_name, float lwhite);

/* Self-test (no reference counterpart): the shading code's division helper -- IEEE division with a branch-free
 * shortcut for zero numerators, pt_math.cuh:fdiv -- and the plain hardware division, element-wise over host arrays. */
int ptb_selftest_fdiv(ptb_context* ctx, const float* a, const float* b, float* out_shortcut, float* out_plain, int n);

/* ---- multi-GPU merge (no reference counterpart: the reference is single-device).  One process per GPU,
 * scene and BVH replicated, every rank renders its share of the samples (ptb_set_partition for image tiles, or
 * different frame indices for a sample-index split) into its own accumulation buffers; the ONE collective of the
 * path sums  sum(w L) (float4[P]) || sum(w) (float[P])  -- a single 20 B/pixel allocation, "AccumulationBuffer" --
 * over the ranks with NCCL (NVLink / NVSwitch) on the context's stream, and the division of ReconstructionPass
 * (reconstruction.cl:50-58) follows in one kernel:
 *   root >= 0  ncclReduce to `root`, in place.  The root's buffers then hold the sums over all ranks and all frames
 *              so far and "NextFrameImageCL" on the root is the merged image; every other rank's buffers are zeroed
 *              (its samples have been handed over), so rendering simply goes on and the next merge adds the new
 *              samples.  Call it every frame for a progressive display or once at the end of a batch.
 *   root < 0   ncclAllReduce into a side buffer; every rank gets the merged image, rank-local sums are kept.
 * `nccl_comm` is an ncclComm_t whose ranks all make this call (NULL = single rank: only the division runs).  Under
 * a sample-index split only the rank that renders frame 0 restarts its sums by itself (ptb_accumulate); a host that
 * restarts the image calls ptb_reset_accumulation on every rank.
 * The ptb_nccl_* helpers wrap ncclGetUniqueId / ncclCommInitRank / ncclCommDestroy for hosts that do not link
 * NCCL themselves: rank 0 makes the id, the host's launcher carries its 128 bytes to the other ranks. */
#define PTB_NCCL_UNIQUE_ID_BYTES 128
int ptb_nccl_unique_id(void* id_out);
int ptb_nccl_comm_create(ptb_context* ctx, const void* id_in, int rank, int num_ranks, void** nccl_comm_out);
int ptb_nccl_comm_destroy(void* nccl_comm);
int ptb_merge_accumulation(ptb_context* ctx, void* nccl_comm, int root);

/* Self-test (no reference counterpart): the shading code's sin / cos / log (csrc/pt_libm.h, which restate the host C
 * library's algorithms so that sampled directions are bit-identical to the oracle's), evaluated on the device
 * element-wise over a host array; fn 0 = sin, 1 = cos, 2 = log. */
int ptb_selftest_libm(ptb_context* ctx, int fn, const float* in, float* out, int n);

int ptb_get_buffer(ptb_context* ctx, const char* name, void** device_ptr, size_t* bytes);
/* Blocking host->device write of a whole named buffer: restores a checkpoint of the accumulation state
 * ("AccumulationBuffer" = sum(w L) || sum(w), RTReconstructionPass.cpp:212-213; for BDPT also "BDPTSampledLightBuffer",
 * whose records are read before they are rewritten) so that a render resumes bit for bit. */
int ptb_write_buffer(ptb_context* ctx, const char* name, const void* host_src, size_t bytes);
/* Blocking device->host read of a named buffer (the step's result for the end-to-end leg). */
int ptb_read_buffer(ptb_context* ctx, const char* name, void* host_dst, size_t bytes);
/* Pipelined read-back for a consumer that wants the image every frame (the role of RTDisplayPass): snapshots the
 * buffer on the context's stream (device-to-device), then copies the snapshot to `host_dst` (pinned memory) on a
 * copy stream while the next frames render.  Two copies may be in flight per context: `host_dst` is valid after
 * ptb_read_wait, or after the second ptb_read_buffer_async that follows (which waits for the copy issued two calls
 * before it), so a consumer rotates three host buffers -- or two, if it only looks at an image after the wait. */
int ptb_read_buffer_async(ptb_context* ctx, const char* name, void* host_dst, size_t bytes);
int ptb_read_wait(ptb_context* ctx);
int ptb_synchronize(ptb_context* ctx);
int ptb_get_counters(ptb_context* ctx, ptb_counters* out);
int ptb_reset_counters(ptb_context* ctx);

#ifdef __cplusplus
}
#endif
#endif /* PTB_H */
