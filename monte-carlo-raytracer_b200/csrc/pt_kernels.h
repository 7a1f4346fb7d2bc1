// pt_kernels.h -- device state and host-callable launchers of the path-tracer kernels (pt_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

namespace mcrt {

struct SceneDev;  // opaque here; defined in pt_kernels.cu as mcrt::Scene

enum { PT_NUM_BUCKETS = 8, PT_CNT_STRIDE = 32 };   // coherence buckets; uint32 counters per half frame

// Everything a kernel needs, passed by value.  Per-pixel arrays are full-frame (P = width*height)
// and indexed by the GLOBAL pixel index; queues hold pixel indices of live paths.
struct PtState {
    // scene (device pointers, reference layouts)
    const void* shapes;
    const uint32_t* indices;
    const float4* positions;
    const float2* uvs;
    const float4* normals;
    const void* textures2D;
    const unsigned char* texData2D;
    const void* lights;
    const void* materials;
    const void* camera;
    const uint32_t* sobol;   // Sobol generator matrices when that sampler is selected, else NULL
    const float2* material_alpha;   // roughnessToAlpha of every material (launch_material_alpha), may be NULL
    const float4* tri_shade;        // per-triangle shading records (launch_build_tri_shade), 9 float4 each, may be NULL
    int numLights;
    int numShapes, numMaterials, numTextures;   // table sizes (the sorted shade kernel stages the tables in shared memory)
    int width, height;
    // acceleration structure
    const uint4* nodes;
    const float4* tris;
    const float4* instances;   // two-level world (rrb_commit with instances): the shape table, 4 float4 per shape; else NULL
    int numNodes;         // nodes in the tree; the first MCRT_TOP_NODES are staged in shared memory by the traversal CTAs
    int stageTop;         // 0 = every node from global memory (experiment switch MCRT_STAGE_TOP=0)
    int coopIters;        // > 0: cooperative tail of the traversal kernels (pt_wavefront.cuh): a ray still walked this many
                          // rounds after the queue ran dry goes to the straggler list; 0 = off.  Flat trees only.
    unsigned int* strag;  // straggler list of the running traversal launch, 2 entries per pixel (or NULL)
    // SoA wavefront state
    float4* ray_o;        // xyz origin, w tmax
    float4* ray_d;        // xyz direction
    int4* hit;            // shape id, prim id, u bits, v bits
    float4* throughput;   // rgb, w = previous BSDF flags (int bits)
    float4* radiance;     // "RadianceBufferCL": this frame's per-pixel radiance
    float4* direct_weight;  // opt-in MIS only (pt_shade.cuh): path weight for light found along the current ray
    uint32_t* queue[2];   // ping-pong live-path queues (pixel indices)
    float4* sh_o;         // shadow queue: xyz origin, w tmax
    float4* sh_d;         // xyz direction, w = pixel index (int bits)
    float4* sh_L;         // rgb = throughput * L (unshadowed direct light)
    unsigned int* cnt;    // [0],[1] queue counts, [3],[4] work counters, [5],[6] shadow counts by bounce parity,
                          // [8..15] coherence-bucket counts, [16..24] their prefix sums (PT_CNT_STRIDE per half frame)
    uint32_t* bucket;     // coherence staging: PT_NUM_BUCKETS regions per half frame (pt_kernels.cu)
    float4* vtx;          // split shade stage (ptb_set_shade_split): 7 planes of vtx_stride vertex records (pt_kernels.cu)
    int vtx_stride;       // = pixels of the frame
    int shade_split;      // k_shade_surface + k_shade_scatter instead of k_shade (reference estimator only): 0 never,
                          // 1 on wavefronts of at least 393,216 entries, 2 always
    int estimator;        // PT_EST_* flags of the opt-in estimator switches (pt_shade.cuh); 0 = the reference's estimator
    int coherence;        // 0 = queues in push order, 1 = bounces >= 1 bucketed by sampler decisions, 2 = bounce 0 too
    unsigned long long* stats;  // [0] closest rays, [1] shadow rays, [2] shaded vertices, [3] samples
    // Diagnostic record of the fused wavefront (ptb_set_path_trace; NULL = off): one int4 per (bounce, pixel) at
    // [bounce * P + pixel] = {shape id, prim id, flags, shadow ray occluded}.  flags: bit 0 the vertex was shaded,
    // bit 1 a shadow ray was traced for it, bit 2 the path ended on an emitter, bit 3 a next ray was spawned,
    // bits 8.. the sampled BSDF flags.  The parity tests use it to show WHERE a pixel that leaves the tolerance
    // diverged from the oracle (tests/parity_util.py).
    int4* path_trace;
    // reference-layout (AoS) stage buffers
    void* aos_rays;        // RTRay[P]
    void* aos_isect;       // Intersection[P]
    void* aos_shadow_rays; // RTRay[P]
    int* aos_occlusion;    // int[P]
    float4* aos_temp_radiance;
    void* aos_throughput;  // RTThroughput[P]
    // accumulation
    float4* weighted_radiance;
    float* filter_weights;
    float4* image;         // "NextFrameImageCL"
};

// Optional per-stage timing: every mark() records a pooled CUDA event on the launching stream; the
// caller reads the gaps after a synchronise (ptb_profile_collect).  Marks per bounce: before the
// closest-hit trace, before shade, before the shadow trace, after the shadow trace.
struct StageTimer {
    // = ptb_stage.  SHADE: the single shade kernel, or the surface kernel of the split stage; SHADE_SCATTER: its scatter kernel
    enum { NONE = -1, TRACE_CLOSEST = 0, TRACE_MIXED = 1, TRACE_SHADOW = 2, SHADE = 3, SHADE_SCATTER = 4, NUM_STAGES = 5 };
    // one sequence of marks per stream ("lane") the frame is issued on; intervals are taken between consecutive
    // marks of a lane, so kernels of the two lanes that overlap on the device are each timed as they ran
    struct Lane {
        std::vector<cudaEvent_t> events;   // pool, grown on demand
        std::vector<int> tags;             // what ran between the previous mark and this one
        size_t used = 0;
    };
    Lane lanes[2];
    void mark(int lane, cudaStream_t stream, int tag = NONE) {
        Lane& l = lanes[lane];
        if (l.used == l.events.size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            l.events.push_back(e);
            l.tags.push_back(NONE);
        }
        l.tags[l.used] = tag;
        cudaEventRecord(l.events[l.used++], stream);
    }
    void reset() { lanes[0].used = lanes[1].used = 0; }
    void destroy() {
        for (Lane& l : lanes) {
            for (cudaEvent_t e : l.events) cudaEventDestroy(e);
            l.events.clear();
            l.tags.clear();
            l.used = 0;
        }
    }
};

// Second stream of a frame (owned by the context).  With it the frame's pixels are split in two halves whose
// wavefronts are issued on the two streams, so that the drain-out tail of every kernel of one half is filled by
// the other half's kernels (results are per-pixel and unchanged).
struct PtOverlap {
    cudaStream_t stream;
    cudaEvent_t fork, join, stagger;
};

// One frame of the fused wavefront path tracer.  `pixels` (device, may be NULL = identity) lists the
// n_pixels global pixel indices this context owns.  Returns the number of kernels launched.
int launch_pt_sample(const PtState& s, const uint32_t* pixels, int n_pixels, int frame, int max_depth,
                     cudaStream_t stream, cudaError_t* err, StageTimer* timer = nullptr,
                     const PtOverlap* overlap = nullptr);

// Stage-wise, reference layouts.
cudaError_t launch_generate_primary_rays(const PtState& s, cudaStream_t stream);
cudaError_t launch_pt_shade_aos(const PtState& s, int frame, int max_depth, int bounce, cudaStream_t stream);
cudaError_t launch_pt_shadow_resolve_aos(const PtState& s, int bounce, cudaStream_t stream);

// alpha[i] = (roughnessToAlpha(materials[i].roughness.x), roughnessToAlpha(materials[i].roughness.y))
cudaError_t launch_material_alpha(const void* materials, float2* alpha, int num_materials, cudaStream_t stream);

// records[9 * t ..] = the triangle-only half of computeSurfaceInteraction for scene triangle t (index-array order);
// tri_shape[t] = its RTShape.  Re-run after the shapes array changes.
cudaError_t launch_build_tri_shade(const PtState& s, const int* tri_shape, float4* records, int n_tris, cudaStream_t stream);
cudaError_t launch_selftest_libm(int fn, const float* in, float* out, int n, cudaStream_t stream);
cudaError_t launch_selftest_fdiv(const float* a, const float* b, float* shortcut, float* plain, int n, cudaStream_t stream);

struct FilterProps {
    int filter_type;
    float radius[2];
    float mitchell_b, mitchell_c, lanczos_sinc_tau, gaussian_alpha, gaussian_exp_x, gaussian_exp_y;
    float pixel_offset[2];
};
cudaError_t launch_accumulate(const PtState& s, const FilterProps& f, int frame, const uint32_t* pixels, int n_pixels,
                              cudaStream_t stream);

// the same for a BDPT frame of tile partition `part` of `num_parts`: radiance is accumulated over the whole frame
// (splats land anywhere), the filter weight only on the pixels the partition owns
cudaError_t launch_accumulate_splat(const PtState& s, const FilterProps& f, int frame, int part, int num_parts,
                                    cudaStream_t stream);
// image = weighted / weights over n pixels (the merge step of the multi-GPU path, ptb_merge_accumulation)
cudaError_t launch_divide_accumulation(const float4* weighted, const float* weights, float4* image, int n, cudaStream_t stream);

// post-processing (postfx_kernels.cu)
cudaError_t launch_bilateral_denoise(int width, int height, int radius, float sigmaSpatial, float sigmaRange,
                                     const float4* src, float4* dst, cudaStream_t stream);
cudaError_t launch_reinhard_tonemap(int width, int height, float Lwhite, const float4* src, float4* dst,
                                    cudaStream_t stream);

}  // namespace mcrt
