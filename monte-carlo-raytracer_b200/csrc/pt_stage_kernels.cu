// pt_stage_kernels.cu -- the reference's own stage granularity in the reference's own buffer layouts
// (for pass code that keeps the reference's per-stage dispatch), and the accumulation stage.
//
//   k_generate_rays_aos   GeneratePerspectiveRays  assets/kernels/PathTracing.cl:13-35
//   k_pt_shade_aos        PathTracing              assets/kernels/PathTracing.cl:52-184
//   k_shadow_pass_aos     ShadowPass               assets/kernels/PathTracing.cl:186-217
//   k_accumulate          ReconstructionPass       assets/kernels/reconstruction.cl:6-60, filters.cl:12-69
// They share shade_vertex() with the wavefront kernels; buffers they do not write keep their
// previous contents exactly like the reference's (stale shadow rays of paths that did no NEE).
#include "pt_kernels.h"

#include "pt_shade.cuh"

namespace mcrt {

struct AosRay {           // RTRay, kernel_data.h:407-417
    float4 o, d;
    int2 extra;
    int doBackfaceCulling, padding;
};
struct AosIsect {         // RTIntersection, kernel_data.h:394-405
    int shapeid, primid, pad0, pad1;
    float4 uvwt;
};
struct AosThroughput {    // RTThroughput, kernel_data.h:154-160
    float4 throughput;
    int prevBsdfFlags, ignoreOcclusion, pad0, pad1;
};
static_assert(sizeof(AosRay) == 48 && sizeof(AosIsect) == 32 && sizeof(AosThroughput) == 32, "layout");

__device__ __forceinline__ Scene make_scene2(const PtState& s) {
    Scene sc;
    sc.shapes = (const RTShape*)s.shapes;
    sc.indices = s.indices;
    sc.positions = s.positions;
    sc.uvs = s.uvs;
    sc.normals = s.normals;
    sc.textures2D = (const TextureDesc2D*)s.textures2D;
    sc.texData2D = s.texData2D;
    sc.lights = (const RTLight*)s.lights;
    sc.materials = (const RTMaterial*)s.materials;
    sc.camera = (const RTPinholeCamera*)s.camera;
    sc.sobolMatrices = s.sobol;
    sc.materialAlpha = nullptr;   // the stage-wise kernels evaluate roughnessToAlpha per lookup, like the reference
    sc.triShade = nullptr;        // ... and walk shape -> indices -> vertices per vertex, like the reference
    sc.numLights = s.numLights;
    sc.staged = 0;
    return sc;
}

// setRay (kernel_data.h:434-440): o.xyz, o.w, d.xyz and extra are written; d.w and the tail are not
__device__ __forceinline__ void set_ray(AosRay* r, V3 p, float maxRange, V3 d) {
    r->o = make_float4(p.x, p.y, p.z, maxRange);
    r->d.x = d.x;
    r->d.y = d.y;
    r->d.z = d.z;
    r->extra = make_int2(-1, -1);
}

__global__ void __launch_bounds__(256) k_generate_rays_aos(PtState s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.width * s.height) return;
    const RTPinholeCamera* cam = (const RTPinholeCamera*)s.camera;
    const V3 d = primary_ray_dir(cam, i % s.width, i / s.width);
    set_ray((AosRay*)s.aos_rays + i, v3(cam->pos), 1000.0f, d);
}

__global__ void __launch_bounds__(128) k_pt_shade_aos(PtState s, int frame, int maxDepth, int bounce) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.width * s.height) return;
    AosRay* ray = (AosRay*)s.aos_rays + i;
    AosRay* shadow = (AosRay*)s.aos_shadow_rays + i;
    AosThroughput* tp = (AosThroughput*)s.aos_throughput + i;
    const AosIsect isect = ((const AosIsect*)s.aos_isect)[i];
    float4 radiance = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ray->extra.y != 0 && isect.shapeid != -1 && isect.primid != -1 && s.numLights > 0) {
        const Scene sc = make_scene2(s);
        ShadeIn in;
        in.pixel = i;
        in.shapeId = isect.shapeid;
        in.primId = isect.primid;
        in.bu = isect.uvwt.x;
        in.bv = isect.uvwt.y;
        in.rayDir = v3(ray->d);
        in.throughput = v3(tp->throughput);
        in.prevBsdfFlags = tp->prevBsdfFlags;
        ShadeOut out;
        shade_vertex<false>(sc, in, true, frame, s.width, s.height, bounce, maxDepth, &out);
        if (bounce == 0) tp->throughput = make_float4(1.f, 1.f, 1.f, tp->throughput.w);
        if (out.hasEmission) {
            radiance = make_float4(out.emitted.x, out.emitted.y, out.emitted.z, 0.f);
            shadow->extra.y = 0;
            ray->extra.y = 0;
            tp->ignoreOcclusion = 1;
        } else {
            tp->ignoreOcclusion = 0;
            if (out.shadow.written) set_ray(shadow, out.shadow.o, out.shadow.tmax, out.shadow.d);
            radiance = make_float4(out.direct.x, out.direct.y, out.direct.z, 0.f);
            if (bounce + 1 < maxDepth) {
                const int materialId = sc.shapes[isect.shapeid].materialId;
                if (materialId != RT_INVALID_ID) {
                    tp->prevBsdfFlags = out.bsdfFlags;
                    if (out.continues) {
                        tp->throughput = make_float4(out.throughput.x, out.throughput.y, out.throughput.z, tp->throughput.w);
                        set_ray(ray, out.nextO, RT_MAX_TRACE_DISTANCE, out.nextD);
                    } else {
                        ray->extra.y = 0;
                    }
                } else {
                    ray->extra.y = 0;
                }
            }
        }
    } else {
        ray->extra.y = 0;
    }
    s.aos_temp_radiance[i] = radiance;
}

__global__ void __launch_bounds__(256) k_shadow_pass_aos(PtState s, int bounce) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.width * s.height) return;   // the reference's `&&` border test (E11) is not reproduced
    float4 radiance = s.aos_temp_radiance[i];
    const AosThroughput* tp = (const AosThroughput*)s.aos_throughput + i;
    if (tp->ignoreOcclusion == 0) {
        const AosRay* sr = (const AosRay*)s.aos_shadow_rays + i;
        const float V = (sr->extra.y == 0 || s.aos_occlusion[i] != -1) ? 0.0f : 1.0f;
        radiance.x *= V;
        radiance.y *= V;
        radiance.z *= V;
        radiance.w *= V;
    }
    if (bounce == 0) {
        s.radiance[i] = radiance;
    } else {
        float4 r = s.radiance[i];
        r.x += radiance.x;
        r.y += radiance.y;
        r.z += radiance.z;
        r.w += radiance.w;
        s.radiance[i] = r;
    }
}

// ------------------------------------------------------------------ filters.cl:12-69
__device__ __forceinline__ float gaussian1(float d, float alpha, float expv) { return fmaxf(0.0f, expf(-alpha * d * d) - expv); }
__device__ __forceinline__ float mitchell1D(float x, float B, float C) {
    x = fabsf(2.0f * x);
    if (x > 1.0f)
        return ((-B - 6 * C) * x * x * x + (6 * B + 30 * C) * x * x + (-12 * B - 48 * C) * x + (8 * B + 24 * C)) * (1.f / 6.f);
    else
        return ((12 - 9 * B - 6 * C) * x * x * x + (-18 + 12 * B + 6 * C) * x * x + (6 - 2 * B)) * (1.f / 6.f);
}
__device__ __forceinline__ float sinc1(float x) {
    x = fabsf(x);
    if (x < 1e-5) return 1.0f;
    return sinf(PT_PI * x) / (PT_PI * x);
}
__device__ __forceinline__ float windowedSinc(float x, float radius, float tau) {
    x = fabsf(x);
    if (x > radius) return 0.0f;
    float lanczos = sinc1(x / tau);
    return sinc1(x) * lanczos;
}
__device__ __forceinline__ float filter_weight(const FilterProps& f) {
    const float px = f.pixel_offset[0], py = f.pixel_offset[1];
    switch (f.filter_type) {
        case 0: return 1.0f;
        case 1: return fmaxf(0.0f, f.radius[0] - fabsf(px)) * fmaxf(0.0f, f.radius[1] - fabsf(py));
        case 2: return gaussian1(px, f.gaussian_alpha, f.gaussian_exp_x) * gaussian1(py, f.gaussian_alpha, f.gaussian_exp_y);
        case 3: return mitchell1D(px / f.radius[0], f.mitchell_b, f.mitchell_c) * mitchell1D(py / f.radius[1], f.mitchell_b, f.mitchell_c);
        case 4: return windowedSinc(px, f.radius[0], f.lanczos_sinc_tau) * windowedSinc(py, f.radius[1], f.lanczos_sinc_tau);
        default: return 1.0f;
    }
}

#define RT_MAX_ALLOWED_RADIANCE 1000.0f   // kernel_data.h:13

__global__ void __launch_bounds__(256)
k_accumulate(PtState s, FilterProps f, int frame, const uint32_t* __restrict__ pixels, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t px = pixels ? pixels[i] : (uint32_t)i;
    float4 r = s.radiance[px];
    r.x = clampf(r.x, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.y = clampf(r.y, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.z = clampf(r.z, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.w = clampf(r.w, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    const float w = filter_weight(f);
    float4 acc;
    float wsum;
    if (frame == 0) {
        acc = make_float4(r.x * w, r.y * w, r.z * w, r.w * w);
        wsum = w;
    } else {
        acc = s.weighted_radiance[px];
        acc.x += r.x * w;
        acc.y += r.y * w;
        acc.z += r.z * w;
        acc.w += r.w * w;
        wsum = s.filter_weights[px] + w;
    }
    s.weighted_radiance[px] = acc;
    s.filter_weights[px] = wsum;
    s.image[px] = make_float4(acc.x / wsum, acc.y / wsum, acc.z / wsum, acc.w / wsum);
}

// ReconstructionPass for a BDPT frame rendered under the tile partition (ptb_set_partition): the rank's light-tracing
// (t = 1) strategies splat into pixels of ANY tile (BDPT.cl:888-899), so sum(w L) is taken over the whole frame,
// while sum(w) counts the sample only on the rank that owns the pixel -- summed over the ranks that gives, per
// pixel, sum over all ranks' contributions / one weight per sample, which is what a single device accumulates.
// (The clamp of reconstruction.cl:17-20 applies to each rank's share, a single device clamps their sum: they differ
// only for a radiance above RT_MAX_ALLOWED_RADIANCE.)
__global__ void __launch_bounds__(256)
k_accumulate_splat(PtState s, FilterProps f, int frame, int part, int num_parts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = s.width * s.height;
    if (i >= n) return;
    const int tw = (s.width + 63) / 64;
    const int tile = ((i / s.width) / 64) * tw + (i % s.width) / 64;
    const bool own = tile % num_parts == part;
    float4 r = s.radiance[i];
    r.x = clampf(r.x, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.y = clampf(r.y, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.z = clampf(r.z, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    r.w = clampf(r.w, 0.0f, RT_MAX_ALLOWED_RADIANCE);
    const float w = filter_weight(f);
    const float wo = own ? w : 0.0f;
    float4 acc;
    float wsum;
    if (frame == 0) {
        acc = make_float4(r.x * w, r.y * w, r.z * w, r.w * w);
        wsum = wo;
    } else {
        acc = s.weighted_radiance[i];
        acc.x += r.x * w;
        acc.y += r.y * w;
        acc.z += r.z * w;
        acc.w += r.w * w;
        wsum = s.filter_weights[i] + wo;
    }
    s.weighted_radiance[i] = acc;
    s.filter_weights[i] = wsum;
    if (own) s.image[i] = make_float4(acc.x / wsum, acc.y / wsum, acc.z / wsum, acc.w / wsum);   // partial until merged
}
cudaError_t launch_accumulate_splat(const PtState& s, const FilterProps& f, int frame, int part, int num_parts,
                                    cudaStream_t stream) {
    const int n = s.width * s.height;
    if (n <= 0) return cudaSuccess;
    k_accumulate_splat<<<(n + 255) / 256, 256, 0, stream>>>(s, f, frame, part, num_parts);
    return cudaGetLastError();
}

cudaError_t launch_generate_primary_rays(const PtState& s, cudaStream_t stream) {
    const int n = s.width * s.height;
    k_generate_rays_aos<<<(n + 255) / 256, 256, 0, stream>>>(s);
    return cudaGetLastError();
}
cudaError_t launch_pt_shade_aos(const PtState& s, int frame, int max_depth, int bounce, cudaStream_t stream) {
    const int n = s.width * s.height;
    k_pt_shade_aos<<<(n + 127) / 128, 128, 0, stream>>>(s, frame, max_depth, bounce);
    return cudaGetLastError();
}
cudaError_t launch_pt_shadow_resolve_aos(const PtState& s, int bounce, cudaStream_t stream) {
    const int n = s.width * s.height;
    k_shadow_pass_aos<<<(n + 255) / 256, 256, 0, stream>>>(s, bounce);
    return cudaGetLastError();
}
cudaError_t launch_accumulate(const PtState& s, const FilterProps& f, int frame, const uint32_t* pixels, int n_pixels,
                              cudaStream_t stream) {
    if (n_pixels <= 0) return cudaSuccess;
    k_accumulate<<<(n_pixels + 255) / 256, 256, 0, stream>>>(s, f, frame, pixels, n_pixels);
    return cudaGetLastError();
}

// The division ReconstructionPass ends with (reconstruction.cl:50-58), alone: after the ranks' accumulation buffers
// have been summed (ptb_merge_accumulation) the image is sum(w L) / sum(w) over the whole frame.
__global__ void __launch_bounds__(256)
k_divide_accumulation(const float4* __restrict__ weighted, const float* __restrict__ weights, float4* __restrict__ image, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 acc = weighted[i];
    const float wsum = weights[i];
    image[i] = make_float4(acc.x / wsum, acc.y / wsum, acc.z / wsum, acc.w / wsum);
}
cudaError_t launch_divide_accumulation(const float4* weighted, const float* weights, float4* image, int n, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    k_divide_accumulation<<<(n + 255) / 256, 256, 0, stream>>>(weighted, weights, image, n);
    return cudaGetLastError();
}

__global__ void k_material_alpha(const RTMaterial* __restrict__ materials, float2* __restrict__ alpha, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    alpha[i] = make_float2(roughnessToAlpha(materials[i].roughness.x), roughnessToAlpha(materials[i].roughness.y));
}
cudaError_t launch_material_alpha(const void* materials, float2* alpha, int num_materials, cudaStream_t stream) {
    if (num_materials <= 0) return cudaSuccess;
    k_material_alpha<<<(num_materials + 127) / 128, 128, 0, stream>>>((const RTMaterial*)materials, alpha, num_materials);
    return cudaGetLastError();
}

// self-test of pt_math.cuh:fdiv against the plain IEEE division (ptb_selftest_fdiv)
__global__ void k_selftest_fdiv(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ shortcut,
                                float* __restrict__ plain, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    shortcut[i] = fdiv(a[i], b[i]);
    plain[i] = __fdiv_rn(a[i], b[i]);
}
cudaError_t launch_selftest_fdiv(const float* a, const float* b, float* shortcut, float* plain, int n, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    k_selftest_fdiv<<<(n + 255) / 256, 256, 0, stream>>>(a, b, shortcut, plain, n);
    return cudaGetLastError();
}

// per-triangle shading records (pt_scene.cuh: TriShade): one thread per scene triangle, `tri_shape` names its RTShape
__global__ void __launch_bounds__(256)
k_build_tri_shade(PtState s, const int* __restrict__ tri_shape, float4* __restrict__ records, int n_tris) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    Scene sc = make_scene2(s);
    sc.triShade = nullptr;
    const int shapeId = tri_shape[t];
    const RTShape* shape = sc.shapes + shapeId;
    storeTriShade(records + 9 * (size_t)t, computeTriShade(sc, shapeId, t - (int)(shape->startIdx / 3u)));
}
cudaError_t launch_build_tri_shade(const PtState& s, const int* tri_shape, float4* records, int n_tris, cudaStream_t stream) {
    if (n_tris <= 0) return cudaSuccess;
    k_build_tri_shade<<<(n_tris + 255) / 256, 256, 0, stream>>>(s, tri_shape, records, n_tris);
    return cudaGetLastError();
}

// self-test of pt_libm.h on the device (ptb_selftest_libm): fn 0 = sin, 1 = cos, 2 = log
__global__ void k_selftest_libm(int fn, const float* __restrict__ in, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = in[i];
    out[i] = fn == 0 ? libm::sinf_glibc(x) : (fn == 1 ? libm::cosf_glibc(x) : libm::logf_glibc(x));
}
cudaError_t launch_selftest_libm(int fn, const float* in, float* out, int n, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    k_selftest_libm<<<(n + 255) / 256, 256, 0, stream>>>(fn, in, out, n);
    return cudaGetLastError();
}

}  // namespace mcrt
