// ref_pt_kernels.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// Compiles the reference's own assets/kernels/PathTracing.cl (and everything it includes:
// rng, samplers, bxdfs, materials, geometry, lights, textures, math, matrix, kernel_data.h),
// read from /root/reference through cl2cpp.py, as C++.  Each extern "C" entry runs one
// NDRange the way the reference host launches it (RTPrimaryRaysPass.cpp:110-120,
// RTPathTracingPass.cpp:127-147,177-190): one work-item per pixel, here one loop iteration.
#include "clshim.hpp"
#include "ref_api.h"
#include <omp.h>

namespace clshim {
extern thread_local size_t g_gid[3];
extern thread_local size_t g_lid[3];
}

namespace refpt {
#include "app/PathTracing.cl"
}

using namespace refpt;

template <typename Fn>
static void launch2d(int width, int height, Fn fn) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            clshim::g_gid[0] = (size_t)x;
            clshim::g_gid[1] = (size_t)y;
            clshim::g_lid[0] = (size_t)(x & 7);
            clshim::g_lid[1] = (size_t)(y & 7);
            fn();
        }
    }
}

extern "C" {

// The OpenMP team size is set by the caller, not inherited from the environment: a launcher such as torchrun exports
// OMP_NUM_THREADS=1 to its children, which would silently turn the CPU baseline into a single-threaded run.
void ref_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
}
// the team size an `omp parallel` region actually gets (counted inside one)
int ref_get_threads(void) {
    int n = 0;
#pragma omp parallel
    {
#pragma omp atomic
        n += 1;
    }
    return n;
}

int ref_pt_struct_sizes(int* out, int n) {
    const int sizes[] = {(int)sizeof(RTShape),      (int)sizeof(RTMaterial),   (int)sizeof(RTLight),
                         (int)sizeof(RTThroughput), (int)sizeof(RTInteraction), (int)sizeof(RTBDPTVertex),
                         (int)sizeof(RTPinholeCamera), (int)sizeof(RTRayDifferentials), (int)sizeof(TextureDesc2D),
                         (int)sizeof(RTRay),        (int)sizeof(RTIntersection)};
    int cnt = (int)(sizeof(sizes) / sizeof(sizes[0]));
    for (int i = 0; i < n && i < cnt; ++i) out[i] = sizes[i];
    return cnt;
}

void ref_pt_generate_rays(void* rays, void* ray_differentials, const void* camera) {
    const RTPinholeCamera* cam = (const RTPinholeCamera*)camera;
    launch2d((int)cam->width, (int)cam->height, [&]() {
        GeneratePerspectiveRays((RTRay*)rays, (RTRayDifferentials*)ray_differentials, cam);
    });
}

void ref_pt_shade(const ref_scene_desc* s, int width, int height, void* shadow_rays, void* rays,
                  const void* isects, int frame, int max_depth, int bounce, void* temp_radiance, void* throughput) {
    launch2d(width, height, [&]() {
        PathTracing((const RTShape*)s->shapes, s->indices, (const float3*)s->positions, (const float2*)s->uvs,
                    (const float3*)s->normals, nullptr, nullptr, nullptr, (const TextureDesc2D*)s->textures,
                    s->texdata, s->sobol, (const RTLight*)s->lights, s->num_lights, (const RTMaterial*)s->materials,
                    (const RTPinholeCamera*)s->camera, width, height, (RTRay*)shadow_rays, (RTRay*)rays,
                    (RTIntersection*)isects, frame, max_depth, bounce, (float4*)temp_radiance,
                    (RTThroughput*)throughput);
    });
}

void ref_pt_shadow_pass(const void* shadow_rays, const int* occlusion, int width, int height, int bounce,
                        void* throughput, void* temp_radiance, void* radiance) {
    launch2d(width, height, [&]() {
        ShadowPass((const RTRay*)shadow_rays, occlusion, width, height, bounce, (RTThroughput*)throughput,
                   (float4*)temp_radiance, (float4*)radiance);
    });
}
}
