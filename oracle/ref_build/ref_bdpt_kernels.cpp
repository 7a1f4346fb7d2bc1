// ref_bdpt_kernels.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// Compiles the reference's own assets/kernels/BDPT.cl (and its includes), read from
// /root/reference through cl2cpp.py, as C++.  Each extern "C" entry runs one NDRange the way
// RTBDPTPass launches it (renderPasses/RTBDPTPass.cpp:130-448): one work-item per pixel.
// ConnectVertices adds into other pixels' accumulators with its CAS-loop float atomic, which the
// shim maps to a real atomic, so the OpenMP loop is race-free (summation order is not fixed, as
// in the reference).
#include "clshim.hpp"
#include "ref_api.h"
#include <omp.h>

namespace clshim {
extern thread_local size_t g_gid[3];
extern thread_local size_t g_lid[3];
}

namespace refbdpt {
#include "app/BDPT.cl"
}

using namespace refbdpt;

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

#define SCENE_ARGS(s)                                                                                             \
    (const RTShape*)(s)->shapes, (s)->indices, (const float3*)(s)->positions, (const float2*)(s)->uvs,            \
        (const float3*)(s)->normals, nullptr, nullptr, nullptr, (const TextureDesc2D*)(s)->textures, (s)->texdata, \
        (s)->sobol, (const RTLight*)(s)->lights, (s)->num_lights, (const RTMaterial*)(s)->materials,              \
        (const RTPinholeCamera*)(s)->camera

extern "C" {

int ref_bdpt_vertex_size(void) { return (int)sizeof(RTBDPTVertex); }

void ref_bdpt_generate_start_vertices(const ref_scene_desc* s, int width, int height, int frame, int max_depth,
                                      void* cam_vertices, void* cam_rays, void* cam_throughputs, float* cam_fwd_pdfs,
                                      int* cam_counts, void* light_vertices, void* light_rays,
                                      void* light_throughputs, float* light_fwd_pdfs, int* light_counts,
                                      float* final_radiance) {
    launch2d(width, height, [&]() {
        GenerateStartVertices(SCENE_ARGS(s), width, height, frame, max_depth, (RTBDPTVertex*)cam_vertices,
                              (RTRay*)cam_rays, (float3*)cam_throughputs, cam_fwd_pdfs, cam_counts,
                              (RTBDPTVertex*)light_vertices, (RTRay*)light_rays, (float3*)light_throughputs,
                              light_fwd_pdfs, light_counts, final_radiance);
    });
}

void ref_bdpt_generate_secondary_vertices(const ref_scene_desc* s, int width, int height, int frame, int max_depth,
                                          int is_camera_path, void* vertices, void* rays, const void* isects,
                                          void* throughputs, float* fwd_pdfs, int* counts, int cur_depth) {
    launch2d(width, height, [&]() {
        GenerateSecondaryVertices(SCENE_ARGS(s), width, height, frame, max_depth, is_camera_path,
                                  (RTBDPTVertex*)vertices, (RTRay*)rays, (const RTIntersection*)isects,
                                  (float3*)throughputs, fwd_pdfs, counts, cur_depth);
    });
}

void ref_bdpt_prepare_connections(const ref_scene_desc* s, int width, int height, int frame, int max_depth,
                                  void* cam_vertices, void* light_vertices, void* sampled_cam, void* sampled_light,
                                  void* connection_rays, const int* cam_counts, const int* light_counts,
                                  void* temp_radiance) {
    launch2d(width, height, [&]() {
        PrepareConnections(SCENE_ARGS(s), width, height, frame, max_depth, (RTBDPTVertex*)cam_vertices,
                           (RTBDPTVertex*)light_vertices, (RTBDPTVertex*)sampled_cam, (RTBDPTVertex*)sampled_light,
                           (RTRay*)connection_rays, cam_counts, light_counts, (float4*)temp_radiance);
    });
}

void ref_bdpt_connect_vertices(const ref_scene_desc* s, int width, int height, int frame, int max_depth,
                               void* cam_vertices, void* light_vertices, void* sampled_cam, void* sampled_light,
                               void* connection_rays, const int* visibilities, const int* cam_counts,
                               const int* light_counts, void* temp_radiance, float* final_radiance) {
    launch2d(width, height, [&]() {
        ConnectVertices(SCENE_ARGS(s), width, height, frame, max_depth, (RTBDPTVertex*)cam_vertices,
                        (RTBDPTVertex*)light_vertices, (RTBDPTVertex*)sampled_cam, (RTBDPTVertex*)sampled_light,
                        (RTRay*)connection_rays, visibilities, cam_counts, light_counts, (float4*)temp_radiance,
                        final_radiance);
    });
}

void ref_bdpt_copy_buffer(int width, int height, const float* src, void* dst) {
    launch2d(width, height, [&]() { CopyBuffer(width, height, src, (float4*)dst); });
}
}
