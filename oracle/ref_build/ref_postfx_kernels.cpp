// ref_postfx_kernels.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// The reference's post-processing kernels compiled as C++: ReinhardToneMapping
// (assets/kernels/ToneMapping.cl:41-63) and BilateralDenoise (assets/kernels/Denoise.cl:6-47), launched
// over W x H like RTToneMappingPass.cpp:60-70 / RTDenoisePass.cpp:48-62 do.  Images are RGBA32F buffers.
#include "clshim.hpp"
#include "ref_api.h"
#include <omp.h>

namespace clshim {
extern thread_local size_t g_gid[3];
extern thread_local size_t g_lid[3];
}

namespace refpost {
#include "app/ToneMapping.cl"
#include "app/Denoise.cl"
}

using namespace refpost;

template <typename Fn>
static void launch2d(int width, int height, Fn fn) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            clshim::g_gid[0] = (size_t)x;
            clshim::g_gid[1] = (size_t)y;
            fn();
        }
}

extern "C" {

void ref_reinhard_tonemap(int width, int height, float lwhite, void* in_rgba, void* out_rgba) {
    clshim_image src{(float4*)in_rgba, width, height}, dst{(float4*)out_rgba, width, height};
    launch2d(width, height, [&]() { ReinhardToneMapping(width, height, lwhite, &src, &dst); });
}

void ref_bilateral_denoise(int width, int height, int radius, float sigma_spatial, float sigma_range, void* in_rgba,
                           void* out_rgba) {
    clshim_image src{(float4*)in_rgba, width, height}, dst{(float4*)out_rgba, width, height};
    launch2d(width, height,
             [&]() { BilateralDenoise(width, height, radius, sigma_spatial, sigma_range, &src, &dst); });
}
}
