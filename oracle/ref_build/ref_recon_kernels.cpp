// ref_recon_kernels.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// Compiles the reference's assets/kernels/reconstruction.cl (+ filters.cl) as C++ and runs its
// ReconstructionPassAllFilters kernel (reconstruction.cl:62-121), which accumulates the five pixel
// filters side by side into plain buffers (the single-filter ReconstructionPass, :6-60, computes the
// same weight and sums but writes a GL-interop image, which does not exist here).
#include "clshim.hpp"
#include "ref_api.h"
#include <omp.h>

namespace clshim {
extern thread_local size_t g_gid[3];
extern thread_local size_t g_lid[3];
}

namespace refrecon {
#include "app/reconstruction.cl"
}

using namespace refrecon;

extern "C" {

int ref_filter_props_size(void) { return (int)sizeof(RTFilterProperties); }

// props: radius x/y, mitchell B/C, lanczos tau, gaussian alpha/expX/expY, pixel offset x/y (10 floats)
void ref_reconstruction_all_filters(int width, int height, int frame, const float* props, void* radiance,
                                    void* weighted /* float4[5P] */, float* weights /* float[5P] */) {
    RTFilterProperties fp;
    fp.filterType = RT_TRIANGLE_FILTER;
    fp.radius = float2{props[0], props[1]};
    fp.mitchellB = props[2];
    fp.mitchellC = props[3];
    fp.lanczosSincTau = props[4];
    fp.gaussianAlpha = props[5];
    fp.gaussianExpX = props[6];
    fp.gaussianExpY = props[7];
    fp.pixelOffset = float2{props[8], props[9]};
#pragma omp parallel for schedule(static)
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            clshim::g_gid[0] = (size_t)x;
            clshim::g_gid[1] = (size_t)y;
            ReconstructionPassAllFilters(width, height, frame, &fp, (float4*)radiance, (float4*)weighted, weights);
        }
}
}
