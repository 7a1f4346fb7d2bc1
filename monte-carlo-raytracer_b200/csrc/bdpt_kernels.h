// bdpt_kernels.h -- device state and launcher of the bidirectional path tracer (bdpt_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "pt_kernels.h"

namespace mcrt {

// Per-pixel BDPT buffers (RTBDPTPass::createBuffers, RTBDPTPass.cpp:451-478), sized by maxDepth.
struct BdptState {
    void* cam_vertices;      // RTBDPTVertex[P * (D+2)]
    void* light_vertices;    // RTBDPTVertex[P * (D+1)]
    float4* cam_ray_o;       // camera / light subpath rays (SoA)
    float4* cam_ray_d;
    float4* light_ray_o;
    float4* light_ray_d;
    int4* cam_hit;
    int4* light_hit;
    float4* cam_beta;        // rgb throughput, w = pdfFwd of the next vertex (solid angle)
    float4* light_beta;
    int* cam_count;          // vertices on each subpath
    int* light_count;
    float4* conn_o;          // connection shadow rays, P * maxConn
    float4* conn_d;
    float4* conn_L;          // unweighted contribution rgb, w = MIS weight
    int* conn_pixel;         // target pixel of the contribution (t = 1 splats elsewhere)
    unsigned int* conn_count;
    float* splat;            // float[3P] RGB accumulator (atomic adds)
};

struct BdptBufferSpec {
    const char* name;
    size_t bytes_per_pixel;
    void** dst;
};

const std::vector<const char*>& bdpt_buffer_names();
std::vector<BdptBufferSpec> bdpt_buffer_specs(BdptState& b, int max_depth);

// One frame of RTBDPTPass::update (RTBDPTPass.cpp:67-128).  Returns the number of kernels launched.
int launch_bdpt_sample(const PtState& s, const BdptState& b, int frame, int max_depth, cudaStream_t stream,
                       cudaError_t* err);

}  // namespace mcrt
