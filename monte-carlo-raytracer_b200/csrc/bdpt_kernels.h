// bdpt_kernels.h -- device state and launcher of the bidirectional path tracer (bdpt_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "pt_kernels.h"

namespace mcrt {

// Buffers of the BDPT integrator.  They replace RTBDPTPass::createBuffers (RTBDPTPass.cpp:451-478):
// same lifetime (sized by maxDepth, zero-filled at allocation, persistent across frames -- the
// sampled-light records are read before they are rewritten), different layout:
//  * path vertices are 128-byte records (8 x float4, see bdpt_kernels.cu) stored depth-major,
//    vertex i of pixel px at [i * P + px], so a warp extending one depth writes 4 KB contiguously
//    (the reference: 240-byte RTBDPTVertex, pixel-major);
//  * connections are a compacted queue of records instead of P * maxConn fixed slots.
struct BdptState {
    float4* cam_vertices;    // [(D+2) * P * 8]
    float4* light_vertices;  // [(D+1) * P * 8]
    float4* sampled_light;   // [D * P * 2]: {p.xyz, pdfFwd}, {gn.xyz, lightIdx | flags << 24}
    float4* cam_ray_o;       // subpath rays (xyz + tmax), directions, hits and running throughput
    float4* cam_ray_d;
    float4* light_ray_o;
    float4* light_ray_d;
    int4* cam_hit;
    int4* light_hit;
    float4* cam_beta;        // rgb throughput, w = solid-angle pdf of the direction just sampled
    float4* light_beta;
    int* cam_count;          // vertices on each subpath
    int* light_count;
    uint32_t* cam_queue[2];  // live-subpath queues (pixel indices), ping-pong
    uint32_t* light_queue[2];
    float4* conn_o;          // connection records [P * maxConn]: ray origin + length
    float4* conn_d;          //   direction, w = (t | s << 8) bits
    float4* conn_C;          //   unweighted contribution rgb, w = target pixel bits
    int* conn_src;           //   pixel whose subpaths are connected
    int* conn_vis;           //   1 = unoccluded (or no ray needed)
    float* conn_w;           //   MIS weight the record was accumulated with (0 when occluded), written by k_bdpt_connect
    uint32_t* conn_rays;     // indices of the records that need a shadow ray
    float* splat;            // float[3P] RGB accumulator (atomic adds; t = 1 strategies splat elsewhere)
    unsigned int* cnt;       // [0..1] cam queue counts, [2..3] light queue counts, [4] cam work, [5] light work,
                             // [6] records, [7] rays, [8] conn work
    int max_depth;
    int max_conn;
};

struct BdptBufferSpec {
    const char* name;
    size_t bytes_per_pixel;
    void** dst;
};

const std::vector<const char*>& bdpt_buffer_names();
std::vector<BdptBufferSpec> bdpt_buffer_specs(BdptState& b, int max_depth);

// One frame of RTBDPTPass::update (RTBDPTPass.cpp:67-128); the result lands in s.radiance
// ("RadianceBufferCL").  `pixels`/`n_pixels`: the rank's pixel list (NULL = all).  Returns the number
// of kernels launched.
int launch_bdpt_sample(const PtState& s, const BdptState& b, const uint32_t* pixels, int n_pixels, int frame,
                       int max_depth, cudaStream_t stream, cudaError_t* err);

}  // namespace mcrt
