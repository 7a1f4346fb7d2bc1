// lbvh_build.h -- device-side builder of the compressed 8-wide BVH (lbvh_build.cu).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "rrb_kernels.h"

namespace mcrt {

struct DeviceBvh8 {
    // Breadth-first levels the traversal's 40-entry stack is safe for (at most two entries per level).
    enum { kMaxLevels = 19 };
    void* nodes = nullptr;      // uint4[5 * num_nodes]   (bvh8.h layout), cudaMalloc'ed, owned by the caller on success
    void* tris = nullptr;       // float4[3 * num_tris]
    int* tri_mesh = nullptr;    // [num_tris] mesh index of each emitted triangle (refit input)
    float* boxes = nullptr;     // [6 * num_nodes] exact node bounds (refit scratch)
    int num_nodes = 0, num_tris = 0;
    std::vector<int> level_offsets;   // [levels + 1]
    int launches = 0;
};

// Builds the tree over the pooled object-space geometry already on the device (the refit inputs of rrb_api.cu):
// `face_prefix[m]` = number of faces of the meshes before m, `face_prefix.back()` = total.  Returns false with
// *err set when the input is outside what the device builder handles (fewer than 4 triangles, a tree deeper than
// kMaxLevels); the caller then uses the host builder.
bool build_bvh8_device(const float* d_vertices, const int* d_indices, const RefitMesh* d_meshes,
                       const std::vector<int>& face_prefix, DeviceBvh8* out, cudaStream_t stream, std::string* err);

}  // namespace mcrt
