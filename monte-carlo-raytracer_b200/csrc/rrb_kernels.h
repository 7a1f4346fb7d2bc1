// rrb_kernels.h -- host-callable launchers of the traversal kernels (rrb_kernels.cu).
#pragma once
#include <cuda_runtime.h>

namespace mcrt {

// Enqueues one closest-hit (any_hit=false) or any-hit query over rays[0..numrays) in the
// reference's API layouts.  `numrays_dev` (optional) is a device int holding the actual count
// (the indirect overloads, radeon_rays.h:274-277); `work_counter` is a device uint owned by the
// API object; `counts` is only used by the instrumented closest-hit variant.
cudaError_t launch_query(bool any_hit, bool counted, const void* nodes, const void* tris, const void* rays,
                         const int* numrays_dev, int numrays_max, void* out, unsigned int* work_counter,
                         int* counts, cudaStream_t stream, int num_nodes, const void* instances = nullptr);

// ---- device refit (refit_kernels.cu)
struct RefitMesh {
    float m[12];          // rows 0..2 of the row-major object -> world matrix
    int vertex_base;      // first vertex of the mesh in the pooled object-space vertex array
    int index_base;       // first index of the mesh in the pooled index array
    int shape_id;
    int pad;
};
// Recomputes the triangle records and re-fits all nodes (levels given on the host, breadth-first node order).
// Returns the number of kernels launched through *launches when not NULL.
cudaError_t refit_bvh8(void* nodes, void* tris, float* boxes, const int* tri_mesh, const RefitMesh* meshes,
                       const float* vertices, const int* indices, int num_tris, const int* level_offsets_host,
                       int num_levels, cudaStream_t stream);

}  // namespace mcrt
