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
                         int* counts, cudaStream_t stream);

}  // namespace mcrt
