// rrb_kernels.cu -- the query kernels behind rrb_query_intersection / rrb_query_occlusion.
//
// Persistent warps: the grid is sized to the SM count times the resident CTAs per SM; every warp
// pulls rays from a global counter, refilling idle lanes whenever fewer than
// MCRT_DYNAMIC_FETCH_THRESHOLD lanes are still traversing (warp ballot + one atomic per refill).
// I/O uses the reference's API layouts: ray 48 B in, Intersection 32 B / int32 out
// (intersect_bvh2_lds.cl:66-77,229-240).
#include "rrb_kernels.h"

#include <cstdlib>
#include "traverse.cuh"

namespace mcrt {

struct ApiRay {      // RadeonRays::ray, include/math/ray.h:103-107
    float4 o;        // xyz origin, w max t
    float4 d;        // xyz direction, w time
    int4 extra;      // x mask, y active, z doBackfaceCulling, w padding
};

template <bool ANY_HIT, bool COUNT, bool TWO_LEVEL>
__global__ void __launch_bounds__(128, TWO_LEVEL ? 6 : 8)
query_kernel(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const ApiRay* __restrict__ rays,
             const int* __restrict__ numrays_ptr, int numrays_max, void* __restrict__ out,
             unsigned int* __restrict__ work_counter, int* __restrict__ counts, int num_nodes, int stage_top,
             const float4* __restrict__ instances) {
    // a negative device-side count means no rays, as the reference's `index < *num_rays` does
    const int numrays = numrays_ptr ? max(0, min(*numrays_ptr, numrays_max)) : numrays_max;
    const unsigned lane = threadIdx.x & 31u;
    __shared__ TopNodes s_top;
    Traverser<ANY_HIT, COUNT, false, TWO_LEVEL> trav{nodes, tris};
    trav.instances = instances;
    trav.top = s_top.nodes;
    trav.top_count = stage_top_nodes(s_top, nodes, num_nodes, stage_top);
    uint2 stack[MCRT_STACK_SIZE];
    TravState st;
    Ray ray;
    int ray_index = -1;      // -1: lane idle
    bool exhausted = false;  // no more rays to fetch
    int cnt[2] = {0, 0};

    while (true) {
        // ---- refill idle lanes
        const unsigned idle = __ballot_sync(0xffffffffu, ray_index < 0 && !exhausted);
        if (idle) {
            const int n_idle = __popc(idle);
            const int my_rank = __popc(idle & ((1u << lane) - 1u));
            unsigned base = 0;
            const int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(work_counter, (unsigned)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (ray_index < 0 && !exhausted) {
                const unsigned idx = base + (unsigned)my_rank;
                if (idx >= (unsigned)numrays) {
                    exhausted = true;
                } else {
                    const ApiRay ar = rays[idx];
                    if (ar.extra.y != 0) {   // inactive rays are skipped and nothing is written
                        ray.ox = ar.o.x; ray.oy = ar.o.y; ray.oz = ar.o.z; ray.tmax = ar.o.w;
                        ray.dx = ar.d.x; ray.dy = ar.d.y; ray.dz = ar.d.z; ray.mask = ar.extra.x;
                        trav.init(st, ray);
                        ray_index = (int)idx;
                        if (COUNT) { cnt[0] = 0; cnt[1] = 0; }
                    }
                }
            }
        }
        if (__all_sync(0xffffffffu, ray_index < 0 && exhausted)) break;

        // ---- traverse
        if (ray_index >= 0) {
            const bool done = trav.run(st, ray, stack, cnt);
            if (done) {
                if (ANY_HIT) {
                    ((int*)out)[ray_index] = st.best_tri >= 0 ? 1 : -1;
                } else {
                    int4* h = (int4*)out + 2 * (size_t)ray_index;
                    if (st.best_tri >= 0) {
                        const float4 v0 = __ldg(tris + 3 * (size_t)st.best_tri + 0);
                        const float4 v1 = __ldg(tris + 3 * (size_t)st.best_tri + 1);
                        const float4 v2 = __ldg(tris + 3 * (size_t)st.best_tri + 2);
                        // (two levels: in the object space the triangle was hit in, intersect_bvh2level_skiplinks.cl:205-222)
                        const Ray hr = TWO_LEVEL ? hit_space_ray(instances, ray, st.best_inst) : ray;
                        const float px = add_(hr.ox, mul_(st.closest, hr.dx));
                        const float py = add_(hr.oy, mul_(st.closest, hr.dy));
                        const float pz = add_(hr.oz, mul_(st.closest, hr.dz));
                        const float2 uv = barycentrics_ref(px, py, pz, v0, v1, v2);
                        // shape_id, prim_id; the two padding ints are left as they are
                        ((int2*)h)[0] = make_int2(TWO_LEVEL ? st.best_sid : __float_as_int(v0.w), __float_as_int(v1.w));
                        ((float4*)h)[1] = make_float4(uv.x, uv.y, 0.0f, st.closest);
                    } else {
                        ((int2*)h)[0] = make_int2(-1, -1);   // uvwt untouched on a miss
                    }
                }
                if (COUNT) {
                    counts[2 * (size_t)ray_index + 0] = cnt[0];
                    counts[2 * (size_t)ray_index + 1] = cnt[1];
                }
                ray_index = -1;
            }
        }
    }
}

static int g_grid_blocks = 0;

static int grid_blocks() {
    if (g_grid_blocks == 0) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        g_grid_blocks = sms * 8;
    }
    return g_grid_blocks;
}

cudaError_t launch_query(bool any_hit, bool counted, const void* nodes, const void* tris, const void* rays,
                         const int* numrays_dev, int numrays_max, void* out, unsigned int* work_counter,
                         int* counts, cudaStream_t stream, int num_nodes, const void* instances) {
    if (numrays_max <= 0) return cudaSuccess;
    static const int stage_top = getenv("MCRT_STAGE_TOP") ? atoi(getenv("MCRT_STAGE_TOP")) : 1;
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const int warps_needed = (numrays_max + 31) / 32;
    int blocks = grid_blocks();
    blocks = blocks < (warps_needed + 3) / 4 ? blocks : (warps_needed + 3) / 4;
    const dim3 grid((unsigned)blocks), block(128);
    const uint4* n = (const uint4*)nodes;
    const float4* t = (const float4*)tris;
    const ApiRay* r = (const ApiRay*)rays;
    const float4* inst = (const float4*)instances;
#define MCRT_LAUNCH_QUERY(A, C, T) \
    query_kernel<A, C, T><<<grid, block, 0, stream>>>(n, t, r, numrays_dev, numrays_max, out, work_counter, counts, num_nodes, stage_top, inst)
    if (inst) {
        if (any_hit) MCRT_LAUNCH_QUERY(true, false, true);
        else if (counted) MCRT_LAUNCH_QUERY(false, true, true);
        else MCRT_LAUNCH_QUERY(false, false, true);
    } else {
        if (any_hit) MCRT_LAUNCH_QUERY(true, false, false);
        else if (counted) MCRT_LAUNCH_QUERY(false, true, false);
        else MCRT_LAUNCH_QUERY(false, false, false);
    }
#undef MCRT_LAUNCH_QUERY
    return cudaGetLastError();
}

}  // namespace mcrt
