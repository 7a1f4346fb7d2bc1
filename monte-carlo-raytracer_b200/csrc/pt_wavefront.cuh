// pt_wavefront.cuh -- device-side building blocks shared by the path-tracer and BDPT kernels:
// the Scene view of a PtState, the persistent warp-cooperative traversal loop over a work queue,
// and warp-ballot stream compaction.
#pragma once
#include "pt_kernels.h"
#include "pt_scene.cuh"
#include "traverse.cuh"

namespace mcrt {

__device__ __forceinline__ Scene make_scene(const PtState& s) {
    Scene sc;
    sc.shapes = (const RTShape*)s.shapes;
    sc.indices = s.indices;
    sc.positions = s.positions;
    sc.uvs = s.uvs;
    sc.normals = s.normals;
    sc.textures2D = (const TextureDesc2D*)s.textures2D;
    sc.texData2D = s.texData2D;
    sc.lights = (const RTLight*)s.lights;
    sc.materials = (const RTMaterial*)s.materials;
    sc.camera = (const RTPinholeCamera*)s.camera;
    sc.sobolMatrices = s.sobol;
    sc.materialAlpha = s.material_alpha;
    sc.triShade = s.tri_shade;
    sc.numLights = s.numLights;
    sc.staged = 0;
    return sc;
}

// ---------------------------------------------------------------------------------------------
// persistent traversal over a queue
// ---------------------------------------------------------------------------------------------
// MIXED: fetch() returns the kind of the ray it loaded (true = any hit); otherwise its result is ignored.
template <bool ANY_HIT, bool MIXED = false, bool TWO_LEVEL = false, typename Fetch, typename Store>
__device__ __forceinline__ void persistent_trace(const uint4* nodes, const float4* tris, unsigned count,
                                                 unsigned int* work_counter, Fetch fetch, Store store,
                                                 int num_nodes = 0, int stage_top = 0, const float4* instances = nullptr) {
    const unsigned lane = threadIdx.x & 31u;
    __shared__ TopNodes s_top;
    Traverser<ANY_HIT, false, MIXED, TWO_LEVEL> trav{nodes, tris};
    trav.instances = instances;
    trav.top = s_top.nodes;
    trav.top_count = stage_top_nodes(s_top, nodes, num_nodes, stage_top);
    uint2 stack[MCRT_STACK_SIZE];
    TravState st;
    Ray ray;
    int item = -1;
    bool exhausted = false;
    while (true) {
        const unsigned idle = __ballot_sync(0xffffffffu, item < 0 && !exhausted);
        if (idle) {
            const int n_idle = __popc(idle);
            const int my_rank = __popc(idle & ((1u << lane) - 1u));
            unsigned base = 0;
            const int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(work_counter, (unsigned)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (item < 0 && !exhausted) {
                const unsigned idx = base + (unsigned)my_rank;
                if (idx >= count) {
                    exhausted = true;
                } else {
                    const bool any = fetch(idx, ray);
                    trav.init(st, ray);
                    if (MIXED) st.any_hit = any;
                    item = (int)idx;
                }
            }
        }
        if (__all_sync(0xffffffffu, item < 0 && exhausted)) break;
        if (item >= 0) {
            if (trav.run(st, ray, stack, nullptr)) {
                store((unsigned)item, ray, st);
                item = -1;
            }
        }
    }
}

// warp-aggregated queue push: one atomic per warp, returns this lane's slot (valid where pred)
__device__ __forceinline__ unsigned warp_push(bool pred, unsigned int* counter) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return 0;
    unsigned base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(counter, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

inline int sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}


}  // namespace mcrt
