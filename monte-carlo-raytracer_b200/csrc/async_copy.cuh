// async_copy.cuh -- bulk asynchronous global -> shared copies (the 1-D form of TMA, cp.async.bulk) with completion on
// an mbarrier, for sm_100a.  Used to stage per-CTA read-only tables: the top levels of the BVH in the traversal
// kernels, the scene's shape / material / light / texture-descriptor tables in the shade kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mcrt {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// The first `count` nodes of the breadth-first node array = the top levels of the tree (root, its <= 8 children,
// their <= 64: 73 nodes), which every ray visits: one bulk copy per CTA, then they are served from shared memory
// (Radeon Rays keeps its per-lane traversal stack in LDS, intersect_bvh2_lds.cl:78,144-153; here it is the hot
// node set).  MCRT_TOP_NODES * 80 B = 6.4 KB per CTA.
#ifndef MCRT_TOP_NODES
#define MCRT_TOP_NODES 80
#endif
struct TopNodes {
    __align__(16) uint4 nodes[5 * MCRT_TOP_NODES];
    uint64_t bar;
};
// every thread of the CTA calls it; returns the number of nodes staged (0 = staging off)
__device__ __forceinline__ unsigned stage_top_nodes(TopNodes& t, const uint4* nodes, int num_nodes, int enable) {
    const unsigned n = enable ? (unsigned)min(num_nodes, MCRT_TOP_NODES) : 0u;
    if (n == 0) return 0;
    if (threadIdx.x == 0) mbar_init(&t.bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&t.bar, n * 80u);
        bulk_copy_g2s(t.nodes, nodes, n * 80u, &t.bar);
    }
    mbar_wait(&t.bar, 0);
    return n;
}

}  // namespace mcrt
