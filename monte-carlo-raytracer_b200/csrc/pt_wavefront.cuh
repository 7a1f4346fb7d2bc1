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
//
// Cooperative tail (flat trees, coop.min_iters > 0).  One ray per lane with dynamic refill leaves a launch waiting for
// its slowest rays: a ray that grazes a finely tessellated surface visits hundreds of nodes and triangles, one
// dependent memory access after the other, while the other lanes of its warp -- and, at the end, the whole GPU -- have
// nothing left to fetch.  So a ray that is still being walked coop.min_iters rounds after the queue ran dry is
// ABANDONED by its lane and put on a list of stragglers; every warp, once it is out of work, pulls rays off that list
// and walks each with all 32 lanes: lane k restarts the ray restricted to the k-th of 32 intervals of t (over the part
// of the ray inside the root node's frame), see Traverser::run<SPLIT>.  Same origin and direction, so every t is the
// value the single walk computes; the intervals partition the accepted hits, hence the smallest (t, shape id, prim id)
// over the lanes IS the single walk's answer, tie rule included (any-hit rays: any lane's hit).  A first version that
// let a warp's idle lanes help inside the main loop cost the main loop 11 % (spills at the 64-register cap) and split
// far too many rays; here the main loop only counts rounds, and the stragglers of ALL warps are balanced over the grid.
#ifndef MCRT_COOP_ROUNDS
#define MCRT_COOP_ROUNDS 8
#endif
struct CoopTail {
    int min_iters;                 // 0 = off
    unsigned int* list;            // ray index + 1 of the launch's stragglers; all zero between launches
    unsigned int* count;           // [0] entries in `list`, [1] entries taken; both zero at launch
};
__device__ __forceinline__ float next_up_nonneg(float x) { return __int_as_float(__float_as_int(x) + 1); }

template <bool ANY_HIT, bool MIXED = false, bool TWO_LEVEL = false, bool COOP = false, typename Fetch, typename Store>
__device__ __forceinline__ void persistent_trace(const uint4* nodes, const float4* tris, unsigned count,
                                                 unsigned int* work_counter, Fetch fetch, Store store,
                                                 int num_nodes = 0, int stage_top = 0, const float4* instances = nullptr,
                                                 CoopTail coop = CoopTail{0, nullptr, nullptr}) {
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
    const bool coop_on = COOP && !TWO_LEVEL && coop.min_iters > 0;      // COOP = false: the code below is not even compiled in
    int tail_iters = 0;          // rounds this lane's ray has been walked since the queue ran dry
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
                    tail_iters = 0;
                }
            }
        }
        if (__all_sync(0xffffffffu, item < 0 && exhausted)) break;
        const bool dry = coop_on && __ballot_sync(0xffffffffu, exhausted) != 0;      // a lane found the queue empty
        if (item >= 0) {
            if (trav.run(st, ray, stack, nullptr)) {
                store((unsigned)item, ray, st);
                item = -1;
            } else if (dry && ++tail_iters >= coop.min_iters) {
                coop.list[atomicAdd(coop.count, 1u)] = (unsigned)item + 1u;     // a straggler: the grid will finish it
                item = -1;
                exhausted = true;
            }
        }
    }
    if (!coop_on) return;
    // ---- the stragglers, one ray per WARP
    while (true) {
        // take the next listed entry, if there is one.  `taken` never runs ahead of `listed`: an entry appended after
        // this warp gave up is found by the warp that appended it, which comes here after its last append.
        unsigned idx = 0xffffffffu;
        if (lane == 0) {
            while (true) {
                const unsigned taken = *(volatile unsigned int*)(coop.count + 1);
                if (taken >= *(volatile unsigned int*)coop.count) break;
                if (atomicCAS(coop.count + 1, taken, taken + 1u) == taken) {
                    unsigned v;      // the slot was counted before it was written: 0 = not there yet
                    while ((v = *(volatile unsigned int*)(coop.list + taken)) == 0u) {
                    }
                    coop.list[taken] = 0u;      // the list is all zero again when the launch ends
                    idx = v - 1u;
                    break;
                }
            }
        }
        idx = __shfl_sync(0xffffffffu, idx, 0);
        if (idx == 0xffffffffu) break;
        const bool any = fetch(idx, ray);
        trav.init(st, ray);
        const bool any_kind = ANY_HIT || (MIXED && any);
        // the stretch of the ray that can hit anything: inside the root node's quantisation frame (origin +
        // [0, 255] * 2^e per axis encloses every child box).  Only the CHOICE of the slice boundaries depends on it;
        // any boundaries give the same union.
        const float cl = ray.tmax;
        float t0 = 0.0f, t1 = cl;
        {
            const uint4 n0 = __ldg(nodes);
            const float ex = 255.0f * __uint_as_float(extract_byte(n0.w, 0) << 23);
            const float ey = 255.0f * __uint_as_float(extract_byte(n0.w, 1) << 23);
            const float ez = 255.0f * __uint_as_float(extract_byte(n0.w, 2) << 23);
            const float ax = (__uint_as_float(n0.x) - ray.ox) * st.idx, bx = (__uint_as_float(n0.x) + ex - ray.ox) * st.idx;
            const float ay = (__uint_as_float(n0.y) - ray.oy) * st.idy, by = (__uint_as_float(n0.y) + ey - ray.oy) * st.idy;
            const float az = (__uint_as_float(n0.z) - ray.oz) * st.idz, bz = (__uint_as_float(n0.z) + ez - ray.oz) * st.idz;
            const float enter = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
            const float leave = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
            if (enter > 0.0f) t0 = enter;
            if (leave < t1) t1 = leave;
            if (!(t1 > t0)) {
                t0 = 0.0f;
                t1 = cl;
            }
        }
        const float span = t1 - t0;
        // Front to back in MCRT_COOP_ROUNDS rounds of 32 slices, lane k on slice 32 * round + k: a ray's cost is
        // rarely spread evenly along it (measured: 32 equal intervals leave one lane with nearly all of the walk), so
        // the slices are thin -- a costly stretch spreads over many lanes of one or two rounds -- and a round that
        // finds a hit ends the sweep, every later slice being farther away.
        float t = __int_as_float(0x7f800000);
        int sid = 0x7fffffff, pid = 0x7fffffff, tri = -1;
        for (int round = 0; round < MCRT_COOP_ROUNDS && tri < 0; ++round) {
            const unsigned q = (unsigned)round * 32u + lane;
            const float n_slices = 32.0f * MCRT_COOP_ROUNDS;
            const float lo = t0 + span * ((float)q / n_slices);
            const float hi = t0 + span * ((float)(q + 1u) / n_slices);      // = the next slice's lo, bit for bit
            trav.init(st, ray);
            if (MIXED) st.any_hit = any;
            st.tlo = q == 0 ? -1.0f : lo;
            st.tlo_box = q == 0 ? 0.0f : lo * 0.9999996f;
            // a lane accepts t < closest: the float above the boundary closes its slice on the right; the last slice
            // keeps the ray's own bound (t < tmax, as the single walk)
            st.closest = q == 32u * MCRT_COOP_ROUNDS - 1u ? cl : next_up_nonneg(fmaxf(hi, 0.0f));
            bool done = false;
            while (true) {
                if (!done) done = trav.template run<true>(st, ray, stack, nullptr);
                if (__ballot_sync(0xffffffffu, done) == 0xffffffffu) break;
                if (any_kind && __ballot_sync(0xffffffffu, done && st.best_tri >= 0) != 0) break;
            }
            if (done && st.best_tri >= 0) {
                t = any_kind ? 0.0f : st.closest;
                sid = st.best_sid;
                pid = st.best_pid;
                tri = st.best_tri;
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                const float t2 = __shfl_xor_sync(0xffffffffu, t, d);
                const int s2 = __shfl_xor_sync(0xffffffffu, sid, d), p2 = __shfl_xor_sync(0xffffffffu, pid, d);
                const int r2 = __shfl_xor_sync(0xffffffffu, tri, d);
                const bool take = r2 >= 0 && (tri < 0 || t2 < t || (t2 == t && (s2 < sid || (s2 == sid && (p2 < pid || (p2 == pid && r2 < tri))))));
                if (take) {
                    t = t2;
                    sid = s2;
                    pid = p2;
                    tri = r2;
                }
            }
        }
        if (lane == 0) {
            trav.init(st, ray);                 // a clean result record: the winner, or a miss
            if (MIXED) st.any_hit = any;
            if (tri >= 0) {
                if (!any_kind) st.closest = t;
                st.best_tri = tri;
                st.best_sid = sid;
                st.best_pid = pid;
            }
            store(idx, ray, st);
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
