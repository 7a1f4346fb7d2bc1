// pt_kernels.cu -- the regular path tracer's kernels for sm_100a.
//
// Wavefront design (replaces the reference's "every stage over all W*H work-items, dead paths
// early-out", RTPathTracingPass.cpp:70-86): per-pixel SoA path state in HBM, queues of live pixel
// indices compacted with a warp ballot + one atomic per warp, persistent warp-cooperative
// traversal kernels that pull work from the queues, the shadow-ray resolve (the reference's
// ShadowPass, PathTracing.cl:186-217) fused into the any-hit traversal, the primary-ray
// generation fused with the zeroing of the radiance buffer.  Per bounce: closest-hit trace ->
// shade -> any-hit trace + resolve.  No host synchronisation inside a frame: queue sizes stay on
// the device.
#include "pt_kernels.h"

#include <cstdlib>

#include "pt_shade.cuh"
#include "pt_wavefront.cuh"

namespace mcrt {

// ---------------------------------------------------------------------------------------------
// queue control: one thread, between stages (keeps all counts on the device)
// ---------------------------------------------------------------------------------------------
// cnt: [0],[1] live-path queue sizes (ping-pong)  [3],[4] work counters of the persistent kernels
//      [5],[6] shadow-ray counts by bounce parity (the shadow rays of bounce b are traced together with the
//      closest-hit rays of bounce b + 1, while nothing writes either set)
//      [7] vertex records written by k_shade_surface for k_shade_scatter (split shade stage)
//      [25],[26] stragglers of the running traversal launch: listed, taken (cooperative tail, pt_wavefront.cuh)
//      [8..15] coherence-bucket counts of the queue being filled, [16..24] prefix sums of the queue about to be read
__global__ void k_bounce_ctl(unsigned int* cnt, unsigned long long* stats, int cur, int bounce, int n_first, int bucketed) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (bucketed) {   // the queue of this bounce sits in the staging regions: k_queue_gather makes it contiguous
        unsigned sum = 0;
        for (int k = 0; k < PT_NUM_BUCKETS; ++k) {
            cnt[16 + k] = sum;
            sum += cnt[8 + k];
            cnt[8 + k] = 0;
        }
        cnt[16 + PT_NUM_BUCKETS] = sum;
        cnt[cur] = sum;
    }
    // stats are shared by the two halves of a frame (two streams): atomic adds
    if (bounce == 0) {
        cnt[cur] = (unsigned)n_first;
        atomicAdd(stats + 3, (unsigned long long)n_first);
    } else {
        atomicAdd(stats + 1, (unsigned long long)cnt[5 + ((bounce - 1) & 1)]);   // shadow rays of the bounce just shaded
    }
    atomicAdd(stats + 0, (unsigned long long)cnt[cur]);     // closest-hit rays of the bounce about to start
    atomicAdd(stats + 2, (unsigned long long)cnt[cur]);     // = path vertices handed to the shade stage
    cnt[1 - cur] = 0;
    cnt[5 + (bounce & 1)] = 0;
    cnt[3] = 0;
    cnt[4] = 0;
    cnt[7] = 0;      // vertex records of the split shade stage
    cnt[25] = 0;     // stragglers of the traversal launch about to start (cooperative tail): listed, taken
    cnt[26] = 0;
}
__global__ void k_sample_end(unsigned int* cnt, unsigned long long* stats, int last_bounce) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    atomicAdd(stats + 1, (unsigned long long)cnt[5 + (last_bounce & 1)]);
    cnt[4] = 0;
    cnt[25] = 0;
    cnt[26] = 0;
}

// ---------------------------------------------------------------------------------------------
// primary rays (GeneratePerspectiveRays, PathTracing.cl:13-35) + radiance clear + queue fill
// ---------------------------------------------------------------------------------------------
// Coherence key of the vertex that pixel `px` will shade at `bounce`: the sampler is a pure function of
// (pixel, frame, bounce) (samplers.cl:75-84), so the light the vertex will pick (first 1D draw) and the half of
// the lobe-selection variate (fourth draw; PathTracing.cl:106-139 draws 1D light pick, 2D light sample, 2D BSDF
// sample) are known before the ray is even traced.  Queues ordered by this key give the shade kernel warps whose
// lanes take the same light-sampling and lobe-sampling branches; per-path results do not depend on queue order.
__device__ __forceinline__ unsigned coherence_key(const PtState& s, uint32_t px, int frame, int bounce) {
    Sampler sp = makeSampler(s.sobol, (int)px, frame, s.width, s.height, bounce);
    uint32_t lightIdx = ((uint32_t)floorf(getSample1D(&sp) * s.numLights));
    lightIdx %= (uint32_t)s.numLights;
    skipSamples(&sp, 2);
    const float u = getSample1D(&sp);
    return (lightIdx < 3u ? lightIdx : 3u) * 2u + (u >= 0.5f ? 1u : 0u);
}
// one atomic per distinct key in the warp; returns the lane's slot within its key's region
__device__ __forceinline__ unsigned warp_push_keyed(bool pred, unsigned key, unsigned int* counters) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned m = __match_any_sync(0xffffffffu, pred ? key : 0xffffffffu);
    const int leader = __ffs(m) - 1;
    unsigned base = 0;
    if (pred && (int)lane == leader) base = atomicAdd(counters + key, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(256)
k_raygen(PtState s, const uint32_t* __restrict__ pixels, int first, int n, int frame) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    uint32_t px = 0;
    if (live) {
        px = pixels ? pixels[first + i] : (uint32_t)(first + i);
        const RTPinholeCamera* cam = (const RTPinholeCamera*)s.camera;
        const int x = (int)(px % (uint32_t)s.width), y = (int)(px / (uint32_t)s.width);
        const V3 d = primary_ray_dir(cam, x, y);
        s.ray_o[px] = make_float4(cam->pos.x, cam->pos.y, cam->pos.z, 1000.0f);
        s.ray_d[px] = make_float4(d.x, d.y, d.z, 0.0f);
        s.radiance[px] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (s.coherence == 2 && s.numLights > 0) {   // without lights no bounce runs (launch_pt_sample)
        const unsigned key = live ? coherence_key(s, px, frame, 0) : 0u;
        const unsigned slot = warp_push_keyed(live, key, s.cnt + 8);
        if (live) s.bucket[(size_t)key * (size_t)n + slot] = px;
    } else if (live) {
        s.queue[0][i] = px;
    }
}

// staging regions -> contiguous queue, bucket by bucket (prefix sums written by k_bounce_ctl)
__global__ void __launch_bounds__(256)
k_queue_gather(PtState s, int cur, int cap) {
    __shared__ unsigned prefix[PT_NUM_BUCKETS + 1];
    if (threadIdx.x <= PT_NUM_BUCKETS) prefix[threadIdx.x] = s.cnt[16 + threadIdx.x];
    __syncthreads();
    const unsigned total = prefix[PT_NUM_BUCKETS];
    uint32_t* __restrict__ q = s.queue[cur];
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        unsigned k = 0;
#pragma unroll
        for (int j = 1; j < PT_NUM_BUCKETS; ++j) k += (i >= prefix[j]) ? 1u : 0u;
        q[i] = s.bucket[(size_t)k * (size_t)cap + (i - prefix[k])];
    }
}

// the cooperative tail of a traversal launch (pt_wavefront.cuh): straggler list and its two counters
__device__ __forceinline__ CoopTail coop_tail(const PtState& s) {
    return CoopTail{s.strag ? s.coopIters : 0, s.strag, s.cnt + 25};
}

// closest-hit ray of queue entry idx / its result
__device__ __forceinline__ void fetch_closest_ray(const PtState& s, const uint32_t* __restrict__ queue, unsigned idx, Ray& r) {
    const uint32_t px = queue[idx];
    const float4 o = s.ray_o[px], d = s.ray_d[px];
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmax = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.mask = -1;   // setRay(): extra = 0xFFFFFFFF (kernel_data.h:434-440)
}
template <bool TWO_LEVEL = false>
__device__ __forceinline__ void store_closest_hit(const PtState& s, const uint32_t* __restrict__ queue, unsigned idx,
                                                  const Ray& rw, const TravState& st) {
    const uint32_t px = queue[idx];
    int4 h = make_int4(-1, -1, 0, 0);
    if (st.best_tri >= 0) {
        const float4 v0 = __ldg(s.tris + 3 * (size_t)st.best_tri + 0);
        const float4 v1 = __ldg(s.tris + 3 * (size_t)st.best_tri + 1);
        const float4 v2 = __ldg(s.tris + 3 * (size_t)st.best_tri + 2);
        // two levels: the hit point and its barycentrics are taken in the object space the triangle was hit in, like
        // the reference's two-level kernel (intersect_bvh2level_skiplinks.cl:205-222)
        const Ray r = TWO_LEVEL ? hit_space_ray(s.instances, rw, st.best_inst) : rw;
        const float hx = add_(r.ox, mul_(st.closest, r.dx));
        const float hy = add_(r.oy, mul_(st.closest, r.dy));
        const float hz = add_(r.oz, mul_(st.closest, r.dz));
        const float2 uv = barycentrics_ref(hx, hy, hz, v0, v1, v2);
        h = make_int4(TWO_LEVEL ? st.best_sid : __float_as_int(v0.w), __float_as_int(v1.w), __float_as_int(uv.x),
                      __float_as_int(uv.y));
    }
    s.hit[px] = h;
}
// shadow ray idx / the reference's ShadowPass on its result: radiance += (throughput * L) * V
__device__ __forceinline__ void fetch_shadow_ray(const PtState& s, unsigned idx, Ray& r) {
    const float4 o = s.sh_o[idx], d = s.sh_d[idx];
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmax = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.mask = -1;
}
__device__ __forceinline__ void resolve_shadow(const PtState& s, unsigned idx, const TravState& st, int bounce) {
    const uint32_t px = (uint32_t)__float_as_int(s.sh_d[idx].w);
    const float V = st.best_tri >= 0 ? 0.0f : 1.0f;
    if (s.path_trace) s.path_trace[(size_t)bounce * ((size_t)s.width * s.height) + px].w = st.best_tri >= 0 ? 1 : 0;
    const float4 L = s.sh_L[idx];
    float4 r = s.radiance[px];
    r.x += L.x * V;
    r.y += L.y * V;
    r.z += L.z * V;
    s.radiance[px] = r;
}

// TWO_LEVEL: the world is a top-level tree over shapes with per-mesh bottom-level trees (rrb_commit with instances);
// separate instantiations, so the flat kernels keep their 64 registers / 8 CTAs per SM.
template <bool TWO_LEVEL, bool COOP = false>
__global__ void __launch_bounds__(128, TWO_LEVEL ? 6 : 8)
k_trace_closest(PtState s, int cur) {
    const unsigned count = s.cnt[cur];
    const uint32_t* __restrict__ queue = s.queue[cur];
    persistent_trace<false, false, TWO_LEVEL, COOP>(
        s.nodes, s.tris, count, s.cnt + 3,
        [&](unsigned idx, Ray& r) {
            fetch_closest_ray(s, queue, idx, r);
            return false;
        },
        [&](unsigned idx, const Ray& r, const TravState& st) { store_closest_hit<TWO_LEVEL>(s, queue, idx, r, st); },
        s.numNodes, s.stageTop, s.instances, coop_tail(s));
}

// any-hit traversal fused with the reference's ShadowPass; shadow rays of bounce parity `sh`
template <bool TWO_LEVEL, bool COOP = false>
__global__ void __launch_bounds__(128, TWO_LEVEL ? 6 : 8)
k_trace_shadow(PtState s, int sh, int bounce) {
    const unsigned count = s.cnt[5 + sh];
    persistent_trace<true, false, TWO_LEVEL, COOP>(
        s.nodes, s.tris, count, s.cnt + 4,
        [&](unsigned idx, Ray& r) {
            fetch_shadow_ray(s, idx, r);
            return true;
        },
        [&](unsigned idx, const Ray&, const TravState& st) { resolve_shadow(s, idx, st, bounce); }, s.numNodes, s.stageTop,
        s.instances, coop_tail(s));
}

// One launch for the two ray sets that are ready after a shade stage: the closest-hit rays of the next bounce
// (queue `cur`) and the shadow rays of the bounce just shaded (parity `sh`).  They are independent of each other, so
// sharing the persistent warps removes one drain-out tail and one launch per bounce.  Long rays first.
template <bool TWO_LEVEL, bool COOP = false>
__global__ void __launch_bounds__(128, TWO_LEVEL ? 6 : 8)
k_trace_mixed(PtState s, int cur, int sh, int sh_bounce) {
    const unsigned n_closest = s.cnt[cur], n_shadow = s.cnt[5 + sh];
    const uint32_t* __restrict__ queue = s.queue[cur];
    persistent_trace<false, true, TWO_LEVEL, COOP>(
        s.nodes, s.tris, n_closest + n_shadow, s.cnt + 3,
        [&](unsigned idx, Ray& r) {
            if (idx < n_closest) {
                fetch_closest_ray(s, queue, idx, r);
                return false;
            }
            fetch_shadow_ray(s, idx - n_closest, r);
            return true;
        },
        [&](unsigned idx, const Ray& r, const TravState& st) {
            if (idx < n_closest) store_closest_hit<TWO_LEVEL>(s, queue, idx, r, st);
            else resolve_shadow(s, idx - n_closest, st, sh_bounce);
        },
        s.numNodes, s.stageTop, s.instances, coop_tail(s));
}

// ---------------------------------------------------------------------------------------------
// shade (PathTracing kernel body) over the live queue, with compaction of both outputs
// ---------------------------------------------------------------------------------------------
// Block size of the shade kernel, and an optional phase synchronisation of its warps (pt_shade.cuh).
// The shading code is ~170 KB of SASS, far beyond the 32 KB L1.5 instruction cache (ncu r01:
// gcc__cache_requests_type_instruction at 84 % of peak, stall no_instructions 22 %).  Measured on
// C3 (profiles/r01_shade_variants.md): walking the phases in lock-step per CTA costs more in
// barrier waits than it saves in instruction fetches (2.25 ms vs 2.00 ms per step), so it is off.
#ifndef PT_SHADE_BLOCK
#define PT_SHADE_BLOCK 512
#endif
#ifndef PT_SHADE_SYNC
#define PT_SHADE_SYNC 0
#endif
#ifndef PT_SHADE_CTAS_PER_SM
#define PT_SHADE_CTAS_PER_SM (512 / PT_SHADE_BLOCK > 0 ? 512 / PT_SHADE_BLOCK : 1)   // 16 resident warps at ~115 registers
#endif
#ifndef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_MIN_BLOCKS 1      // register cap = 64K / (PT_SHADE_BLOCK * PT_SHADE_MIN_BLOCKS)
#endif
// What a shaded vertex leaves behind: the diagnostic record, the shadow-queue entry, the entry of the next live queue
// (ordered per `s.coherence`) and the path state of the pixel.  Called by every thread of the CTA (barriers / ballots).
template <bool EST>
__device__ __forceinline__ void push_vertex_results(const PtState& s, int cur, int frame, int bounce, int cap, uint32_t px,
                                                    bool in_range, bool active, bool pushShadow, bool pushNext,
                                                    const ShadeOut& out) {
    if (s.path_trace && in_range) {
        const int4 h = s.hit[px];
        const int flags = (active ? 1 : 0) | (pushShadow ? 2 : 0) | (active && out.hasEmission ? 4 : 0) |
                          (pushNext ? 8 : 0) | (active ? (out.bsdfFlags << 8) : 0);
        s.path_trace[(size_t)bounce * ((size_t)s.width * s.height) + px] = make_int4(h.x, h.y, flags, 0);
    }
    const unsigned si = warp_push(pushShadow, s.cnt + 5 + (bounce & 1));
    if (pushShadow) {
        s.sh_o[si] = make_float4(out.shadow.o.x, out.shadow.o.y, out.shadow.o.z, out.shadow.tmax);
        s.sh_d[si] = make_float4(out.shadow.d.x, out.shadow.d.y, out.shadow.d.z, __int_as_float((int)px));
        s.sh_L[si] = make_float4(out.direct.x, out.direct.y, out.direct.z, 0.0f);
    }
    if (s.coherence == 3) {
        // CTA-local bucketing: the survivors of this CTA's chunk of the queue go to ONE contiguous chunk of the
        // next queue, ordered by coherence key inside it.  Neighbouring pixels stay neighbours (texture and
        // geometry locality of the next bounce), warps of the next shade launch still see runs of equal keys.
        __shared__ unsigned s_cnt[PT_NUM_BUCKETS], s_pre[PT_NUM_BUCKETS], s_base;
        if (threadIdx.x < PT_NUM_BUCKETS) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        const unsigned key = pushNext ? coherence_key(s, px, frame, bounce + 1) : 0u;
        const unsigned rank = warp_push_keyed(pushNext, key, s_cnt);      // shared-memory atomics
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned sum = 0;
            for (int k = 0; k < PT_NUM_BUCKETS; ++k) {
                s_pre[k] = sum;
                sum += s_cnt[k];
            }
            s_base = sum ? atomicAdd(s.cnt + (1 - cur), sum) : 0u;
        }
        __syncthreads();
        if (pushNext) s.queue[1 - cur][s_base + s_pre[key] + rank] = px;
    } else if (s.coherence >= 1) {   // next bounce's queue goes to the staging regions, by coherence key
        const unsigned key = pushNext ? coherence_key(s, px, frame, bounce + 1) : 0u;
        const unsigned slot = warp_push_keyed(pushNext, key, s.cnt + 8);
        if (pushNext) s.bucket[(size_t)key * (size_t)cap + slot] = px;
    } else {
        const unsigned ni = warp_push(pushNext, s.cnt + (1 - cur));
        if (pushNext) s.queue[1 - cur][ni] = px;
    }
    if (pushNext) {
        s.ray_o[px] = make_float4(out.nextO.x, out.nextO.y, out.nextO.z, RT_MAX_TRACE_DISTANCE);
        s.ray_d[px] = make_float4(out.nextD.x, out.nextD.y, out.nextD.z, EST ? out.nextPdf : 0.0f);
        if (EST && out.nextPdf > 0.0f)
            s.direct_weight[px] = make_float4(out.nextDirectWeight.x, out.nextDirectWeight.y, out.nextDirectWeight.z, 0.0f);
        s.throughput[px] = make_float4(out.throughput.x, out.throughput.y, out.throughput.z,
                                       __int_as_float(out.bsdfFlags));
    }
}

template <bool EST>
__device__ __forceinline__ void shade_queue(const PtState& s, int cur, int frame, int bounce, int maxDepth, int cap,
                                            int estimator) {
    const unsigned count = s.cnt[cur];
    const Scene sc = make_scene(s);
    // CTA-uniform trip count: every thread reaches the barriers inside shade_vertex and the ballots below
    for (unsigned cta_base = blockIdx.x * blockDim.x; cta_base < count; cta_base += gridDim.x * blockDim.x) {
        const unsigned i = cta_base + threadIdx.x;
        bool pushNext = false, pushShadow = false, active = false;
        ShadeOut out;
        ShadeIn in;
        uint32_t px = 0;
        in.pixel = 0;
        in.shapeId = in.primId = 0;
        in.bu = in.bv = 0.0f;
        in.rayDir = in.throughput = v3(0.0f);
        in.prevBsdfFlags = 0;
        in.prevPdf = 0.0f;
        in.prevOrigin = in.prevDirectWeight = v3(0.0f);
        if (i < count) {
            px = s.queue[cur][i];
            const int4 h = s.hit[px];
            if (h.x != -1 && h.y != -1) {
                active = true;
                in.pixel = (int)px;
                in.shapeId = h.x;
                in.primId = h.y;
                in.bu = __int_as_float(h.z);
                in.bv = __int_as_float(h.w);
                const float4 d = s.ray_d[px];
                in.rayDir = v3(d.x, d.y, d.z);
                const float4 tp = s.throughput[px];
                in.throughput = v3(tp.x, tp.y, tp.z);
                in.prevBsdfFlags = __float_as_int(tp.w);
                if (EST) {
                    in.prevPdf = d.w;
                    const float4 o = s.ray_o[px];
                    in.prevOrigin = v3(o.x, o.y, o.z);
                    if (d.w > 0.0f) {
                        const float4 dw = s.direct_weight[px];
                        in.prevDirectWeight = v3(dw.x, dw.y, dw.z);
                    }
                }
            }
        }
        shade_vertex<PT_SHADE_SYNC != 0, EST>(sc, in, active, frame, s.width, s.height, bounce, maxDepth, &out, estimator);
        if (active) {
            // hasEmission: the reference's emission branch, which ends the path; with EST a weighted emission can
            // also come with a path that goes on
            if (out.hasEmission || (EST && (out.emitted.x != 0.0f || out.emitted.y != 0.0f || out.emitted.z != 0.0f))) {
                float4 r = s.radiance[px];
                r.x += out.emitted.x;
                r.y += out.emitted.y;
                r.z += out.emitted.z;
                s.radiance[px] = r;
            }
            if (!out.hasEmission) {
                // a zero contribution adds nothing whatever the visibility: no shadow ray needed
                pushShadow = out.shadow.written &&
                             !(out.direct.x == 0.0f && out.direct.y == 0.0f && out.direct.z == 0.0f);
                pushNext = out.continues;
            }
        }
        push_vertex_results<EST>(s, cur, frame, bounce, cap, px, i < count, active, pushShadow, pushNext, out);
    }
}

__global__ void __launch_bounds__(PT_SHADE_BLOCK, PT_SHADE_MIN_BLOCKS)
k_shade(PtState s, int cur, int frame, int bounce, int maxDepth, int cap) {
    shade_queue<false>(s, cur, frame, bounce, maxDepth, cap, 0);
}
// the same with the opt-in estimator switches (PtState::estimator != 0)
__global__ void __launch_bounds__(PT_SHADE_BLOCK, PT_SHADE_MIN_BLOCKS)
k_shade_est(PtState s, int cur, int frame, int bounce, int maxDepth, int cap) {
    shade_queue<true>(s, cur, frame, bounce, maxDepth, cap, s.estimator);
}

// ---------------------------------------------------------------------------------------------
// shade in two stages (ptb_set_shade_split; pt_shade.cuh: shade_surface / shade_scatter)
// ---------------------------------------------------------------------------------------------
// k_shade_surface: one thread per queue entry -- drops the rays that missed, ends the paths that hit an emitter
// (radiance += throughput * Le), and for every other vertex does the gather-heavy half (geometry, normal map, material
// textures) and writes a 112-byte record.  A CTA's records go to ONE contiguous run of the record array in queue
// order (warp ballots + one atomic per CTA and iteration), so the second stage reads them coalesced and the next
// queue keeps neighbouring pixels together.  Record layout: 7 float4 planes of `vtx_stride` entries,
//   0: p | pixel   1: gn | flags   2: sn | opacity.x   3: sdpdu | opacity.y   4: sdpdv | opacity.z   5: Kd | key   6: Ks | -
#ifndef PT_SURF_BLOCK
#define PT_SURF_BLOCK 256
#endif
#ifndef PT_SURF_MIN_BLOCKS
#define PT_SURF_MIN_BLOCKS 3       // register cap 85: 24 warps per SM
#endif
__global__ void __launch_bounds__(PT_SURF_BLOCK, PT_SURF_MIN_BLOCKS)
k_shade_surface(PtState s, int cur, int frame, int bounce) {
    const unsigned count = s.cnt[cur];
    const Scene sc = make_scene(s);
    __shared__ unsigned s_warp[PT_SURF_BLOCK / 32], s_base;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t P = (size_t)s.width * s.height;
    const size_t stride = (size_t)s.vtx_stride;
    for (unsigned cta_base = blockIdx.x * blockDim.x; cta_base < count; cta_base += gridDim.x * blockDim.x) {
        const unsigned i = cta_base + threadIdx.x;
        int kind = SURF_DEAD;
        VertexRecord rec;
        uint32_t px = 0;
        if (i < count) {
            px = s.queue[cur][i];
            const int4 h = s.hit[px];
            int trace_flags = 0;
            if (h.x != -1 && h.y != -1) {
                const float4 d = s.ray_d[px];
                // the throughput and the previous lobe flags matter here only when the shape is an emitter
                float4 tp = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
                const bool emitter = sc.shapes[h.x].lightID != RT_INVALID_ID;
                if (emitter) tp = s.throughput[px];
                V3 Le;
                kind = shade_surface(sc, h.x, h.y, __int_as_float(h.z), __int_as_float(h.w), v3(d.x, d.y, d.z),
                                     __float_as_int(tp.w), bounce, px, frame, s.width, s.height, &Le, &rec);
                if (kind == SURF_EMITTED) {
                    const V3 t = bounce == 0 ? v3(1.0f) : v3(tp.x, tp.y, tp.z);
                    const V3 e = t * Le;
                    float4 r = s.radiance[px];
                    r.x += e.x;
                    r.y += e.y;
                    r.z += e.z;
                    s.radiance[px] = r;
                    trace_flags = 1 | 4 | (__float_as_int(tp.w) << 8);
                }
            }
            if (s.path_trace && kind != SURF_RECORD)
                s.path_trace[(size_t)bounce * P + px] = make_int4(h.x, h.y, trace_flags, 0);
        }
        const bool has = kind == SURF_RECORD;
        const unsigned m = __ballot_sync(0xffffffffu, has);
        if (lane == 0) s_warp[warp] = (unsigned)__popc(m);
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned sum = 0;
            for (int w = 0; w < PT_SURF_BLOCK / 32; ++w) {
                const unsigned t = s_warp[w];
                s_warp[w] = sum;
                sum += t;
            }
            s_base = sum ? atomicAdd(s.cnt + 7, sum) : 0u;
        }
        __syncthreads();
        if (has) {
            float4* v = s.vtx + (s_base + s_warp[warp] + (unsigned)__popc(m & ((1u << lane) - 1u)));
            v[0 * stride] = make_float4(rec.p.x, rec.p.y, rec.p.z, __int_as_float((int)rec.px));
            v[1 * stride] = make_float4(rec.gn.x, rec.gn.y, rec.gn.z, __int_as_float(rec.flags));
            v[2 * stride] = make_float4(rec.sn.x, rec.sn.y, rec.sn.z, rec.opacity.x);
            v[3 * stride] = make_float4(rec.sdpdu.x, rec.sdpdu.y, rec.sdpdu.z, rec.opacity.y);
            v[4 * stride] = make_float4(rec.sdpdv.x, rec.sdpdv.y, rec.sdpdv.z, rec.opacity.z);
            v[5 * stride] = make_float4(rec.Kd.x, rec.Kd.y, rec.Kd.z, __int_as_float(rec.key));
            v[6 * stride] = make_float4(rec.Ks.x, rec.Ks.y, rec.Ks.z, 0.0f);
        }
        __syncthreads();      // s_warp / s_base are rewritten by the next iteration
    }
}

// k_shade_scatter: one thread per record -- light sampling, BSDF evaluation, lobe sampling, the pushes of k_shade.
// This half gathers nothing but its record and the pixel's ray / throughput and is bound by dependent arithmetic
// (ncu r02l: stall `wait` 30 %, 4 warps per scheduler at 121 registers), so -- unlike the single kernel, where a lower
// register cap only added spills to a kernel that waits for memory -- it gains from more resident warps: capped at
// 85 registers (3 x 256 threads per SM, 24 warps) it is 3 % faster than at 121 (profiles/r02_full_frame.md).
#ifndef PT_SCAT_BLOCK
#define PT_SCAT_BLOCK 256
#endif
#ifndef PT_SCAT_MIN_BLOCKS
#define PT_SCAT_MIN_BLOCKS 3
#endif
__global__ void __launch_bounds__(PT_SCAT_BLOCK, PT_SCAT_MIN_BLOCKS)
k_shade_scatter(PtState s, int cur, int frame, int bounce, int maxDepth, int cap) {
    const unsigned count = s.cnt[7];
    const Scene sc = make_scene(s);
    const size_t stride = (size_t)s.vtx_stride;
    for (unsigned cta_base = blockIdx.x * blockDim.x; cta_base < count; cta_base += gridDim.x * blockDim.x) {
        const unsigned i = cta_base + threadIdx.x;
        const bool active = i < count;
        bool pushNext = false, pushShadow = false;
        ShadeOut out;
        out.hasEmission = false;
        out.bsdfFlags = 0;
        uint32_t px = 0;
        if (active) {
            const float4* v = s.vtx + i;
            const float4 r0 = v[0 * stride], r1 = v[1 * stride], r2 = v[2 * stride], r3 = v[3 * stride],
                         r4 = v[4 * stride], r5 = v[5 * stride], r6 = v[6 * stride];
            VertexRecord rec;
            rec.p = v3(r0.x, r0.y, r0.z);
            rec.px = (uint32_t)__float_as_int(r0.w);
            rec.gn = v3(r1.x, r1.y, r1.z);
            rec.flags = __float_as_int(r1.w);
            rec.sn = v3(r2.x, r2.y, r2.z);
            rec.sdpdu = v3(r3.x, r3.y, r3.z);
            rec.sdpdv = v3(r4.x, r4.y, r4.z);
            rec.opacity = v3(r2.w, r3.w, r4.w);
            rec.Kd = v3(r5.x, r5.y, r5.z);
            rec.Ks = v3(r6.x, r6.y, r6.z);
            px = rec.px;
            const float4 d = s.ray_d[px];
            const float4 tp = s.throughput[px];
            shade_scatter(sc, rec, v3(d.x, d.y, d.z), v3(tp.x, tp.y, tp.z), __float_as_int(tp.w), frame, s.width, s.height,
                          bounce, maxDepth, &out);
            // a zero contribution adds nothing whatever the visibility: no shadow ray needed
            pushShadow = out.shadow.written && !(out.direct.x == 0.0f && out.direct.y == 0.0f && out.direct.z == 0.0f);
            pushNext = out.continues;
        }
        push_vertex_results<false>(s, cur, frame, bounce, cap, px, active, active, pushShadow, pushNext, out);
    }
}

// k_shade_scatter_sorted: the same work in an order that keeps a warp's lanes together.  With dense records the
// second stage is bound by its instruction stream, and ncu counts 16 of 32 lanes active per instruction (r02l): the
// lanes of a warp sample different light types and different lobes.  This stage gathers nothing but its own records
// and the pixel's ray / throughput, so -- unlike the single kernel's consumer-side sort (mode 4), which lost more to
// texture / geometry locality than it won -- the order of the ARITHMETIC is free, as long as the MEMORY traffic keeps
// the record order (a first version that also gathered and scattered in sorted order was slower: 25 of 32 lanes,
// 33 % fewer warp instructions, +15 % time).  So a CTA
//   1. brings a CHUNK of consecutive records into shared memory with one bulk asynchronous copy per record plane
//      (cp.async.bulk, completion on an mbarrier), and the pixels' ray directions / throughputs beside them;
//   2. counting-sorts the chunk's record indices by the 7-bit key the first stage left in the record (scatter_key);
//   3. shades in sorted order out of shared memory, leaving the next ray and throughput in the record's own slot;
//   4. writes the survivors' path state and the next queue's entries in RECORD order -- one contiguous run per chunk,
//      neighbouring pixels stay neighbours for the next bounce's surface stage.
#define PT_SCAT_CHUNK 1024
#define PT_SCAT_PER_THREAD (PT_SCAT_CHUNK / PT_SHADE_BLOCK)
#define PT_SCAT_KEYS 128
#define PT_SCAT_PLANES 9        // 7 record planes + ray direction + throughput
static_assert(PT_SCAT_CHUNK % PT_SHADE_BLOCK == 0 && PT_SCAT_KEYS <= PT_SHADE_BLOCK && PT_SCAT_CHUNK <= 65536, "scatter sort geometry");
static size_t scatter_sorted_smem() {
    return (size_t)PT_SCAT_PLANES * PT_SCAT_CHUNK * 16 + PT_SCAT_CHUNK * 2 + PT_SCAT_CHUNK + PT_SCAT_KEYS * 4 +
           (4 + PT_SHADE_BLOCK / 32) * 4 + 8;
}
__global__ void __launch_bounds__(PT_SHADE_BLOCK, PT_SHADE_MIN_BLOCKS)
k_shade_scatter_sorted(PtState s, int cur, int frame, int bounce, int maxDepth) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* s_rec = (float4*)smem;                                            // [PLANES][CHUNK]
    unsigned short* s_idx = (unsigned short*)(s_rec + PT_SCAT_PLANES * PT_SCAT_CHUNK);   // sorted position -> record offset
    unsigned char* s_next = (unsigned char*)(s_idx + PT_SCAT_CHUNK);          // record offset -> the path goes on
    unsigned* s_bin = (unsigned*)(s_next + PT_SCAT_CHUNK);
    unsigned* s_misc = s_bin + PT_SCAT_KEYS;                                  // [0] queue base, [1] run counter, [4..] warp sums
    uint64_t* s_bar = (uint64_t*)(s_misc + 4 + PT_SHADE_BLOCK / 32);
    const unsigned count = s.cnt[7];
    const Scene sc = make_scene(s);
    const size_t stride = (size_t)s.vtx_stride;
    const size_t P = (size_t)s.width * s.height;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    unsigned chunk = PT_SCAT_CHUNK;
    const unsigned per_cta = (count + gridDim.x - 1) / gridDim.x;
    if (per_cta < PT_SCAT_CHUNK) chunk = max((unsigned)PT_SHADE_BLOCK, (per_cta + PT_SHADE_BLOCK - 1) / PT_SHADE_BLOCK * PT_SHADE_BLOCK);
    const unsigned nchunks = (count + chunk - 1) / chunk;
    if (threadIdx.x == 0) mbar_init(s_bar, 1);
    __syncthreads();
    unsigned parity = 0;
    for (unsigned c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const unsigned base = c * chunk;
        const unsigned n = min(chunk, count - base);
        // ---- 1. the chunk's records: one bulk copy per plane
        if (threadIdx.x == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the slots were written by the last chunk
            mbar_expect_tx(s_bar, 7u * n * 16u);
            for (int f = 0; f < 7; ++f)
                bulk_copy_g2s(s_rec + f * PT_SCAT_CHUNK, s.vtx + f * stride + base, n * 16u, s_bar);
        }
        if (threadIdx.x < PT_SCAT_KEYS) s_bin[threadIdx.x] = 0;
        if (threadIdx.x == 0) s_misc[1] = 0;      // run counter of step 3
        __syncthreads();
        mbar_wait(s_bar, parity);
        parity ^= 1u;
        // ---- 2. keys and ranks; the pixels' ray directions and throughputs (record order: neighbouring pixels)
        unsigned e_key[PT_SCAT_PER_THREAD], e_rank[PT_SCAT_PER_THREAD];
#pragma unroll
        for (int j = 0; j < PT_SCAT_PER_THREAD; ++j) {
            const unsigned e = threadIdx.x + j * PT_SHADE_BLOCK;
            e_key[j] = 0xffffffffu;
            if (e < n) {
                const uint32_t px = (uint32_t)__float_as_int(s_rec[0 * PT_SCAT_CHUNK + e].w);
                s_rec[7 * PT_SCAT_CHUNK + e] = s.ray_d[px];
                s_rec[8 * PT_SCAT_CHUNK + e] = s.throughput[px];
                s_next[e] = 0;
                e_key[j] = (unsigned)__float_as_int(s_rec[5 * PT_SCAT_CHUNK + e].w) & (PT_SCAT_KEYS - 1);
                e_rank[j] = atomicAdd(s_bin + e_key[j], 1u);
            }
        }
        __syncthreads();
        {   // exclusive scan of the bins
            unsigned v = threadIdx.x < PT_SCAT_KEYS ? s_bin[threadIdx.x] : 0u;
            unsigned incl = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)lane >= d) incl += t;
            }
            if (lane == 31 && warp < PT_SCAT_KEYS / 32) s_misc[4 + warp] = incl;
            __syncthreads();
            if (threadIdx.x < PT_SCAT_KEYS) {
                unsigned before = 0;
                for (unsigned w = 0; w < warp; ++w) before += s_misc[4 + w];
                s_bin[threadIdx.x] = before + incl - v;
            }
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < PT_SCAT_PER_THREAD; ++j)
            if (e_key[j] != 0xffffffffu) s_idx[s_bin[e_key[j]] + e_rank[j]] = (unsigned short)(threadIdx.x + j * PT_SHADE_BLOCK);
        __syncthreads();
        // ---- 3. shade in sorted order.  Warps PULL runs of 32 sorted records from a shared counter: the cost of a run
        // depends on its key (a microfacet lobe costs several Lambertian ones), a static assignment leaves the CTA
        // waiting at the barrier below for its slowest warp.
        for (;;) {
            unsigned r = 0;
            if (lane == 0) r = atomicAdd(s_misc + 1, 32u);
            r = __shfl_sync(0xffffffffu, r, 0);
            if (r >= n) break;
            const unsigned j = r + lane;
            const bool active = j < n;
            bool pushNext = false, pushShadow = false;
            ShadeOut out;
            out.hasEmission = false;
            out.bsdfFlags = 0;
            uint32_t px = 0;
            unsigned e = 0;
            if (active) {
                e = s_idx[j];
                const float4* v = s_rec + e;
                const float4 r0 = v[0 * PT_SCAT_CHUNK], r1 = v[1 * PT_SCAT_CHUNK], r2 = v[2 * PT_SCAT_CHUNK],
                             r3 = v[3 * PT_SCAT_CHUNK], r4 = v[4 * PT_SCAT_CHUNK], r5 = v[5 * PT_SCAT_CHUNK],
                             r6 = v[6 * PT_SCAT_CHUNK], d = v[7 * PT_SCAT_CHUNK], tp = v[8 * PT_SCAT_CHUNK];
                VertexRecord rec;
                rec.p = v3(r0.x, r0.y, r0.z);
                rec.px = (uint32_t)__float_as_int(r0.w);
                rec.gn = v3(r1.x, r1.y, r1.z);
                rec.flags = __float_as_int(r1.w);
                rec.sn = v3(r2.x, r2.y, r2.z);
                rec.sdpdu = v3(r3.x, r3.y, r3.z);
                rec.sdpdv = v3(r4.x, r4.y, r4.z);
                rec.opacity = v3(r2.w, r3.w, r4.w);
                rec.Kd = v3(r5.x, r5.y, r5.z);
                rec.Ks = v3(r6.x, r6.y, r6.z);
                rec.key = 0;
                px = rec.px;
                shade_scatter(sc, rec, v3(d.x, d.y, d.z), v3(tp.x, tp.y, tp.z), __float_as_int(tp.w), frame, s.width, s.height,
                              bounce, maxDepth, &out);
                pushShadow = out.shadow.written && !(out.direct.x == 0.0f && out.direct.y == 0.0f && out.direct.z == 0.0f);
                pushNext = out.continues;
                if (s.path_trace) {
                    const int4 h = s.hit[px];
                    const int flags = 1 | (pushShadow ? 2 : 0) | (pushNext ? 8 : 0) | (out.bsdfFlags << 8);
                    s.path_trace[(size_t)bounce * P + px] = make_int4(h.x, h.y, flags, 0);
                }
            }
            const unsigned si = warp_push(pushShadow, s.cnt + 5 + (bounce & 1));
            if (pushShadow) {
                s.sh_o[si] = make_float4(out.shadow.o.x, out.shadow.o.y, out.shadow.o.z, out.shadow.tmax);
                s.sh_d[si] = make_float4(out.shadow.d.x, out.shadow.d.y, out.shadow.d.z, __int_as_float((int)px));
                s.sh_L[si] = make_float4(out.direct.x, out.direct.y, out.direct.z, 0.0f);
            }
            if (pushNext) {      // the next path state waits in the record's own slot (plane 0 keeps the pixel)
                s_next[e] = 1;
                s_rec[1 * PT_SCAT_CHUNK + e] = make_float4(out.nextO.x, out.nextO.y, out.nextO.z, RT_MAX_TRACE_DISTANCE);
                s_rec[2 * PT_SCAT_CHUNK + e] = make_float4(out.nextD.x, out.nextD.y, out.nextD.z, 0.0f);
                s_rec[3 * PT_SCAT_CHUNK + e] = make_float4(out.throughput.x, out.throughput.y, out.throughput.z,
                                                           __int_as_float(out.bsdfFlags));
            }
        }
        __syncthreads();
        // ---- 4. survivors in record order: thread t owns PER_THREAD consecutive records
        {
            const unsigned first = threadIdx.x * PT_SCAT_PER_THREAD;
            unsigned mine = 0;
#pragma unroll
            for (int j = 0; j < PT_SCAT_PER_THREAD; ++j) mine += (first + j < n && s_next[first + j]) ? 1u : 0u;
            unsigned incl = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)lane >= d) incl += t;
            }
            if (lane == 31) s_misc[4 + warp] = incl;
            __syncthreads();
            if (threadIdx.x == 0) {
                unsigned sum = 0;
                for (int w = 0; w < PT_SHADE_BLOCK / 32; ++w) {
                    const unsigned t = s_misc[4 + w];
                    s_misc[4 + w] = sum;
                    sum += t;
                }
                s_misc[0] = sum ? atomicAdd(s.cnt + (1 - cur), sum) : 0u;
            }
            __syncthreads();
            unsigned at = s_misc[0] + s_misc[4 + warp] + incl - mine;
#pragma unroll
            for (int j = 0; j < PT_SCAT_PER_THREAD; ++j) {
                const unsigned e = first + j;
                if (e < n && s_next[e]) {
                    const uint32_t px = (uint32_t)__float_as_int(s_rec[0 * PT_SCAT_CHUNK + e].w);
                    s.queue[1 - cur][at++] = px;
                    s.ray_o[px] = s_rec[1 * PT_SCAT_CHUNK + e];
                    s.ray_d[px] = s_rec[2 * PT_SCAT_CHUNK + e];
                    s.throughput[px] = s_rec[3 * PT_SCAT_CHUNK + e];
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// shade, sorted (coherence mode 4; opt-in -- measured slower than mode 3 on the benchmark, see the end of this comment)
// ---------------------------------------------------------------------------------------------
// One CTA takes a CHUNK of consecutive queue entries (neighbouring pixels: they share geometry and texels), SORTS it
// in shared memory and shades it in sorted order, warp by warp:
//   * entries whose ray missed are dropped before any shading code runs (in k_shade they idle through all of it);
//   * the key is what the lanes of a warp would otherwise diverge on -- the TYPE of the light the vertex will sample
//     and the half of its lobe-selection variate, both known from the sampler alone (samplers.cl:75-84: a pure
//     function of pixel, frame, bounce), major; the material id, minor (a warp's texel fetches go to one texture set).
// r01 measured 21 of 32 lanes active on k_shade even on the coherent first bounce, where the queue is in pixel order
// and every warp walks all three light types and both lobes.  Sorting at CONSUMPTION (here) instead of at production
// (r01's mode 3: buckets inside the producer's 512-entry chunk, whose 8 boundaries cut through half of its 16 warps)
// knows hit / miss and the material, covers bounce 0, and needs no key evaluation in the producer.
// The small scene tables every vertex reads -- shapes, materials, lights, texture descriptors, per-material alpha --
// are copied into shared memory once per CTA by bulk asynchronous copies (cp.async.bulk, completion on an mbarrier)
// when they fit; the reference copies a 160-byte RTShape and a 128-byte RTMaterial by value several times per vertex
// (geometry.cl:181, lights.cl:104, materials.cl:78).
// Survivors are collected in shared memory and appended to the next queue as ONE contiguous run per chunk, so
// neighbouring pixels stay neighbours.  Per-path results do not depend on the order (bitwise: the coherence tests).
// MEASURED (C3 1080p, profiles/r02_shade.md): the sort does what it was built for -- 26.7 of 32 lanes active instead
// of 17.9, 29 % fewer warp instructions per launch -- and the launch still takes LONGER (219 vs 189 us, bounce-1 half
// frame): a warp's lanes now come from all over the 2048-pixel chunk, the L1 hit rate falls from 53 % to 31 % and the
// exposed memory latency (long_scoreboard + barrier + no_instructions: 62 % of the stall samples) grows by more than
// the instruction stream shrinks; shade 1.94 -> 2.29 ms per step.  Staging the tables changes nothing measurable
// (2.34 vs 2.41 ms without / with).  So it is an option (ptb_set_coherence(4)), not the default.
#ifndef PT_SORT_CHUNK
#define PT_SORT_CHUNK 2048
#endif
#define PT_SORT_PER_THREAD (PT_SORT_CHUNK / PT_SHADE_BLOCK)
#define PT_SORT_KEYS 256        // (light type 4) x (lobe half 2) x (material id mod 32)
static_assert(PT_SORT_CHUNK % PT_SHADE_BLOCK == 0 && PT_SORT_KEYS <= PT_SHADE_BLOCK, "sort geometry");

// byte sizes of the staged tables, rounded to the 16 bytes a bulk copy moves
struct StagedTables {
    unsigned shapes, materials, lights, textures, alpha;
    __host__ __device__ unsigned total() const { return shapes + materials + lights + textures + alpha; }
};
__host__ __device__ inline StagedTables staged_tables(const PtState& s) {
    StagedTables t;
    t.shapes = (unsigned)s.numShapes * 160u;
    t.materials = (unsigned)s.numMaterials * 128u;
    t.lights = (unsigned)s.numLights * 80u;
    t.textures = (unsigned)s.numTextures * 16u;
    t.alpha = (((unsigned)s.numMaterials * 8u) + 15u) & ~15u;
    return t;
}
#define PT_STAGE_MAX_BYTES (48 * 1024)

template <bool EST>
__device__ __forceinline__ void shade_sorted(const PtState& s, int cur, int frame, int bounce, int maxDepth, int estimator,
                                             int stage_tables) {
    extern __shared__ __align__(16) unsigned char smem[];
    int4* s_hit = (int4*)smem;                                                   // [CHUNK] sorted hits
    uint32_t* s_px = (uint32_t*)(smem + PT_SORT_CHUNK * 16);                     // [CHUNK] sorted pixels
    uint32_t* s_next = s_px + PT_SORT_CHUNK;                                     // [CHUNK] survivors of the chunk
    unsigned* s_bin = s_next + PT_SORT_CHUNK;                                    // [KEYS] counts, then offsets
    unsigned* s_misc = s_bin + PT_SORT_KEYS;                                     // [0] live, [1] survivors, [2] queue base, [4..11] warp sums
    uint64_t* s_bar = (uint64_t*)(s_misc + 16);
    unsigned char* s_tab = (unsigned char*)(s_bar + 2);

    Scene sc = make_scene(s);
    if (stage_tables) {
        const StagedTables t = staged_tables(s);
        if (threadIdx.x == 0) mbar_init(s_bar, 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            mbar_expect_tx(s_bar, t.total());
            unsigned char* d = s_tab;
            if (t.shapes) bulk_copy_g2s(d, s.shapes, t.shapes, s_bar);
            d += t.shapes;
            if (t.materials) bulk_copy_g2s(d, s.materials, t.materials, s_bar);
            d += t.materials;
            if (t.lights) bulk_copy_g2s(d, s.lights, t.lights, s_bar);
            d += t.lights;
            if (t.textures) bulk_copy_g2s(d, s.textures2D, t.textures, s_bar);
            d += t.textures;
            if (t.alpha) bulk_copy_g2s(d, s.material_alpha, t.alpha, s_bar);
        }
        unsigned char* d = s_tab;
        sc.shapes = (const RTShape*)d;
        d += t.shapes;
        sc.materials = (const RTMaterial*)d;
        d += t.materials;
        sc.lights = (const RTLight*)d;
        d += t.lights;
        sc.textures2D = (const TextureDesc2D*)d;
        d += t.textures;
        sc.materialAlpha = (const float2*)d;
        sc.staged = 1;
        mbar_wait(s_bar, 0);
    }

    const unsigned count = s.cnt[cur];
    // chunk: full-size while there is a chunk for every CTA, else shrunk (in whole blocks) so that all SMs shade
    unsigned chunk = PT_SORT_CHUNK;
    const unsigned per_cta = (count + gridDim.x - 1) / gridDim.x;
    if (per_cta < PT_SORT_CHUNK) chunk = max((unsigned)PT_SHADE_BLOCK, (per_cta + PT_SHADE_BLOCK - 1) / PT_SHADE_BLOCK * PT_SHADE_BLOCK);
    const unsigned nchunks = (count + chunk - 1) / chunk;
    const uint32_t* __restrict__ queue = s.queue[cur];
    const size_t P = (size_t)s.width * s.height;

    for (unsigned c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const unsigned base = c * chunk;
        const unsigned n = min(chunk, count - base);
        if (threadIdx.x < PT_SORT_KEYS) s_bin[threadIdx.x] = 0;
        if (threadIdx.x == 0) s_misc[1] = 0;
        __syncthreads();
        // ---- 1. keys and ranks
        uint32_t e_px[PT_SORT_PER_THREAD];
        int4 e_hit[PT_SORT_PER_THREAD];
        unsigned e_key[PT_SORT_PER_THREAD], e_rank[PT_SORT_PER_THREAD];
#pragma unroll
        for (int j = 0; j < PT_SORT_PER_THREAD; ++j) {
            const unsigned e = threadIdx.x + j * PT_SHADE_BLOCK;
            e_key[j] = 0xffffffffu;
            if (e < n) {
                const uint32_t px = queue[base + e];
                const int4 h = s.hit[px];
                e_px[j] = px;
                e_hit[j] = h;
                if (h.x != -1 && h.y != -1) {
                    Sampler sp = makeSampler(s.sobol, (int)px, frame, s.width, s.height, bounce);
                    uint32_t lightIdx = ((uint32_t)floorf(getSample1D(&sp) * s.numLights));
                    lightIdx %= (uint32_t)s.numLights;
                    skipSamples(&sp, 2);
                    const unsigned lobe = getSample1D(&sp) >= 0.5f ? 1u : 0u;
                    const unsigned ltype = (unsigned)sc.lights[lightIdx].type & 3u;
                    const unsigned mat = (unsigned)sc.shapes[h.x].materialId & 31u;
                    e_key[j] = ((ltype * 2u + lobe) << 5) | mat;
                    e_rank[j] = atomicAdd(s_bin + e_key[j], 1u);
                } else if (s.path_trace) {
                    s.path_trace[(size_t)bounce * P + px] = make_int4(h.x, h.y, 0, 0);
                }
            }
        }
        __syncthreads();
        // ---- 2. exclusive scan of the bins (one value per thread of the first KEYS threads)
        {
            const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
            unsigned v = threadIdx.x < PT_SORT_KEYS ? s_bin[threadIdx.x] : 0u;
            unsigned incl = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)lane >= d) incl += t;
            }
            if (lane == 31 && warp < PT_SORT_KEYS / 32) s_misc[4 + warp] = incl;
            __syncthreads();
            if (threadIdx.x < PT_SORT_KEYS) {
                unsigned before = 0;
                for (unsigned w = 0; w < warp; ++w) before += s_misc[4 + w];
                s_bin[threadIdx.x] = before + incl - v;
                if (threadIdx.x == PT_SORT_KEYS - 1) s_misc[0] = before + incl;
            }
            __syncthreads();
        }
        // ---- 3. scatter into sorted order
#pragma unroll
        for (int j = 0; j < PT_SORT_PER_THREAD; ++j)
            if (e_key[j] != 0xffffffffu) {
                const unsigned pos = s_bin[e_key[j]] + e_rank[j];
                s_px[pos] = e_px[j];
                s_hit[pos] = e_hit[j];
            }
        __syncthreads();
        const unsigned n_live = s_misc[0];
        // ---- 4. shade, a block of sorted entries at a time (whole warps drop out at the tail)
        for (unsigned r = (threadIdx.x & ~31u); r < n_live; r += PT_SHADE_BLOCK) {
            const unsigned j = r + (threadIdx.x & 31u);
            const bool active = j < n_live;
            bool pushNext = false, pushShadow = false;
            ShadeOut out;
            ShadeIn in;
            uint32_t px = 0;
            in.pixel = 0;
            in.shapeId = in.primId = 0;
            in.bu = in.bv = 0.0f;
            in.rayDir = in.throughput = v3(0.0f);
            in.prevBsdfFlags = 0;
            in.prevPdf = 0.0f;
            in.prevOrigin = in.prevDirectWeight = v3(0.0f);
            if (active) {
                px = s_px[j];
                const int4 h = s_hit[j];
                in.pixel = (int)px;
                in.shapeId = h.x;
                in.primId = h.y;
                in.bu = __int_as_float(h.z);
                in.bv = __int_as_float(h.w);
                const float4 d = s.ray_d[px];
                in.rayDir = v3(d.x, d.y, d.z);
                const float4 tp = s.throughput[px];
                in.throughput = v3(tp.x, tp.y, tp.z);
                in.prevBsdfFlags = __float_as_int(tp.w);
                if (EST) {
                    in.prevPdf = d.w;
                    const float4 o = s.ray_o[px];
                    in.prevOrigin = v3(o.x, o.y, o.z);
                    if (d.w > 0.0f) {
                        const float4 dw = s.direct_weight[px];
                        in.prevDirectWeight = v3(dw.x, dw.y, dw.z);
                    }
                }
            }
            shade_vertex<false, EST>(sc, in, active, frame, s.width, s.height, bounce, maxDepth, &out, estimator);
            if (active) {
                if (out.hasEmission || (EST && (out.emitted.x != 0.0f || out.emitted.y != 0.0f || out.emitted.z != 0.0f))) {
                    float4 rr = s.radiance[px];
                    rr.x += out.emitted.x;
                    rr.y += out.emitted.y;
                    rr.z += out.emitted.z;
                    s.radiance[px] = rr;
                }
                if (!out.hasEmission) {
                    pushShadow = out.shadow.written &&
                                 !(out.direct.x == 0.0f && out.direct.y == 0.0f && out.direct.z == 0.0f);
                    pushNext = out.continues;
                }
                if (s.path_trace) {
                    const int flags = 1 | (pushShadow ? 2 : 0) | (out.hasEmission ? 4 : 0) | (pushNext ? 8 : 0) | (out.bsdfFlags << 8);
                    s.path_trace[(size_t)bounce * P + px] = make_int4(in.shapeId, in.primId, flags, 0);
                }
            }
            const unsigned si = warp_push(pushShadow, s.cnt + 5 + (bounce & 1));
            if (pushShadow) {
                s.sh_o[si] = make_float4(out.shadow.o.x, out.shadow.o.y, out.shadow.o.z, out.shadow.tmax);
                s.sh_d[si] = make_float4(out.shadow.d.x, out.shadow.d.y, out.shadow.d.z, __int_as_float((int)px));
                s.sh_L[si] = make_float4(out.direct.x, out.direct.y, out.direct.z, 0.0f);
            }
            const unsigned ni = warp_push(pushNext, s_misc + 1);      // shared-memory counter
            if (pushNext) {
                s_next[ni] = px;
                s.ray_o[px] = make_float4(out.nextO.x, out.nextO.y, out.nextO.z, RT_MAX_TRACE_DISTANCE);
                s.ray_d[px] = make_float4(out.nextD.x, out.nextD.y, out.nextD.z, EST ? out.nextPdf : 0.0f);
                if (EST && out.nextPdf > 0.0f)
                    s.direct_weight[px] = make_float4(out.nextDirectWeight.x, out.nextDirectWeight.y, out.nextDirectWeight.z, 0.0f);
                s.throughput[px] = make_float4(out.throughput.x, out.throughput.y, out.throughput.z,
                                               __int_as_float(out.bsdfFlags));
            }
        }
        __syncthreads();
        // ---- 5. the chunk's survivors become one contiguous run of the next queue
        const unsigned n_next = s_misc[1];
        if (threadIdx.x == 0) s_misc[2] = n_next ? atomicAdd(s.cnt + (1 - cur), n_next) : 0u;
        __syncthreads();
        const unsigned qbase = s_misc[2];
        for (unsigned k = threadIdx.x; k < n_next; k += PT_SHADE_BLOCK) s.queue[1 - cur][qbase + k] = s_next[k];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(PT_SHADE_BLOCK, PT_SHADE_MIN_BLOCKS)
k_shade_sorted(PtState s, int cur, int frame, int bounce, int maxDepth, int stage_tables) {
    shade_sorted<false>(s, cur, frame, bounce, maxDepth, 0, stage_tables);
}
__global__ void __launch_bounds__(PT_SHADE_BLOCK, PT_SHADE_MIN_BLOCKS)
k_shade_sorted_est(PtState s, int cur, int frame, int bounce, int maxDepth, int stage_tables) {
    shade_sorted<true>(s, cur, frame, bounce, maxDepth, s.estimator, stage_tables);
}
// dynamic shared memory of the sorted shade kernels: sort arrays + bins + barrier + staged tables
static size_t shade_sorted_smem(unsigned table_bytes) {
    return (size_t)PT_SORT_CHUNK * 16 + 2 * (size_t)PT_SORT_CHUNK * 4 + PT_SORT_KEYS * 4 + 16 * 4 + 16 + table_bytes;
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

int launch_pt_sample(const PtState& s, const uint32_t* pixels, int n_pixels, int frame, int max_depth,
                     cudaStream_t stream, cudaError_t* err, StageTimer* timer, const PtOverlap* overlap) {
    int launches = 0;
    *err = cudaSuccess;
    if (n_pixels <= 0) return 0;
    const int sms = sm_count();
    // CTAs per SM of the persistent kernels; the environment overrides exist for co-scheduling experiments
    // and are read once.
    static const int trace_per_sm = env_int("MCRT_TRACE_CTAS_PER_SM", 8);
    static const int shade_per_sm = env_int("MCRT_SHADE_CTAS_PER_SM", PT_SHADE_CTAS_PER_SM);
    static const int surf_per_sm = env_int("MCRT_SURF_CTAS_PER_SM", PT_SURF_MIN_BLOCKS);
    static const int scat_per_sm = env_int("MCRT_SCAT_CTAS_PER_SM", PT_SCAT_MIN_BLOCKS);
    // 1 = k_shade_scatter_sorted (25 of 32 lanes active instead of 16, a third fewer warp instructions, 2 % faster than
    // record order at the same occupancy -- and 3 % slower than record order at 24 warps per SM, which its 151 KB of
    // shared memory rules out: profiles/r02_full_frame.md)
    static const int scatter_sort = env_int("MCRT_SCATTER_SORT", 0);
    static const int stagger = env_int("MCRT_OVERLAP_STAGGER", 0);   // measured: no gain from starting the halves a stage apart

    // the frame as one wavefront, or as two half-frame wavefronts on two streams (PtOverlap)
    struct Half {
        PtState s;
        int first, n;
        cudaStream_t stream;
    } half[2];
    const int nh = (overlap && n_pixels >= 2 * 32768) ? 2 : 1;
    half[0] = {s, 0, nh == 2 ? n_pixels / 2 : n_pixels, stream};
    if (nh == 2) {
        const int nA = half[0].n;
        PtState t = s;           // compact (queue-indexed) arrays and the counters of the second half
        t.queue[0] += nA;
        t.queue[1] += nA;
        t.sh_o += nA;
        t.sh_d += nA;
        t.sh_L += nA;
        t.cnt += PT_CNT_STRIDE;
        t.bucket += (size_t)PT_NUM_BUCKETS * nA;
        if (t.vtx) t.vtx += nA;      // record planes keep the full-frame stride
        if (t.strag) t.strag += 2 * (size_t)nA;      // a mixed launch serves up to 2 rays per pixel
        half[1] = {t, nA, n_pixels - nA, overlap->stream};
        cudaEventRecord(overlap->fork, stream);
        cudaStreamWaitEvent(overlap->stream, overlap->fork, 0);
    }
    int cur[2] = {0, 0};
    const bool two_level = s.instances != nullptr;
    const bool coop = !two_level && s.coopIters > 0 && s.strag != nullptr;      // kernels with the cooperative tail compiled in
    // the sorted shade kernel stages the small scene tables in shared memory when they fit
    const unsigned table_bytes = staged_tables(s).total();
    static const int stage_enabled = env_int("MCRT_STAGE_TABLES", 1);
    const int stage_tables = (stage_enabled && table_bytes > 0 && table_bytes <= PT_STAGE_MAX_BYTES) ? 1 : 0;
    const size_t sorted_smem = shade_sorted_smem(stage_tables ? table_bytes : 0);
    static bool smem_attr_set = false;
    if (!smem_attr_set) {
        const int max_smem = (int)shade_sorted_smem(PT_STAGE_MAX_BYTES);
        cudaFuncSetAttribute(k_shade_sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(k_shade_sorted_est, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        cudaFuncSetAttribute(k_shade_scatter_sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scatter_sorted_smem());
        smem_attr_set = true;
    }
    // The split shade stage pays for its second launch and its record traffic only on large wavefronts: 1 M entries
    // (a 1080p half frame) +4.4 % paths/s, 131 K entries (a 512^2 half frame) -12 % (profiles/r02_full_frame.md).
    static const int split_min = env_int("MCRT_SPLIT_MIN_ENTRIES", 393216);
    auto use_split = [&](int n) { return s.shade_split && !s.estimator && s.vtx != nullptr && (s.shade_split == 2 || n >= split_min); };
    auto blocks_trace = [&](int n) { return min(sms * trace_per_sm, (n + 127) / 128); };
    auto blocks_shade = [&](int n) { return min(sms * shade_per_sm, (n + PT_SHADE_BLOCK - 1) / PT_SHADE_BLOCK); };
    for (int h = 0; h < nh; ++h) {
        k_raygen<<<(half[h].n + 255) / 256, 256, 0, half[h].stream>>>(half[h].s, pixels, half[h].first, half[h].n, frame);
        ++launches;
    }
    if (s.numLights > 0) {   // RTPathTracingPass::update returns early without lights (:42-43); PathTracing.cl:72
        // bounce b: [closest-hit trace of b  +  any-hit trace / resolve of b - 1] -> shade b;  then the last shadow set
        for (int b = 0; b < max_depth; ++b) {
            for (int h = 0; h < nh; ++h) {
                const Half& H = half[h];
                const bool bucketed = s.coherence == 2 || (s.coherence == 1 && b >= 1);
                k_bounce_ctl<<<1, 32, 0, H.stream>>>(H.s.cnt, H.s.stats, cur[h], b, H.n, bucketed ? 1 : 0);
                if (bucketed) {
                    k_queue_gather<<<min(sms * 4, (H.n + 255) / 256), 256, 0, H.stream>>>(H.s, cur[h], H.n);
                    ++launches;
                }
                if (timer) timer->mark(h, H.stream);
                if (b == 0) {
                    // experiment knob: start the second half one stage behind the first
                    if (h == 1 && stagger) cudaStreamWaitEvent(H.stream, overlap->stagger, 0);
                    if (two_level) k_trace_closest<true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h]);
                    else if (coop) k_trace_closest<false, true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h]);
                    else k_trace_closest<false><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h]);
                    if (h == 0 && nh == 2 && stagger) cudaEventRecord(overlap->stagger, H.stream);
                } else {
                    if (two_level) k_trace_mixed<true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h], (b - 1) & 1, b - 1);
                    else if (coop) k_trace_mixed<false, true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h], (b - 1) & 1, b - 1);
                    else k_trace_mixed<false><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, cur[h], (b - 1) & 1, b - 1);
                }
                if (timer) timer->mark(h, H.stream, b == 0 ? StageTimer::TRACE_CLOSEST : StageTimer::TRACE_MIXED);
                if (s.coherence == 4) {
                    const int nb = min(sms, (H.n + PT_SHADE_BLOCK - 1) / PT_SHADE_BLOCK);
                    if (s.estimator)
                        k_shade_sorted_est<<<nb, PT_SHADE_BLOCK, sorted_smem, H.stream>>>(H.s, cur[h], frame, b, max_depth, stage_tables);
                    else
                        k_shade_sorted<<<nb, PT_SHADE_BLOCK, sorted_smem, H.stream>>>(H.s, cur[h], frame, b, max_depth, stage_tables);
                } else if (use_split(H.n)) {
                    const int nb = min(sms * surf_per_sm, (H.n + PT_SURF_BLOCK - 1) / PT_SURF_BLOCK);
                    k_shade_surface<<<nb, PT_SURF_BLOCK, 0, H.stream>>>(H.s, cur[h], frame, b);
                    if (timer) timer->mark(h, H.stream, StageTimer::SHADE);
                    if (scatter_sort)
                        k_shade_scatter_sorted<<<min(sms, (H.n + PT_SHADE_BLOCK - 1) / PT_SHADE_BLOCK), PT_SHADE_BLOCK, scatter_sorted_smem(), H.stream>>>(
                            H.s, cur[h], frame, b, max_depth);
                    else
                        k_shade_scatter<<<min(sms * scat_per_sm, (H.n + PT_SCAT_BLOCK - 1) / PT_SCAT_BLOCK), PT_SCAT_BLOCK, 0, H.stream>>>(
                            H.s, cur[h], frame, b, max_depth, H.n);
                    ++launches;
                } else if (s.estimator)
                    k_shade_est<<<blocks_shade(H.n), PT_SHADE_BLOCK, 0, H.stream>>>(H.s, cur[h], frame, b, max_depth, H.n);
                else
                    k_shade<<<blocks_shade(H.n), PT_SHADE_BLOCK, 0, H.stream>>>(H.s, cur[h], frame, b, max_depth, H.n);
                if (timer) timer->mark(h, H.stream, (s.coherence != 4 && use_split(H.n)) ? StageTimer::SHADE_SCATTER : StageTimer::SHADE);
                launches += 3;
                cur[h] = 1 - cur[h];
            }
        }
        for (int h = 0; h < nh; ++h) {
            const Half& H = half[h];
            k_sample_end<<<1, 32, 0, H.stream>>>(H.s.cnt, H.s.stats, max_depth - 1);
            if (timer) timer->mark(h, H.stream);
            if (two_level) k_trace_shadow<true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, (max_depth - 1) & 1, max_depth - 1);
            else if (coop) k_trace_shadow<false, true><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, (max_depth - 1) & 1, max_depth - 1);
            else k_trace_shadow<false><<<blocks_trace(H.n), 128, 0, H.stream>>>(H.s, (max_depth - 1) & 1, max_depth - 1);
            if (timer) timer->mark(h, H.stream, StageTimer::TRACE_SHADOW);
            launches += 2;
        }
    }
    if (nh == 2) {
        cudaEventRecord(overlap->join, overlap->stream);
        cudaStreamWaitEvent(stream, overlap->join, 0);
    }
    *err = cudaGetLastError();
    return launches;
}

}  // namespace mcrt
