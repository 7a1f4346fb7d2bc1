// bdpt_kernels.cu -- the bidirectional path tracer for sm_100a.
//
// What it computes is the reference's BDPT (assets/kernels/BDPT.cl:240-932, host loop
// renderPasses/RTBDPTPass.cpp:67-448): camera and light subpaths of up to D+2 / D+1 vertices, every
// (t, s) connection strategy of depth <= D, balance-heuristic MIS over all strategies, light-tracing
// strategies (t = 1) splatted to the pixel the light vertex projects to.  How it is organised is
// different:
//
//  * subpath extension is a wavefront over queues of live pixel indices (warp-ballot compaction),
//    traced by the persistent traversal loop shared with the path tracer -- dead subpaths cost
//    nothing (the reference re-dispatches all W*H work-items 2D+1 times and traces every slot);
//  * PrepareConnections is STRATEGY-parallel: one thread per (pixel, t, s), pixel fastest, instead of
//    one work-item looping over <= 26 strategies; vertices are depth-major so a warp's loads coalesce.
//    The sampler position of an s = 1 strategy (its draws follow those of the connectible camera
//    vertices before it) is recovered by skipping the stream ahead;
//  * connections with a non-zero unweighted contribution become compact records; only those are
//    traced (any-hit) and weighted.  The reference traces P * maxConn slots including stale ones;
//  * ConnectVertices is one thread per surviving record.  The MIS weight is evaluated on register
//    copies of the (at most four) vertices whose densities change, never by patching and restoring
//    vertices in global memory (BDPT.cl:756-872);
//  * the per-pixel sum uses native float atomics (the reference: CAS loop, BDPT.cl:654-669).
//    Summation order is unspecified in both.
#include "bdpt_kernels.h"

#include "pt_emit.cuh"
#include "pt_shade.cuh"
#include "pt_wavefront.cuh"

namespace mcrt {

enum { RT_BDPT_CAMERA_VERTEX = 0, RT_BDPT_LIGHT_VERTEX = 1, RT_BDPT_SURFACE_VERTEX = 2 };   // kernel_data.h:202-207
enum {                                                                                        // kernel_data.h:209-218
    VF_CONNECTIBLE = 1, VF_DELTA_LIGHT = 1 << 1, VF_DELTA = 1 << 2, VF_INFINITE_LIGHT = 1 << 3
};

// ---------------------------------------------------------------------------------------------
// vertex record: 8 x float4
//   q0  pdfFwd, pdfRev, bits(type | flags << 8), pdfPos
//   q1  p.xyz, traceErrorOffset          q2  gn.xyz, bits(lightIdx)      q3  sn.xyz, bits(materialIdx)
//   q4  sdpdu.xyz, uv.x                  q5  sdpdv.xyz, uv.y             q6  wo.xyz, -      q7  throughput.xyz, -
// The MIS chain walk reads q0 only; a density evaluation reads q0-q5; a BSDF evaluation q0-q6.
// ---------------------------------------------------------------------------------------------
struct BVertex {
    float pdfFwd, pdfRev, pdfPos;
    int type, flags, lightIdx, materialIdx;
    RTInteraction si;      // wo, p, uv, traceErrorOffset, gn, sn, sdpdu, sdpdv (dpdu/dpdv are not kept)
    V3 beta;
};

// The two vertex evaluators inline the whole uber material (texture-modulated properties + BSDF value or density)
// and have four call sites each: inlined, k_bdpt_prepare / k_bdpt_connect are 250 / 200 KB of SASS and stall on
// instruction fetch (ncu r01: no_instructions 35 % / 59 % of the samples).  Out of line they shrink to 156 / 73 KB
// (connect: 123 -> 67 registers), but the vertex then travels through local memory: C4 9.91 ms per frame against
// 9.67 ms inlined (same run), so inlining stays the default.
#ifndef BDPT_NOINLINE_EVAL
#define BDPT_NOINLINE_EVAL 0
#endif
#if BDPT_NOINLINE_EVAL
#define BD_EVAL static __device__ __noinline__
#else
#define BD_EVAL HD
#endif

HD bool vOnSurface(const BVertex& v) {     // BDPT.cl:39-42 isVertexOnSurface
    return isNotNearZero(v.si.gn.x) || isNotNearZero(v.si.gn.y) || isNotNearZero(v.si.gn.z);
}
HD bool vInfiniteLight(const BVertex& v) { return (v.flags & VF_INFINITE_LIGHT) != 0; }
HD bool vConnectible(const BVertex& v) { return (v.flags & VF_CONNECTIBLE) != 0; }
HD bool vIsLight(const BVertex& v) { return v.type == RT_BDPT_LIGHT_VERTEX || v.lightIdx != RT_INVALID_ID; }

HD float4 f4(V3 a, float w) { return make_float4(a.x, a.y, a.z, w); }

HD void store_vertex(float4* __restrict__ base, const BVertex& v) {
    base[0] = make_float4(v.pdfFwd, v.pdfRev, __int_as_float(v.type | (v.flags << 8)), v.pdfPos);
    base[1] = f4(v.si.p, v.si.traceErrorOffset);
    base[2] = f4(v.si.gn, __int_as_float(v.lightIdx));
    base[3] = f4(v.si.sn, __int_as_float(v.materialIdx));
    base[4] = f4(v.si.sdpdu, v.si.uv.x);
    base[5] = f4(v.si.sdpdv, v.si.uv.y);
    base[6] = f4(v.si.wo, 0.0f);
    base[7] = f4(v.beta, 0.0f);
}
// LEVEL 0: q0-q2 (densities, position, normal, light)   1: + q3-q5 (shading frame, uv, material)
//       2: + q6 (wo)                                     3: + q7 (throughput)
template <int LEVEL>
HD BVertex load_vertex(const float4* __restrict__ base) {
    BVertex v;
    const float4 q0 = base[0], q1 = base[1], q2 = base[2];
    v.pdfFwd = q0.x;
    v.pdfRev = q0.y;
    const int tf = __float_as_int(q0.z);
    v.type = tf & 0xFF;
    v.flags = tf >> 8;
    v.pdfPos = q0.w;
    v.si.p = v3(q1);
    v.si.traceErrorOffset = q1.w;
    v.si.gn = v3(q2);
    v.lightIdx = __float_as_int(q2.w);
    v.si.shapeIdx = 0;
    v.si.dpdu = v.si.dpdv = v3(0.0f);
    v.materialIdx = RT_INVALID_ID;
    v.si.sn = v.si.sdpdu = v.si.sdpdv = v.si.wo = v.beta = v3(0.0f);
    v.si.uv = v2(0.0f, 0.0f);
    if (LEVEL >= 1) {
        const float4 q3 = base[3], q4 = base[4], q5 = base[5];
        v.si.sn = v3(q3);
        v.materialIdx = __float_as_int(q3.w);
        v.si.sdpdu = v3(q4);
        v.si.sdpdv = v3(q5);
        v.si.uv = v2(q4.w, q5.w);
    }
    if (LEVEL >= 2) v.si.wo = v3(base[6]);
    if (LEVEL >= 3) v.beta = v3(base[7]);
    return v;
}
HD const float4* vptr(const float4* verts, int P, int depth, uint32_t px) { return verts + ((size_t)depth * P + px) * 8; }
HD float4* vptr(float4* verts, int P, int depth, uint32_t px) { return verts + ((size_t)depth * P + px) * 8; }

// sampled light vertex of an s = 1 strategy (createLightVertex, BDPT.cl:176-202), 32 bytes
HD BVertex unpack_sampled_light(const float4 a, const float4 b) {
    BVertex v;
    v.si.p = v3(a);
    v.pdfFwd = a.w;
    v.si.gn = v3(b);
    const int bits = __float_as_int(b.w);
    v.lightIdx = bits & 0xFFFFFF;
    v.flags = (bits >> 24) & 0xFF;
    v.type = RT_BDPT_LIGHT_VERTEX;
    v.pdfRev = 0.0f;
    v.pdfPos = 0.0f;
    v.materialIdx = RT_INVALID_ID;
    v.si.sn = v.si.gn;
    v.si.sdpdu = v.si.sdpdv = v.si.wo = v.si.dpdu = v.si.dpdv = v.beta = v3(0.0f);
    v.si.uv = v2(0.0f, 0.0f);
    v.si.traceErrorOffset = RT_TRACE_OFFSET;
    v.si.shapeIdx = 0;
    return v;
}
HD int light_vertex_flags(int lightFlags) {
    if ((lightFlags & RT_LIGHT_FLAG_DELTA_DIRECTION) != 0) return VF_DELTA_LIGHT | VF_INFINITE_LIGHT;
    int f = VF_CONNECTIBLE;
    if ((lightFlags & RT_LIGHT_FLAG_DELTA_POSITION) != 0) f |= VF_DELTA_LIGHT;
    return f;
}

// ---------------------------------------------------------------------------------------------
// densities (BDPT.cl:23-154, 215-235)
// ---------------------------------------------------------------------------------------------
HD float computeShadingNormalCorrection(const RTInteraction* inter, V3 wo, V3 wi, int mode) {
    if (mode == TRANSPORT_MODE_IMPORTANCE) {
        float denom = absDot(wo, inter->gn) * absDot(wi, inter->sn);
        if (isNearZero(denom)) return 0.0f;
        return (absDot(wo, inter->sn) * absDot(wi, inter->gn)) / denom;
    }
    return 1.0f;
}

HD float convertVertexDensity(float pdf, const BVertex& thisV, const BVertex& nextV) {
    if (vInfiniteLight(nextV)) return pdf;
    V3 w = nextV.si.p - thisV.si.p;
    float lenSq = dot(w, w);
    if (isNearZero(lenSq)) return 0.0f;
    float invDistSq = 1.0f / lenSq;
    if (vOnSurface(nextV)) pdf *= absDot(nextV.si.gn, w * sqrtf(invDistSq));
    return pdf * invDistSq;
}

HD float evalVertexPdfLight(const Scene& sc, const BVertex& thisV, const BVertex& v) {
    V3 w = v.si.p - thisV.si.p;
    float lenSq = dot(w, w);
    if (isNearZero(lenSq)) return 0.0f;
    float invDistSq = 1.0f / lenSq;
    w *= sqrtf(invDistSq);
    float pdf;
    if (vInfiniteLight(thisV)) {
        float radius = sc.lights[thisV.lightIdx].radius;
        pdf = 1.0f / (PT_PI * radius * radius);
    } else {
        float pdfPos, pdfDir;
        evalLightPdfLe(thisV.lightIdx, sc, w, thisV.si.gn, &pdfPos, &pdfDir);
        pdf = pdfDir * invDistSq;
    }
    if (vOnSurface(v)) pdf *= absDot(v.si.gn, w);
    return pdf;
}

// thisV must carry LEVEL >= 1 fields when it is a surface vertex
BD_EVAL float evalVertexPdf(const Scene& sc, const BVertex& thisV, const BVertex* prevV, const BVertex& nextV) {
    if (thisV.type == RT_BDPT_LIGHT_VERTEX) return evalVertexPdfLight(sc, thisV, nextV);
    V3 wn = nextV.si.p - thisV.si.p;
    float lenSq = dot(wn, wn);
    if (isNearZero(lenSq)) return 0.0f;
    wn /= sqrtf(lenSq);
    float pdf = 0.0f, unused;
    if (thisV.type == RT_BDPT_CAMERA_VERTEX) {
        evalPinholeCameraPdfWe(sc.camera, thisV.si.p, wn, &unused, &pdf);
    } else if (thisV.type == RT_BDPT_SURFACE_VERTEX) {
        V3 wp = v3(0.0f);
        if (prevV) {
            wp = prevV->si.p - thisV.si.p;
            lenSq = dot(wp, wp);
            if (isNearZero(lenSq)) return 0.0f;
            wp /= sqrtf(lenSq);
        }
        UberProps um;
        getUberMaterialProperties(sc, thisV.materialIdx, &thisV.si, &um);
        pdf = evaluateUberBSDF_Pdf(um, &thisV.si, wp, wn, BSDF_ALL);
    }
    return convertVertexDensity(pdf, thisV, nextV);
}

HD float evalVertexPdfLightOrigin(const Scene& sc, const BVertex& thisV, V3 nextVertexPos) {
    V3 w = nextVertexPos - thisV.si.p;
    float lenSq = dot(w, w);
    if (isNearZero(lenSq)) return 0.0f;
    w *= sqrtf(1.0f / lenSq);
    if (vInfiniteLight(thisV)) return 0.0f;
    float pdfPos, pdfDir;
    evalLightPdfLe(thisV.lightIdx, sc, w, thisV.si.gn, &pdfPos, &pdfDir);
    return pdfPos * sc.lights[thisV.lightIdx].choicePdf;
}

// thisV: surface vertex with LEVEL >= 2 fields
BD_EVAL V3 evalVertex_f(const Scene& sc, const BVertex& thisV, V3 nextP, int mode) {
    V3 wi = nextP - thisV.si.p;
    float lenSq = dot(wi, wi);
    if (isNearZero(lenSq)) return v3(0.0f);
    wi /= sqrtf(lenSq);
    if (thisV.type == RT_BDPT_SURFACE_VERTEX) {
        UberProps um;
        getUberMaterialProperties(sc, thisV.materialIdx, &thisV.si, &um);
        V3 f = evaluateUberBSDF(um, &thisV.si, thisV.si.wo, wi, mode);
        return f * computeShadingNormalCorrection(&thisV.si, thisV.si.wo, wi, mode);
    }
    return v3(1.0f, 0.0784f, 0.5765f);   // "shouldn't happen" colour of the reference (BDPT.cl:233-234)
}

// ---------------------------------------------------------------------------------------------
// queue control (one thread between stages; all counts stay on the device)
// ---------------------------------------------------------------------------------------------
// cnt: [0..1] cam queue counts, [2..3] light queue counts, [4] cam work, [5] light work, [6] records, [7] rays, [8] conn work
__global__ void k_bdpt_begin(unsigned int* cnt, unsigned long long* stats, int n) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    cnt[0] = (unsigned)n;
    cnt[2] = (unsigned)n;
    cnt[1] = cnt[3] = cnt[4] = cnt[5] = cnt[6] = cnt[7] = cnt[8] = 0;
    stats[0] += 2ull * (unsigned long long)n;     // first camera + light rays
    stats[3] += (unsigned long long)n;
}
// after an extension step: `nxt` queues become current
__global__ void k_bdpt_swap(unsigned int* cnt, unsigned long long* stats, int nxt, int light_too) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    cnt[1 - nxt] = 0;
    cnt[4] = 0;
    stats[0] += cnt[nxt];
    stats[2] += cnt[nxt];
    if (light_too) {
        cnt[2 + (1 - nxt)] = 0;
        cnt[5] = 0;
        stats[0] += cnt[2 + nxt];
        stats[2] += cnt[2 + nxt];
    }
}
__global__ void k_bdpt_conn_stats(unsigned int* cnt, unsigned long long* stats) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    stats[1] += cnt[7];
}

// ---------------------------------------------------------------------------------------------
// GenerateStartVertices (BDPT.cl:240-312)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_bdpt_start(PtState s, BdptState b, const uint32_t* __restrict__ pixels, int n, int frame) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t px = pixels ? pixels[i] : (uint32_t)i;
    const int P = s.width * s.height;
    const Scene sc = make_scene(s);
    const RTPinholeCamera* cam = sc.camera;
    b.cam_count[px] = 1;
    b.light_count[px] = 1;

    // camera ray and vertex
    const int x = (int)(px % (uint32_t)s.width), y = (int)(px / (uint32_t)s.width);
    const V3 d = primary_ray_dir(cam, x, y);
    const V3 cpos = v3(cam->pos);
    b.cam_ray_o[px] = f4(cpos, RT_MAX_TRACE_DISTANCE);
    b.cam_ray_d[px] = f4(d, 0.0f);
    BVertex cv;
    cv.pdfFwd = cv.pdfRev = cv.pdfPos = 0.0f;
    cv.type = RT_BDPT_CAMERA_VERTEX;
    cv.flags = VF_CONNECTIBLE;
    cv.lightIdx = RT_INVALID_ID;
    cv.materialIdx = RT_INVALID_ID;
    cv.si.p = cpos;
    cv.si.gn = cv.si.sn = cv.si.sdpdu = cv.si.sdpdv = cv.si.wo = v3(0.0f);
    cv.si.uv = v2(0.0f, 0.0f);
    cv.si.traceErrorOffset = 0.0f;
    cv.beta = v3(1.0f);
    store_vertex(vptr(b.cam_vertices, P, 0, px), cv);
    float pdfPos, pdfDir;
    evalPinholeCameraPdfWe(cam, cpos, d, &pdfPos, &pdfDir);
    b.cam_beta[px] = make_float4(1.0f, 1.0f, 1.0f, pdfDir);

    // light ray and vertex
    Sampler sampler = makeSampler(sc.sobolMatrices, (int)px, frame, s.width, s.height, 0);
    const int chosen = (int)(((uint32_t)floorf(getSample1D(&sampler) * sc.numLights)) % (uint32_t)sc.numLights);
    const float lightPdf = sc.lights[chosen].choicePdf;
    const V2 u1 = getSample2D(&sampler);
    const V2 u2 = getSample2D(&sampler);
    V3 ro, rd, ln;
    const V3 Le = sampleLightLe(chosen, sc, u1, u2, &ro, &rd, &ln, &pdfPos, &pdfDir);
    b.light_ray_o[px] = f4(ro, RT_MAX_TRACE_DISTANCE);
    b.light_ray_d[px] = f4(rd, 0.0f);
    BVertex lv;
    lv.pdfFwd = pdfPos * lightPdf;
    lv.pdfRev = 0.0f;
    lv.pdfPos = pdfPos;
    lv.type = RT_BDPT_LIGHT_VERTEX;
    lv.flags = light_vertex_flags(sc.lights[chosen].flags);
    lv.lightIdx = chosen;
    lv.materialIdx = RT_INVALID_ID;
    lv.si.p = ro;
    lv.si.gn = lv.si.sn = ln;
    lv.si.sdpdu = lv.si.sdpdv = lv.si.wo = v3(0.0f);
    lv.si.uv = v2(0.0f, 0.0f);
    lv.si.traceErrorOffset = RT_TRACE_OFFSET;
    lv.beta = Le;
    store_vertex(vptr(b.light_vertices, P, 0, px), lv);
    const V3 lb = Le * absDot(ln, rd) / (lightPdf * pdfPos * pdfDir);
    b.light_beta[px] = f4(lb, pdfDir);

    b.cam_queue[0][i] = px;
    b.light_queue[0][i] = px;
    b.splat[3 * (size_t)px + 0] = 0.0f;
    b.splat[3 * (size_t)px + 1] = 0.0f;
    b.splat[3 * (size_t)px + 2] = 0.0f;
}

// zero the accumulator of pixels outside the rank's partition (t = 1 strategies splat anywhere)
__global__ void k_bdpt_clear_splat(float* splat, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) splat[i] = 0.0f;
}

// ---------------------------------------------------------------------------------------------
// closest-hit trace of one subpath queue
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 8)
k_bdpt_trace(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const uint32_t* __restrict__ queue,
             const unsigned int* __restrict__ count_ptr, unsigned int* work, const float4* __restrict__ ray_o,
             const float4* __restrict__ ray_d, int4* __restrict__ hit) {
    const unsigned count = *count_ptr;
    persistent_trace<false>(
        nodes, tris, count, work,
        [&](unsigned idx, Ray& r) {
            const uint32_t px = queue[idx];
            const float4 o = ray_o[px], d = ray_d[px];
            r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmax = o.w;
            r.dx = d.x; r.dy = d.y; r.dz = d.z;
            r.mask = -1;
            return false;
        },
        [&](unsigned idx, const Ray& r, const TravState& st) {
            const uint32_t px = queue[idx];
            int4 h = make_int4(-1, -1, 0, 0);
            if (st.best_tri >= 0) {
                const float4 v0 = __ldg(tris + 3 * (size_t)st.best_tri + 0);
                const float4 v1 = __ldg(tris + 3 * (size_t)st.best_tri + 1);
                const float4 v2 = __ldg(tris + 3 * (size_t)st.best_tri + 2);
                const float hx = add_(r.ox, mul_(st.closest, r.dx));
                const float hy = add_(r.oy, mul_(st.closest, r.dy));
                const float hz = add_(r.oz, mul_(st.closest, r.dz));
                const float2 uv = barycentrics_ref(hx, hy, hz, v0, v1, v2);
                h = make_int4(__float_as_int(v0.w), __float_as_int(v1.w), __float_as_int(uv.x), __float_as_int(uv.y));
            }
            hit[px] = h;
        });
}

// ---------------------------------------------------------------------------------------------
// GenerateSecondaryVertices (BDPT.cl:317-458): one subpath extension over the live queue
// ---------------------------------------------------------------------------------------------
template <bool IS_CAMERA>
__global__ void __launch_bounds__(128)
k_bdpt_extend(PtState s, BdptState b, int cur, int frame, int curDepth) {
    const int maxDepth = b.max_depth;
    const int P = s.width * s.height;
    const unsigned count = b.cnt[(IS_CAMERA ? 0 : 2) + cur];
    const uint32_t* __restrict__ queue = IS_CAMERA ? b.cam_queue[cur] : b.light_queue[cur];
    uint32_t* __restrict__ next_queue = IS_CAMERA ? b.cam_queue[1 - cur] : b.light_queue[1 - cur];
    float4* __restrict__ verts = IS_CAMERA ? b.cam_vertices : b.light_vertices;
    float4* __restrict__ ray_o = IS_CAMERA ? b.cam_ray_o : b.light_ray_o;
    float4* __restrict__ ray_d = IS_CAMERA ? b.cam_ray_d : b.light_ray_d;
    const int4* __restrict__ hits = IS_CAMERA ? b.cam_hit : b.light_hit;
    float4* __restrict__ betas = IS_CAMERA ? b.cam_beta : b.light_beta;
    int* __restrict__ counts = IS_CAMERA ? b.cam_count : b.light_count;
    const int mode = IS_CAMERA ? TRANSPORT_MODE_RADIANCE : TRANSPORT_MODE_IMPORTANCE;

    for (unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < count; base += gridDim.x * blockDim.x) {
        const unsigned i = base + (threadIdx.x & 31u);
        bool pushNext = false;
        uint32_t px = 0;
        V3 nextO = v3(0.0f), nextD = v3(0.0f), newBeta = v3(0.0f);
        float newPdfFwd = 0.0f;
        if (i < count) {
            px = queue[i];
            const int4 h = hits[px];
            const Scene sc = make_scene(s);
            const int materialIdx = (h.x != -1 && h.y != -1) ? sc.shapes[h.x].materialId : RT_INVALID_ID;
            if (materialIdx != RT_INVALID_ID) {
                counts[px] = curDepth + 1;     // vertexCounts[bufferIdx]++ (every earlier depth hit, or the path died)
                RTInteraction si = computeSurfaceInteraction(sc, h.x, h.y, __int_as_float(h.z), __int_as_float(h.w));
                const V3 rayDir = v3(ray_d[px]);
                si.wo = -rayDir;
                const bool isBackfacing = dot(si.gn, si.wo) < 0.0f;
                si.traceErrorOffset = isBackfacing ? -RT_TRACE_OFFSET : RT_TRACE_OFFSET;
                applyNormalMapping(sc, materialIdx, &si);

                float4* curPtr = vptr(verts, P, curDepth, px);
                float4* prevPtr = vptr(verts, P, curDepth - 1, px);
                BVertex prev = load_vertex<0>(prevPtr);
                const float4 bt = betas[px];
                float pdfFwd = bt.w;

                BVertex v;                               // setSurfaceVertex (BDPT.cl:204-213)
                v.beta = v3(bt);
                v.si = si;
                v.type = RT_BDPT_SURFACE_VERTEX;
                v.materialIdx = materialIdx;
                v.flags = 0;
                v.pdfRev = 0.0f;
                v.pdfPos = 0.0f;
                v.lightIdx = sc.shapes[h.x].lightID;
                v.pdfFwd = convertVertexDensity(pdfFwd, prev, v);
                bool prevDirty = false;
                if (!IS_CAMERA && curDepth == 1 && vInfiniteLight(prev)) {     // infinite-light correction (:374-384)
                    v.pdfFwd = prev.pdfPos;
                    if (vOnSurface(v)) v.pdfFwd *= absDot(rayDir, si.gn);
                    prev.pdfFwd = 0.0f;
                    prevDirty = true;
                }
                if (curDepth == maxDepth + (IS_CAMERA ? 1 : 0)) {              // last vertex (:387-394)
                    if (hasMaterialNonDeltaComponents(sc, materialIdx, &si)) v.flags |= VF_CONNECTIBLE;
                } else {
                    Sampler sampler = makeSampler(sc.sobolMatrices, (int)px, frame, s.width, s.height,
                                                  curDepth + (maxDepth + 1) * (IS_CAMERA ? 1 : 0));
                    const V2 bsdfSample = getSample2D(&sampler);
                    UberProps um;
                    getUberMaterialProperties(sc, materialIdx, &si, &um);
                    V3 wi = v3(0.0f);
                    const V3 wo = si.wo;
                    int sampledType, numNonDelta;
                    const V3 f = sampleUberBSDF(um, &si, bsdfSample, mode, BSDF_ALL, wo, &wi, &pdfFwd, &numNonDelta,
                                                &sampledType);
                    if (numNonDelta > 0) v.flags |= VF_CONNECTIBLE;
                    if (!(isBlack(f) || isNearZero(pdfFwd))) {
                        V3 beta = v.beta * (f * absDot(wi, si.sn) / pdfFwd);
                        float pdfRev;
                        if ((sampledType & BSDF_SPECULAR) != 0) {
                            v.flags |= VF_DELTA;
                            pdfFwd = 0.0f;
                            pdfRev = 0.0f;
                        } else {
                            pdfRev = evaluateUberBSDF_Pdf(um, &si, wi, wo, BSDF_ALL);
                        }
                        float traceErrorOffset = si.traceErrorOffset;
                        if ((sampledType & BSDF_TRANSMISSION) != 0 && dot(si.gn, wi) * signf(traceErrorOffset) < 0.0f)
                            traceErrorOffset *= -1.0f;
                        nextO = si.p + si.gn * traceErrorOffset;
                        nextD = wi;
                        beta *= computeShadingNormalCorrection(&si, wo, wi, mode);
                        prev.pdfRev = convertVertexDensity(pdfRev, v, prev);
                        prevDirty = true;
                        newBeta = beta;
                        newPdfFwd = pdfFwd;
                        pushNext = true;
                    }
                }
                store_vertex(curPtr, v);
                if (prevDirty)
                    prevPtr[0] = make_float4(prev.pdfFwd, prev.pdfRev, __int_as_float(prev.type | (prev.flags << 8)), prev.pdfPos);
            }
        }
        const unsigned ni = warp_push(pushNext, b.cnt + (IS_CAMERA ? 0 : 2) + (1 - cur));
        if (pushNext) {
            ray_o[px] = f4(nextO, RT_MAX_TRACE_DISTANCE);
            ray_d[px] = f4(nextD, 0.0f);
            betas[px] = f4(newBeta, newPdfFwd);
            next_queue[ni] = px;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// PrepareConnections (BDPT.cl:460-646), strategy-parallel
// ---------------------------------------------------------------------------------------------
// Strategy k of the reference's loop order (t = 1..D+2 outer, s = 0..D+1 inner, depth t+s-2 in [0, D], (1,1)
// skipped): t = 1 has s = 2..D+1, t >= 2 has s = 0..D+2-t.
HD void strategy_of(int k, int D, int* t, int* s) {
    if (k < D) {
        *t = 1;
        *s = k + 2;
        return;
    }
    k -= D;
    int tt = 2, row = D + 1;
    while (k >= row) {
        k -= row;
        --row;
        ++tt;
    }
    *t = tt;
    *s = k;
}

// Strategy classes: 0 = s == 0 (a camera vertex on a light), 1 = t == 1 (light tracing), 2 = s == 1 (next-event
// estimation), 3 = full connection.  k_bdpt_prepare and k_bdpt_connect are instantiated per class and launched four
// times over the same items: each instance carries one class's code only (ncu r02: the all-in-one kernels waited for
// instruction fetch 33 % / 58 % of the time), and the records of a class end up next to each other.
HD int strategy_class(int t, int st) { return st == 0 ? 0 : (t == 1 ? 1 : (st == 1 ? 2 : 3)); }

template <int CLS>
__global__ void __launch_bounds__(128)
k_bdpt_prepare(PtState s, BdptState b, const uint32_t* __restrict__ pixels, int n, int frame) {
    const int D = b.max_depth;
    const int P = s.width * s.height;
    const unsigned long long total = (unsigned long long)n * (unsigned)b.max_conn;
    for (unsigned long long base = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; base < total;
         base += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long g = base + (threadIdx.x & 31u);
        bool push = false, needRay = false;
        V3 C = v3(0.0f), ro = v3(0.0f), rd = v3(0.0f);
        float rlen = 0.0f;
        uint32_t px = 0;
        int target = 0, t = 0, st = 0;
        if (g < total) {
            const int i = (int)(g % (unsigned)n);
            const int k = (int)(g / (unsigned)n);
            px = pixels ? pixels[i] : (uint32_t)i;
            strategy_of(k, D, &t, &st);
            target = (int)px;
            const bool valid = strategy_class(t, st) == CLS && t <= b.cam_count[px] && st <= b.light_count[px];
            const Scene sc = make_scene(s);
            if (CLS == 0 && valid) {
                // emitted radiance of a camera vertex that lies on a light (ConnectVertices, BDPT.cl:722-730)
                const BVertex cv = load_vertex<3>(vptr(b.cam_vertices, P, t - 1, px));
                if (vIsLight(cv)) {
                    const V3 Le = evalLightLe(sc.lights + cv.lightIdx, cv.si.gn, cv.si.wo);
                    C = Le * cv.beta;
                }
                push = !isBlack(C);
            } else if (CLS == 1 && valid) {
                // light tracing: project light vertex s-1 to the camera (:517-561)
                const BVertex lv = load_vertex<3>(vptr(b.light_vertices, P, st - 1, px));
                if (vConnectible(lv)) {
                    V3 wi;
                    float pdf;
                    const int x = (int)(px % (uint32_t)s.width), y = (int)(px / (uint32_t)s.width);
                    V2 npos = v2((float)x / s.width, (float)y / s.height);
                    const V3 importance = samplePinholeCameraWi(sc.camera, lv.si.p, &wi, &pdf, &npos);
                    if (pdf > 0.0f && isNotBlack(importance)) {
                        const V3 sampledBeta = importance / pdf;
                        int ix = (int)floorf(npos.x * s.width + 0.5f), iy = (int)floorf(npos.y * s.height + 0.5f);
                        ix = min(max(ix, 0), s.width - 1);
                        iy = min(max(iy, 0), s.height - 1);
                        target = ix + iy * s.width;
                        const V3 cpos = v3(sc.camera->pos);
                        C = lv.beta * sampledBeta * evalVertex_f(sc, lv, cpos, TRANSPORT_MODE_IMPORTANCE);
                        if (vOnSurface(lv)) C *= absDot(wi, lv.si.sn);
                        ro = lv.si.p + lv.si.gn * lv.si.traceErrorOffset;
                        rlen = distance(ro, cpos);
                        rd = (cpos - ro) / rlen;
                        needRay = true;
                        push = isNotBlack(C);
                    }
                }
            } else if (CLS == 2 && valid) {
                // next-event estimation at camera vertex t-1 with a fresh light sample (:563-597)
                const BVertex cv = load_vertex<3>(vptr(b.cam_vertices, P, t - 1, px));
                if (vConnectible(cv)) {
                    Sampler sampler = makeSampler(sc.sobolMatrices, (int)px, frame, s.width, s.height, (D + 1) + (D + 2));
                    for (int j = 1; j <= t - 2; ++j) {      // draws of the connectible camera vertices before this one
                        const int tf = __float_as_int(vptr(b.cam_vertices, P, j, px)[0].z);
                        if (((tf >> 8) & VF_CONNECTIBLE) != 0) skipSamples(&sampler, 3);
                    }
                    const int chosen = min((int)floorf(getSample1D(&sampler) * sc.numLights), sc.numLights - 1);
                    const float lightPdf = sc.lights[chosen].choicePdf;
                    V3 lpos = v3(0.0f), lnrm = v3(0.0f), wi = v3(0.0f);
                    float pdf = 0.0f;
                    ShadowRay sr;
                    const V2 u = getSample2D(&sampler);
                    const V3 Li = sampleLightLi(chosen, sc, &cv.si, u, &lpos, &lnrm, &wi, &pdf, &sr);
                    if (isNotNearZero(pdf) && isNotBlack(Li)) {
                        float4* slot = b.sampled_light + ((size_t)(t - 2) * P + px) * 2;
                        // the reference evaluates the light-origin density on the slot's PREVIOUS contents
                        // (read before write, BDPT.cl:585-586; zero-filled on the first frame)
                        const BVertex stale = unpack_sampled_light(slot[0], slot[1]);
                        const float pdfFwd = evalVertexPdfLightOrigin(sc, stale, cv.si.p);
                        const int vflags = light_vertex_flags(sc.lights[chosen].flags);
                        slot[0] = f4(lpos, pdfFwd);
                        slot[1] = f4(lnrm, __int_as_float((chosen & 0xFFFFFF) | (vflags << 24)));
                        const V3 sampledBeta = Li / (lightPdf * pdf);
                        UberProps um;
                        getUberMaterialProperties(sc, cv.materialIdx, &cv.si, &um);
                        const V3 f = evaluateUberBSDF(um, &cv.si, cv.si.wo, wi, TRANSPORT_MODE_RADIANCE);
                        C = cv.beta * sampledBeta * f;
                        if (vOnSurface(cv)) C *= absDot(wi, cv.si.sn);
                        ro = sr.o;
                        rd = sr.d;
                        rlen = sr.tmax;
                        needRay = true;
                        push = sr.written && !isBlack(C);
                    }
                }
            } else if (CLS == 3 && valid) {
                // full connection of camera vertex t-1 and light vertex s-1 (:599-632)
                const BVertex cv = load_vertex<3>(vptr(b.cam_vertices, P, t - 1, px));
                const BVertex lv = load_vertex<3>(vptr(b.light_vertices, P, st - 1, px));
                if (vConnectible(cv) && vConnectible(lv)) {
                    const V3 lvf = evalVertex_f(sc, lv, cv.si.p, TRANSPORT_MODE_IMPORTANCE);
                    const V3 cvf = evalVertex_f(sc, cv, lv.si.p, TRANSPORT_MODE_RADIANCE);
                    const V3 lp = lv.si.p + lv.si.gn * lv.si.traceErrorOffset;
                    const V3 cp = cv.si.p + cv.si.gn * cv.si.traceErrorOffset;
                    V3 w = cp - lp;
                    const float sqDist = dot(w, w);
                    const float dist = sqrtf(sqDist);
                    w /= dist;
                    if (isNotNearZero(sqDist)) {
                        const float gterm = absDot(cv.si.sn, w) * absDot(lv.si.sn, w) / sqDist;
                        C = lv.beta * cv.beta * lvf * cvf * gterm;
                    }
                    ro = lp;
                    rd = w;
                    rlen = dist;
                    needRay = true;
                    push = isNotBlack(C);
                }
            }
        }
        const unsigned ri = warp_push(push, b.cnt + 6);
        if (push) {
            b.conn_o[ri] = f4(ro, rlen);
            b.conn_d[ri] = f4(rd, __int_as_float(t | (st << 8)));
            b.conn_C[ri] = f4(C, __int_as_float(target));
            b.conn_src[ri] = (int)px;
            b.conn_vis[ri] = needRay ? 0 : 1;
        }
        const bool pushRay = push && needRay;
        const unsigned qi = warp_push(pushRay, b.cnt + 7);
        if (pushRay) b.conn_rays[qi] = ri;
    }
}

// any-hit trace of the connection rays; writes 1 for unoccluded
__global__ void __launch_bounds__(128, 8)
k_bdpt_trace_conn(PtState s, BdptState b) {
    const unsigned count = b.cnt[7];
    persistent_trace<true>(
        s.nodes, s.tris, count, b.cnt + 8,
        [&](unsigned idx, Ray& r) {
            const uint32_t rec = b.conn_rays[idx];
            const float4 o = b.conn_o[rec], d = b.conn_d[rec];
            r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmax = o.w;
            r.dx = d.x; r.dy = d.y; r.dz = d.z;
            r.mask = -1;
            return true;
        },
        [&](unsigned idx, const Ray&, const TravState& st) { b.conn_vis[b.conn_rays[idx]] = st.best_tri >= 0 ? 0 : 1; });
}

// ---------------------------------------------------------------------------------------------
// ConnectVertices (BDPT.cl:671-913): MIS weight + accumulation, one thread per surviving record
// ---------------------------------------------------------------------------------------------
HD float remap0(float f) { return f == 0.0f ? 1.0f : f; }

template <int CLS>
HD float mis_weight(const Scene& sc, const BdptState& b, int P, uint32_t px, int t, int s) {
    // what the class says about (t, s), for the compiler: the branches below that cannot be taken disappear
    if (CLS == 0) s = 0;
    if (CLS == 1) t = 1;
    if (CLS == 2) s = 1;
    __builtin_assume(t >= 1 && s >= 0);
    if (CLS != 1) __builtin_assume(t >= 2);
    if (CLS == 1 || CLS == 3) __builtin_assume(s >= 2);
    if (s + t == 2) return 1.0f;
    BVertex pt, qs, ptPrev, qsPrev;
    if (t == 1) {      // the sampled camera vertex (createCameraVertex at the camera position, BDPT.cl:159-174)
        pt = load_vertex<0>(vptr(b.cam_vertices, P, 0, px));
        pt.si.p = v3(sc.camera->pos);
        pt.type = RT_BDPT_CAMERA_VERTEX;
        pt.flags = VF_CONNECTIBLE;
        pt.lightIdx = RT_INVALID_ID;
        pt.pdfFwd = pt.pdfRev = 0.0f;
        pt.si.gn = v3(0.0f);
    } else {
        pt = load_vertex<1>(vptr(b.cam_vertices, P, t - 1, px));
    }
    if (s == 1) {
        const float4* slot = b.sampled_light + ((size_t)(t - 2) * P + px) * 2;
        qs = unpack_sampled_light(slot[0], slot[1]);
    } else if (s > 1) {
        qs = load_vertex<1>(vptr(b.light_vertices, P, s - 1, px));
    }
    if (t > 1) ptPrev = load_vertex<0>(vptr(b.cam_vertices, P, t - 2, px));
    if (s > 1) qsPrev = load_vertex<0>(vptr(b.light_vertices, P, s - 2, px));

    float ptRev, ptPrevRev = 0.0f, qsRev = 0.0f, qsPrevRev = 0.0f;
    if (s > 0) ptRev = evalVertexPdf(sc, qs, s > 1 ? &qsPrev : nullptr, pt);
    else ptRev = evalVertexPdfLightOrigin(sc, pt, ptPrev.si.p);
    if (t > 1) ptPrevRev = s > 0 ? evalVertexPdf(sc, pt, &qs, ptPrev) : evalVertexPdfLight(sc, pt, ptPrev);
    if (s > 0) qsRev = evalVertexPdf(sc, pt, t > 1 ? &ptPrev : nullptr, qs);
    if (s > 1) qsPrevRev = evalVertexPdf(sc, qs, &pt, qsPrev);

    float sumRi = 0.0f;
    float ri = 1.0f;
    // camera subpath: vertices t-1 .. 1; the DELTA flag of pt is cleared for the connection
    bool deltaCur = false;
    for (int i = t - 1; i > 0; --i) {
        float rev, fwd;
        if (i == t - 1) {
            rev = ptRev;
            fwd = pt.pdfFwd;
        } else {
            const float4 q0 = vptr(b.cam_vertices, P, i, px)[0];
            rev = (i == t - 2) ? ptPrevRev : q0.y;
            fwd = q0.x;
        }
        const int tfPrev = __float_as_int(vptr(b.cam_vertices, P, i - 1, px)[0].z);
        const bool deltaPrev = ((tfPrev >> 8) & VF_DELTA) != 0;
        ri *= remap0(rev) / remap0(fwd);
        if (!deltaCur && !deltaPrev) sumRi += ri;
        deltaCur = deltaPrev;
    }
    // light subpath: vertices s-1 .. 0
    ri = 1.0f;
    deltaCur = false;
    for (int i = s - 1; i >= 0; --i) {
        float rev, fwd;
        if (i == s - 1) {
            rev = qsRev;
            fwd = qs.pdfFwd;
        } else {
            const float4 q0 = vptr(b.light_vertices, P, i, px)[0];
            rev = (i == s - 2) ? qsPrevRev : q0.y;
            fwd = q0.x;
        }
        bool deltaPrev;
        if (i > 0) {
            const int tfPrev = __float_as_int(vptr(b.light_vertices, P, i - 1, px)[0].z);
            deltaPrev = ((tfPrev >> 8) & VF_DELTA) != 0;
        } else {      // isVertexDeltaLight of light vertex 0 -- which is the sampled vertex when s == 1
            const int f0 = (s == 1) ? qs.flags : (__float_as_int(vptr(b.light_vertices, P, 0, px)[0].z) >> 8);
            deltaPrev = (f0 & VF_DELTA_LIGHT) != 0;
        }
        ri *= remap0(rev) / remap0(fwd);
        if (!deltaCur && !deltaPrev) sumRi += ri;
        deltaCur = (i > 0) ? deltaPrev : false;
    }
    return 1.0f / (1.0f + sumRi);
}

template <int CLS>
__global__ void __launch_bounds__(128)
k_bdpt_connect(PtState s, BdptState b) {
    const unsigned count = b.cnt[6];
    const int P = s.width * s.height;
    const Scene sc = make_scene(s);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int ts = __float_as_int(b.conn_d[i].w);
        const int t = ts & 0xFF, st = ts >> 8;
        if (strategy_class(t, st) != CLS) continue;
        if (b.conn_vis[i] == 0) {
            b.conn_w[i] = 0.0f;
            continue;
        }
        const float4 Cq = b.conn_C[i];
        const uint32_t px = (uint32_t)b.conn_src[i];
        const int target = __float_as_int(Cq.w);
        const V3 C = v3(Cq);
        const float w = isBlack(C) ? 0.0f : mis_weight<CLS>(sc, b, P, px, t, st);
        b.conn_w[i] = w;
        atomicAdd(b.splat + 3 * (size_t)target + 0, C.x * w);
        atomicAdd(b.splat + 3 * (size_t)target + 1, C.y * w);
        atomicAdd(b.splat + 3 * (size_t)target + 2, C.z * w);
    }
}

// CopyBuffer (BDPT.cl:916-932)
__global__ void k_bdpt_finish(const float* __restrict__ splat, float4* __restrict__ radiance, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) radiance[i] = make_float4(splat[3 * (size_t)i], splat[3 * (size_t)i + 1], splat[3 * (size_t)i + 2], 0.0f);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
const std::vector<const char*>& bdpt_buffer_names() {
    static const std::vector<const char*> names = {
        "bdpt_cam_vertices", "bdpt_light_vertices", "bdpt_sampled_light", "bdpt_cam_ray_o", "bdpt_cam_ray_d",
        "bdpt_light_ray_o", "bdpt_light_ray_d", "bdpt_cam_hit", "bdpt_light_hit", "bdpt_cam_beta", "bdpt_light_beta",
        "bdpt_cam_count", "bdpt_light_count", "bdpt_cam_queue0", "bdpt_cam_queue1", "bdpt_light_queue0",
        "bdpt_light_queue1", "bdpt_conn_o", "bdpt_conn_d", "bdpt_conn_C", "bdpt_conn_src", "bdpt_conn_vis",
        "bdpt_conn_rays", "bdpt_conn_w", "bdpt_splat", "bdpt_cnt"};
    return names;
}

std::vector<BdptBufferSpec> bdpt_buffer_specs(BdptState& b, int D) {
    const size_t maxConn = (size_t)(D + 2) * (D + 3) / 2 - 2;      // RTBDPTPass::getMaxPossibleConnectionsCount
    b.max_depth = D;
    b.max_conn = (int)maxConn;
    return {
        {"bdpt_cam_vertices", (size_t)(D + 2) * 128, (void**)&b.cam_vertices},
        {"bdpt_light_vertices", (size_t)(D + 1) * 128, (void**)&b.light_vertices},
        {"bdpt_sampled_light", (size_t)D * 32, (void**)&b.sampled_light},
        {"bdpt_cam_ray_o", 16, (void**)&b.cam_ray_o},
        {"bdpt_cam_ray_d", 16, (void**)&b.cam_ray_d},
        {"bdpt_light_ray_o", 16, (void**)&b.light_ray_o},
        {"bdpt_light_ray_d", 16, (void**)&b.light_ray_d},
        {"bdpt_cam_hit", 16, (void**)&b.cam_hit},
        {"bdpt_light_hit", 16, (void**)&b.light_hit},
        {"bdpt_cam_beta", 16, (void**)&b.cam_beta},
        {"bdpt_light_beta", 16, (void**)&b.light_beta},
        {"bdpt_cam_count", 4, (void**)&b.cam_count},
        {"bdpt_light_count", 4, (void**)&b.light_count},
        {"bdpt_cam_queue0", 4, (void**)&b.cam_queue[0]},
        {"bdpt_cam_queue1", 4, (void**)&b.cam_queue[1]},
        {"bdpt_light_queue0", 4, (void**)&b.light_queue[0]},
        {"bdpt_light_queue1", 4, (void**)&b.light_queue[1]},
        {"bdpt_conn_o", maxConn * 16, (void**)&b.conn_o},
        {"bdpt_conn_d", maxConn * 16, (void**)&b.conn_d},
        {"bdpt_conn_C", maxConn * 16, (void**)&b.conn_C},
        {"bdpt_conn_src", maxConn * 4, (void**)&b.conn_src},
        {"bdpt_conn_vis", maxConn * 4, (void**)&b.conn_vis},
        {"bdpt_conn_rays", maxConn * 4, (void**)&b.conn_rays},
        {"bdpt_conn_w", maxConn * 4, (void**)&b.conn_w},
        {"bdpt_splat", 12, (void**)&b.splat},
        {"bdpt_cnt", 0, (void**)&b.cnt},      // 0 = fixed 64-byte block, not per pixel
    };
}

int launch_bdpt_sample(const PtState& s, const BdptState& b, const uint32_t* pixels, int n_pixels, int frame,
                       int max_depth, cudaStream_t stream, cudaError_t* err) {
    int launches = 0;
    *err = cudaSuccess;
    if (n_pixels <= 0 || s.numLights <= 0) return 0;      // RTBDPTPass::update returns early without lights (:69)
    const int P = s.width * s.height;
    const int sms = sm_count();
    const int trace_blocks = min(sms * 8, (n_pixels + 127) / 128);
    const int work_blocks = min(sms * 16, (n_pixels + 127) / 128);
    const int D = max_depth;

    if (pixels) {      // a partition owns a subset of pixels but may splat anywhere
        const size_t n3 = 3 * (size_t)P;
        k_bdpt_clear_splat<<<(unsigned)((n3 + 255) / 256), 256, 0, stream>>>(b.splat, n3);
        ++launches;
    }
    k_bdpt_begin<<<1, 32, 0, stream>>>(b.cnt, s.stats, n_pixels);
    k_bdpt_start<<<(n_pixels + 127) / 128, 128, 0, stream>>>(s, b, pixels, n_pixels, frame);
    launches += 2;
    int cur = 0;
    k_bdpt_trace<<<trace_blocks, 128, 0, stream>>>(s.nodes, s.tris, b.cam_queue[cur], b.cnt + cur, b.cnt + 4, b.cam_ray_o,
                                                    b.cam_ray_d, b.cam_hit);
    k_bdpt_trace<<<trace_blocks, 128, 0, stream>>>(s.nodes, s.tris, b.light_queue[cur], b.cnt + 2 + cur, b.cnt + 5,
                                                    b.light_ray_o, b.light_ray_d, b.light_hit);
    launches += 2;
    for (int depth = 1; depth <= D + 1; ++depth) {      // generateSecondaryVertices (RTBDPTPass.cpp:243-300)
        const bool light = depth <= D;
        k_bdpt_extend<true><<<work_blocks, 128, 0, stream>>>(s, b, cur, frame, depth);
        ++launches;
        if (light) {
            k_bdpt_extend<false><<<work_blocks, 128, 0, stream>>>(s, b, cur, frame, depth);
            ++launches;
        }
        if (depth == D + 1) break;      // the last camera vertex spawns no ray
        cur = 1 - cur;
        k_bdpt_swap<<<1, 32, 0, stream>>>(b.cnt, s.stats, cur, light ? 1 : 0);
        k_bdpt_trace<<<trace_blocks, 128, 0, stream>>>(s.nodes, s.tris, b.cam_queue[cur], b.cnt + cur, b.cnt + 4,
                                                        b.cam_ray_o, b.cam_ray_d, b.cam_hit);
        launches += 2;
        if (light && depth < D) {       // the last light vertex (depth D) spawns no ray either
            k_bdpt_trace<<<trace_blocks, 128, 0, stream>>>(s.nodes, s.tris, b.light_queue[cur], b.cnt + 2 + cur,
                                                            b.cnt + 5, b.light_ray_o, b.light_ray_d, b.light_hit);
            ++launches;
        }
    }
    const unsigned long long total = (unsigned long long)n_pixels * (unsigned)b.max_conn;
    const int prep_blocks = (int)std::min<unsigned long long>((total + 127) / 128, (unsigned long long)sms * 32);
    k_bdpt_prepare<0><<<prep_blocks, 128, 0, stream>>>(s, b, pixels, n_pixels, frame);
    k_bdpt_prepare<1><<<prep_blocks, 128, 0, stream>>>(s, b, pixels, n_pixels, frame);
    k_bdpt_prepare<2><<<prep_blocks, 128, 0, stream>>>(s, b, pixels, n_pixels, frame);
    k_bdpt_prepare<3><<<prep_blocks, 128, 0, stream>>>(s, b, pixels, n_pixels, frame);
    k_bdpt_conn_stats<<<1, 32, 0, stream>>>(b.cnt, s.stats);
    k_bdpt_trace_conn<<<sms * 8, 128, 0, stream>>>(s, b);
    k_bdpt_connect<0><<<sms * 16, 128, 0, stream>>>(s, b);
    k_bdpt_connect<1><<<sms * 16, 128, 0, stream>>>(s, b);
    k_bdpt_connect<2><<<sms * 16, 128, 0, stream>>>(s, b);
    k_bdpt_connect<3><<<sms * 16, 128, 0, stream>>>(s, b);
    k_bdpt_finish<<<(P + 255) / 256, 256, 0, stream>>>(b.splat, s.radiance, P);
    launches += 11;
    *err = cudaGetLastError();
    return launches;
}

}  // namespace mcrt
