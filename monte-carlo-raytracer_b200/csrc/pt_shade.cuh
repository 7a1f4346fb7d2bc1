// pt_shade.cuh -- one path vertex of the regular path tracer: surface interaction, emission,
// next-event estimation and BSDF sampling.  Restates the body of the reference's `PathTracing`
// kernel (assets/kernels/PathTracing.cl:52-184) as a device function on registers, so that both the
// queue-based wavefront kernel and the reference-layout stage kernel share it.
//
// Reference semantics kept: emission only on the first hit or after
// a specular bounce, and it terminates the path; one light sample per vertex; sampler re-seeded
// per bounce with draws in the fixed order 1D light pick, 2D light sample, 2D BSDF sample; no
// Russian roulette, no MIS, no environment term.
//
// Opt-in estimator switches (EST = true, `estimator` flags; BASELINE north_star asks for them, the reference has
// neither, so they are off by default and excluded from parity runs -- SURVEY App. B.4):
//   PT_EST_NEE_MIS           power-heuristic MIS between light sampling and BSDF sampling for triangle-mesh
//                            area lights (the only lights a BSDF-sampled ray can hit): the NEE term is weighted by
//                            p_l^2 / (p_l^2 + p_b^2), and a non-specular bounce that lands on an emitter adds
//                            throughput * Le * p_b^2 / (p_b^2 + p_l^2) where the reference adds nothing.  Same
//                            expectation as the reference's estimator when the light pdf is exact (E10: meshes whose
//                            triangles have equal areas), lower variance on glossy surfaces.
//   PT_EST_RUSSIAN_ROULETTE  from the third vertex on, a path survives with probability
//                            min(0.95, max throughput component) and is re-weighted; sixth draw of the bounce's
//                            sampler, so the five draws of the reference's estimator are unchanged.
#pragma once
#include "pt_material.cuh"

namespace mcrt {

enum { PT_EST_RUSSIAN_ROULETTE = 1, PT_EST_NEE_MIS = 2 };   // = PTB_EST_* (include/ptb.h)

struct ShadeIn {
    int pixel;            // GLOBAL buffer index (seeds the sampler)
    int shapeId, primId;
    float bu, bv;         // hit barycentrics
    V3 rayDir;
    V3 throughput;        // ignored on bounce 0 (set to 1)
    int prevBsdfFlags;
    // EST only: density with which the previous vertex's lobe mixture samples rayDir (0 after a delta lobe), the
    // path weight for light found along rayDir (throughput * f * cos / that density), and the ray origin
    float prevPdf;
    V3 prevOrigin, prevDirectWeight;
};

struct ShadeOut {
    V3 emitted;           // throughput * Le, added unconditionally (ignoreOcclusion = 1)
    bool hasEmission;
    V3 direct;            // throughput * L, to be multiplied by the shadow ray's visibility
    ShadowRay shadow;     // .written = the reference's setRay() on the shadow ray ran
    bool continues;       // a next ray was spawned
    float nextPdf;        // EST only: mixture density of nextD (0 for a delta lobe), and the weight that goes with it
    V3 nextDirectWeight;
    V3 nextO, nextD;
    V3 throughput;
    int bsdfFlags;
};

// SYNC = true: the CTA's warps are kept in the same phase of this (very long) function with
// __syncthreads() so that they share the instruction stream in the SM's instruction cache; every
// thread of the CTA must then call it, `active` telling whether it has a vertex to shade.
template <bool SYNC>
HD void phase_barrier() {
    if (SYNC) __syncthreads();
}

HD float power_heuristic(float a, float b) {
    const float a2 = a * a, b2 = b * b;
    return a2 + b2 > 0.0f ? a2 / (a2 + b2) : 0.0f;
}

template <bool SYNC, bool EST = false>
HD void shade_vertex(const Scene& sc, const ShadeIn& in, bool active, int frame, int width, int height, int bounce,
                     int maxDepth, ShadeOut* out, int estimator = 0) {
    out->hasEmission = false;
    out->nextPdf = 0.0f;
    out->emitted = v3(0.0f);
    out->direct = v3(0.0f);
    out->shadow.written = false;
    out->continues = false;
    out->bsdfFlags = in.prevBsdfFlags;
    V3 throughput = bounce == 0 ? v3(1.0f) : in.throughput;
    out->throughput = throughput;

    RTInteraction si;
    int materialId = RT_INVALID_ID, lightID = RT_INVALID_ID;
    if (active) {
        si = computeSurfaceInteraction(sc, in.shapeId, in.primId, in.bu, in.bv);
        si.wo = -in.rayDir;
        const bool isBackfacing = dot(si.gn, si.wo) < 0.0f;
        si.traceErrorOffset = isBackfacing ? -RT_TRACE_OFFSET : RT_TRACE_OFFSET;
        const RTShape* shape = sc.shapes + in.shapeId;
        materialId = shape->materialId;
        lightID = shape->lightID;
    }
    phase_barrier<SYNC>();
    if (active) applyNormalMapping(sc, materialId, &si);

    if (active) {
        const bool isEmitter = lightID != RT_INVALID_ID;
        const bool sampledSpecular = (BSDF_SPECULAR & in.prevBsdfFlags) == BSDF_SPECULAR;
        if (isEmitter && (bounce == 0 || sampledSpecular)) {
            V3 Le = evalLightLe(sc.lights + lightID, si.gn, si.wo);
            out->emitted = throughput * Le;
            out->hasEmission = true;
            active = false;      // the path ends here (PathTracing.cl:91-101)
        } else if (EST && isEmitter && (estimator & PT_EST_NEE_MIS) && in.prevPdf > 0.0f &&
                   sc.lights[lightID].type == RT_TRIANGLE_MESH_AREA_LIGHT) {
            // BSDF-sampled hit of an area light: the share of direct light that light sampling at the previous
            // vertex did not take.  p_l as sampleLightLi computes it for this point (lights.cl:104-131).
            const RTLight* light = sc.lights + lightID;
            const float c = absDot(si.gn, si.wo);
            const float pl = isNearZero(c) ? 0.0f
                                           : light->choicePdf * (1.0f / light->area) * distanceSquared(si.p, in.prevOrigin) / c;
            out->emitted = in.prevDirectWeight * evalLightLe(light, si.gn, si.wo) * power_heuristic(in.prevPdf, pl);
        }
    }
    phase_barrier<SYNC>();

    Sampler sampler = makeSampler(sc.sobolMatrices, in.pixel, frame, width, height, bounce);
    UberProps um;
    const bool hasMaterial = materialId != RT_INVALID_ID;
    if (active && hasMaterial) getUberMaterialProperties(sc, materialId, &si, &um);
    phase_barrier<SYNC>();

    // direct lighting: one light, one sample (PathTracing.cl:106-134)
    float lightPdf = 0.0f;
    V3 wiL = v3(0.0f), Li = v3(0.0f);
    bool lightIsMesh = false;
    if (active) {
        uint32_t lightIdx = ((uint32_t)floorf(getSample1D(&sampler) * sc.numLights));
        lightIdx %= (uint32_t)sc.numLights;
        if (EST) lightIsMesh = sc.lights[lightIdx].type == RT_TRIANGLE_MESH_AREA_LIGHT;
        V2 u = getSample2D(&sampler);
        V3 unusedN, unusedP;
        Li = sampleLightLi((int)lightIdx, sc, &si, u, &unusedP, &unusedN, &wiL, &lightPdf, &out->shadow);
        lightPdf *= sc.lights[lightIdx].choicePdf;
    }
    phase_barrier<SYNC>();
    if (active) {
        V3 L = v3(0.0f);
        if (hasMaterial) {
            V3 bsdf = evaluateUberBSDF(um, &si, si.wo, wiL, TRANSPORT_MODE_RADIANCE);
            bsdf *= absDot(wiL, si.sn);
            if (!isNearZero(lightPdf)) L = Li * bsdf / lightPdf;
            // (the last vertex of a path spawns no ray, so light sampling is its only strategy: weight 1)
            if (EST && (estimator & PT_EST_NEE_MIS) && !isNearZero(lightPdf) && lightIsMesh && bounce + 1 < maxDepth)
                L *= power_heuristic(lightPdf, evaluateUberBSDF_Pdf(um, &si, si.wo, wiL, BSDF_ALL));
        }
        out->direct = throughput * L;
    }
    phase_barrier<SYNC>();

    if (active && bounce + 1 < maxDepth) {   // extend the path (PathTracing.cl:137-172)
        V2 bsdfSample = getSample2D(&sampler);
        if (hasMaterial) {
            float pdf = 0.0f;
            int sampledType, unused;
            V3 wi = v3(0.0f);
            V3 f = sampleUberBSDF(um, &si, bsdfSample, TRANSPORT_MODE_RADIANCE, BSDF_ALL, si.wo, &wi, &pdf, &unused,
                                  &sampledType);
            out->bsdfFlags = sampledType;
            if (!(isNearZero(pdf) || isBlack(f))) {
                f /= pdf;
                V3 tp = f * absDot(wi, si.sn);
                out->throughput = throughput * tp;
                float traceErrorOffset = si.traceErrorOffset;
                if ((sampledType & BSDF_TRANSMISSION) != 0 && dot(si.gn, wi) * signf(traceErrorOffset) < 0.0f)
                    traceErrorOffset *= -1.0f;
                out->nextO = si.p + si.gn * traceErrorOffset;
                out->nextD = wi;
                out->continues = true;
                if (EST) {
                    // The reference's sampler divides by a branch-dependent "pdf" (the lobe walk of bxdfs.cl:892-1053
                    // skips the Lambertian term when the microfacet lobe was chosen, E7), which is fine for its own
                    // estimator but is not the density of wi.  The MIS share of direct light found along wi uses
                    // the true mixture density and the full BSDF value instead.
                    out->nextPdf = 0.0f;
                    out->nextDirectWeight = v3(0.0f);
                    if ((estimator & PT_EST_NEE_MIS) && !(sampledType & BSDF_SPECULAR)) {
                        const float pmix = evaluateUberBSDF_Pdf(um, &si, si.wo, wi, BSDF_ALL);
                        if (!isNearZero(pmix)) {
                            out->nextPdf = pmix;
                            out->nextDirectWeight = throughput *
                                (evaluateUberBSDF(um, &si, si.wo, wi, TRANSPORT_MODE_RADIANCE) * absDot(wi, si.sn) / pmix);
                        }
                    }
                    if ((estimator & PT_EST_RUSSIAN_ROULETTE) && bounce >= 2) {
                        const V3 t = out->throughput;
                        const float survive = fminf(0.95f, fmaxf(t.x, fmaxf(t.y, t.z)));
                        if (!(getSample1D(&sampler) < survive)) {
                            out->continues = false;
                        } else {
                            out->throughput = t / survive;
                            out->nextDirectWeight = out->nextDirectWeight / survive;
                        }
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The same vertex in TWO stages (ptb_set_shade_split; the reference's estimator only).  shade_vertex above needs
// ~119 registers and 13,000 SASS instructions, so the shade kernel runs 16 warps per SM and waits: on the dependent
// gathers of the surface half (hit -> shape -> indices -> vertices -> uv -> texture descriptors -> texels) and on
// instruction fetch.  Split, the gather-heavy half (surface interaction, normal map, emission, material textures)
// runs at a higher occupancy and drops the rays that missed; the arithmetic half (light sampling, BSDF evaluation,
// lobe sampling) reads a dense 112-byte record per live vertex and keeps no lane idle.  Every expression is the one
// shade_vertex evaluates, on the same values, so per-sample results are bit-identical (tests: the coherence /
// split equality test and the full-size parity tests run either way).
//
// Which material properties travel in the record: those a texture can change here -- Kd, Ks and the opacity
// (diffuse, glossy and opacity textures, the diffuse texture's alpha).  Kr, Kt, eta and the roughness are
// re-read from the material table by the second stage, which is exact when the material binds no specular-
// reflection / transmission / ior / roughness texture; the host uses the split only for scenes where none does
// (split_supported in ptb_upload_scene).
struct VertexRecord {
    V3 p, gn, sn, sdpdu, sdpdv, Kd, Ks, opacity;
    uint32_t px;
    int flags;      // bit 0: traceErrorOffset is negative (back face), bit 1: the shape has a material, bits 8..: material id
    int key;        // what the lanes of the second stage would diverge on (scatter_key); a scheduling hint only
};

// Divergence key of the second stage, 7 bits: the TYPE of the light the vertex will sample (2 bits), the lobe its
// BSDF sample will pick (3 bits: the walk of sampleUberBSDF over the material's components with the sampler's
// fourth draw of the bounce) and whether the Lambertian / microfacet-reflection components exist (2 bits: they are
// evaluated for the other lobe's direction).  The sampler is a pure function of (pixel, frame, bounce)
// (samplers.cl:75-84), so all of it is known here.  The second stage may order its records by this key; no result
// depends on it.
HD int scatter_key(const Scene& sc, const UberProps& um, bool hasMaterial, uint32_t px, int frame, int width, int height,
                   int bounce) {
    Sampler sp = makeSampler(sc.sobolMatrices, (int)px, frame, width, height, bounce);
    uint32_t lightIdx = ((uint32_t)floorf(getSample1D(&sp) * sc.numLights));
    lightIdx %= (uint32_t)sc.numLights;
    const int ltype = sc.lights[lightIdx].type & 3;
    if (!hasMaterial) return ltype;
    skipSamples(&sp, 2);
    const float ux = getSample1D(&sp);
    const V3 t = v3(1.0f) - um.opacity;
    const V3 kd = um.Kd * um.opacity, ks = um.Ks * um.opacity, kt = um.Kt * um.opacity, kr = um.Kr * um.opacity;
    const bool perfectT = um.KtW < 0.5f;
    // component order of the walk: T, perfect-specular Kt, Kr, microfacet Kt, Kd, Ks
    const bool has[6] = {isNotBlack(t), perfectT && isNotBlack(kt), isNotBlack(kr), isNotBlack(kt) && !perfectT,
                         isNotBlack(kd), isNotBlack(ks)};
    int n = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) n += has[i] ? 1 : 0;
    int lobe = 7;
    if (n > 0) {
        int chosen = min((int)floorf(ux * n), n - 1);
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (has[i] && chosen-- == 0) lobe = i;
    }
    return ltype | (lobe << 2) | (has[4] ? 32 : 0) | (has[5] ? 64 : 0);
}
enum { SURF_DEAD = 0, SURF_EMITTED = 1, SURF_RECORD = 2 };

// first stage.  SURF_EMITTED: the path ends on an emitter, *Le is its radiance towards the ray origin.
HD int shade_surface(const Scene& sc, int shapeId, int primId, float bu, float bv, V3 rayDir, int prevBsdfFlags,
                     int bounce, uint32_t px, int frame, int width, int height, V3* Le, VertexRecord* rec) {
    RTInteraction si = computeSurfaceInteraction(sc, shapeId, primId, bu, bv);
    si.wo = -rayDir;
    const bool isBackfacing = dot(si.gn, si.wo) < 0.0f;
    si.traceErrorOffset = isBackfacing ? -RT_TRACE_OFFSET : RT_TRACE_OFFSET;
    const RTShape* shape = sc.shapes + shapeId;
    const int materialId = shape->materialId;
    const int lightID = shape->lightID;
    applyNormalMapping(sc, materialId, &si);
    const bool isEmitter = lightID != RT_INVALID_ID;
    const bool sampledSpecular = (BSDF_SPECULAR & prevBsdfFlags) == BSDF_SPECULAR;
    if (isEmitter && (bounce == 0 || sampledSpecular)) {
        *Le = evalLightLe(sc.lights + lightID, si.gn, si.wo);
        return SURF_EMITTED;
    }
    const bool hasMaterial = materialId != RT_INVALID_ID;
    UberProps um;
    um.Kd = um.Ks = um.opacity = um.Kr = um.Kt = v3(0.0f);
    um.KtW = 0.0f;
    if (hasMaterial) getUberMaterialProperties(sc, materialId, &si, &um);
    rec->p = si.p;
    rec->gn = si.gn;
    rec->sn = si.sn;
    rec->sdpdu = si.sdpdu;
    rec->sdpdv = si.sdpdv;
    rec->Kd = um.Kd;
    rec->Ks = um.Ks;
    rec->opacity = um.opacity;
    rec->flags = (isBackfacing ? 1 : 0) | (hasMaterial ? 2 : 0) | (hasMaterial ? (materialId << 8) : 0);
    rec->px = px;
    rec->key = scatter_key(sc, um, hasMaterial, px, frame, width, height, bounce);
    return SURF_RECORD;
}

// the properties getUberMaterialProperties takes from the material table alone (no texture bound)
HD void untexturedUberProperties(const Scene& sc, int materialIdx, UberProps* p) {
    const RTMaterial* m = sc.materials + materialIdx;
    p->Kr = v3(1.0f) * v3(m->kr);
    p->Kt = v3(1.0f) * v3(m->kt);
    p->KtW = m->kt.w;
    p->eta = m->eta;
    if (sc.materialAlpha) {
        const float2 a = sc.staged ? sc.materialAlpha[materialIdx] : __ldg(sc.materialAlpha + materialIdx);
        p->roughness = v2(a.x, a.y);
    } else {
        p->roughness = v2(roughnessToAlpha(m->roughness.x), roughnessToAlpha(m->roughness.y));
    }
}

// second stage: shade_vertex from the sampler on, for a vertex that is alive (not an emitter hit that ended the path)
HD void shade_scatter(const Scene& sc, const VertexRecord& rec, V3 rayDir, V3 throughput, int prevBsdfFlags, int frame,
                      int width, int height, int bounce, int maxDepth, ShadeOut* out) {
    out->hasEmission = false;
    out->nextPdf = 0.0f;
    out->emitted = v3(0.0f);
    out->direct = v3(0.0f);
    out->shadow.written = false;
    out->continues = false;
    out->bsdfFlags = prevBsdfFlags;
    if (bounce == 0) throughput = v3(1.0f);
    out->throughput = throughput;

    RTInteraction si;
    si.p = rec.p;
    si.gn = rec.gn;
    si.sn = rec.sn;
    si.sdpdu = rec.sdpdu;
    si.sdpdv = rec.sdpdv;
    si.wo = -rayDir;
    si.traceErrorOffset = (rec.flags & 1) ? -RT_TRACE_OFFSET : RT_TRACE_OFFSET;
    const bool hasMaterial = (rec.flags & 2) != 0;
    UberProps um;
    if (hasMaterial) {
        um.Kd = rec.Kd;
        um.Ks = rec.Ks;
        um.opacity = rec.opacity;
        untexturedUberProperties(sc, rec.flags >> 8, &um);
    }

    Sampler sampler = makeSampler(sc.sobolMatrices, (int)rec.px, frame, width, height, bounce);
    // direct lighting: one light, one sample (PathTracing.cl:106-134)
    float lightPdf = 0.0f;
    V3 wiL = v3(0.0f), Li = v3(0.0f);
    {
        uint32_t lightIdx = ((uint32_t)floorf(getSample1D(&sampler) * sc.numLights));
        lightIdx %= (uint32_t)sc.numLights;
        V2 u = getSample2D(&sampler);
        V3 unusedN, unusedP;
        Li = sampleLightLi((int)lightIdx, sc, &si, u, &unusedP, &unusedN, &wiL, &lightPdf, &out->shadow);
        lightPdf *= sc.lights[lightIdx].choicePdf;
    }
    {
        V3 L = v3(0.0f);
        if (hasMaterial) {
            V3 bsdf = evaluateUberBSDF(um, &si, si.wo, wiL, TRANSPORT_MODE_RADIANCE);
            bsdf *= absDot(wiL, si.sn);
            if (!isNearZero(lightPdf)) L = Li * bsdf / lightPdf;
        }
        out->direct = throughput * L;
    }
    if (bounce + 1 < maxDepth) {   // extend the path (PathTracing.cl:137-172)
        V2 bsdfSample = getSample2D(&sampler);
        if (hasMaterial) {
            float pdf = 0.0f;
            int sampledType, unused;
            V3 wi = v3(0.0f);
            V3 f = sampleUberBSDF(um, &si, bsdfSample, TRANSPORT_MODE_RADIANCE, BSDF_ALL, si.wo, &wi, &pdf, &unused,
                                  &sampledType);
            out->bsdfFlags = sampledType;
            if (!(isNearZero(pdf) || isBlack(f))) {
                f /= pdf;
                V3 tp = f * absDot(wi, si.sn);
                out->throughput = throughput * tp;
                float traceErrorOffset = si.traceErrorOffset;
                if ((sampledType & BSDF_TRANSMISSION) != 0 && dot(si.gn, wi) * signf(traceErrorOffset) < 0.0f)
                    traceErrorOffset *= -1.0f;
                out->nextO = si.p + si.gn * traceErrorOffset;
                out->nextD = wi;
                out->continues = true;
            }
        }
    }
}

// PathTracing.cl:13-35 GeneratePerspectiveRays: pixel CORNER uv, no +0.5 (E15)
HD V3 primary_ray_dir(const RTPinholeCamera* cam, int x, int y) {
    const float rx = 1.0f / cam->width, ry = 1.0f / cam->height;
    const float u = x * rx, v = y * ry;
    return lerpDirection(v3(cam->r00), v3(cam->r10), v3(cam->r11), v3(cam->r01), u, v);
}

}  // namespace mcrt
