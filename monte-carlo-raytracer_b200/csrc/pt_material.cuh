// pt_material.cuh -- uber material (texture-modulated properties + BSDF evaluate / sample / pdf) and
// light sampling of the shading stages.
//
// Restates assets/kernels/materials.cl:68-156, bxdfs.cl:804-1053 and lights.cl:29-146 (PT) /
// :148-254 (BDPT emission sampling).  One deliberate, rounding-neutral difference from the
// reference: the texture-modulated properties are fetched ONCE per interaction and reused by
// evaluate and sample (the reference re-reads up to 7 textures in each, materials.cl:101-128); the
// fetched values and every later operation are identical.
#pragma once
#include "pt_bxdf.cuh"

namespace mcrt {

struct UberProps {          // materials.cl:68-74 RTUberMaterialProperties
    V3 Kd, Ks, Kr, opacity;
    V3 Kt;
    float KtW;
    V2 roughness;
    float eta;
};

// materials.cl:76-91
HD void getUberMaterialProperties(const Scene& sc, int materialIdx, const RTInteraction* si, UberProps* p) {
    const RTMaterial* m = sc.materials + materialIdx;
    float4 Kd_opacity = make_float4(1.f, 1.f, 1.f, 1.f);
    if (m->diffuseTexId != RT_INVALID_ID) Kd_opacity = readTexture2Df(sc, m->diffuseTexId, si->uv);
    p->Kd = v3(Kd_opacity) * v3(m->kd);
    V3 t = v3(1.0f);
    if (m->glossyTexId != RT_INVALID_ID) t = v3(readTexture2Df(sc, m->glossyTexId, si->uv));
    p->Ks = t * v3(m->ks);
    t = v3(1.0f);
    if (m->specReflectionTexId != RT_INVALID_ID) t = v3(readTexture2Df(sc, m->specReflectionTexId, si->uv));
    p->Kr = t * v3(m->kr);
    t = v3(1.0f);
    if (m->transmissionTexId != RT_INVALID_ID) t = v3(readTexture2Df(sc, m->transmissionTexId, si->uv));
    p->Kt = t * v3(m->kt);
    p->KtW = m->kt.w;
    t = v3(1.0f);
    if (m->opacityTexId != RT_INVALID_ID) t = v3(readTexture2Df(sc, m->opacityTexId, si->uv));
    p->opacity = t * v3(m->opacity) * Kd_opacity.w;
    p->eta = m->eta;
    if (m->iorTexId != RT_INVALID_ID) p->eta = readTexture2Df(sc, m->iorTexId, si->uv).x;
    if (m->roughnessTexId == RT_INVALID_ID && sc.materialAlpha) {
        const float2 a = sc.staged ? sc.materialAlpha[materialIdx] : __ldg(sc.materialAlpha + materialIdx);   // = the two roughnessToAlpha() below, precomputed
        p->roughness = v2(a.x, a.y);
    } else {
        V2 r = v2(m->roughness.x, m->roughness.y);
        if (m->roughnessTexId != RT_INVALID_ID) {
            float4 q = readTexture2Df(sc, m->roughnessTexId, si->uv);
            r = v2(q.x, q.y);
        }
        p->roughness = v2(roughnessToAlpha(r.x), roughnessToAlpha(r.y));
    }
}

// bxdfs.cl:804-827 evaluateUberBSDF
HD V3 evaluateUberBSDF(const UberProps& um, const RTInteraction* si, V3 woWorld, V3 wiWorld, int mode) {
    if (!isReflection(woWorld, wiWorld, si)) {
        bool isPerfectSpecularTransmission = um.KtW < 0.5f;
        if (isPerfectSpecularTransmission) return v3(0.0f);
        V3 wo = dirToShadingSpace(woWorld, si);
        V3 wi = dirToShadingSpace(wiWorld, si);
        V3 kt = um.Kt * um.opacity;
        return evaluateMicrofacetTransmission_TR(kt, mode, um.roughness, 1.0f, um.eta, wo, wi);
    }
    V3 wo = dirToShadingSpace(woWorld, si);
    V3 wi = dirToShadingSpace(wiWorld, si);
    V3 kd = um.Kd * um.opacity;
    V3 ks = um.Ks * um.opacity;
    return evaluateMicrofacetReflection_TR(ks, um.roughness, 1.0f, um.eta, wo, wi) + evaluateLambertianReflection(kd);
}

// bxdfs.cl:829-880 evaluateUberBSDF_Pdf
HD float evaluateUberBSDF_Pdf(const UberProps& um, const RTInteraction* si, V3 woWorld, V3 wiWorld, int type) {
    V3 wo = dirToShadingSpace(woWorld, si);
    V3 wi = dirToShadingSpace(wiWorld, si);
    if (isNearZero(wo.y)) return 0.0f;
    V3 t = v3(1.0f) - um.opacity;
    int numBxDFs = 0;
    V3 kd = um.Kd * um.opacity, ks = um.Ks * um.opacity, kt = um.Kt * um.opacity, kr = um.Kr * um.opacity;
    float pdf = 0.0f;
    if (isNotBlack(t) && BxDFMatchesFlags(BSDF_SPECULAR_TRANSMISSION, type)) ++numBxDFs;
    if (isNotBlack(kr) && BxDFMatchesFlags(BSDF_SPECULAR_REFLECTION, type)) ++numBxDFs;
    if (isNotBlack(kt)) {
        bool isPerfectSpecularTransmission = um.KtW < 0.5f;
        if (isPerfectSpecularTransmission) {
            if (BxDFMatchesFlags(BSDF_SPECULAR_TRANSMISSION, type)) ++numBxDFs;
        } else if (BxDFMatchesFlags(BSDF_MICROFACET_TRANSMISSION, type)) {
            pdf += evalMicrofacetTransmissionPdf_TR(wo, wi, um.roughness, 1.0f, um.eta);
            ++numBxDFs;
        }
    }
    if (isNotBlack(kd) && BxDFMatchesFlags(BSDF_LAMBERTIAN_REFLECTION, type)) {
        pdf += evaluateLambertianReflectionPdf(wo, wi);
        ++numBxDFs;
    }
    if (isNotBlack(ks) && BxDFMatchesFlags(BSDF_MICROFACET_REFLECTION, type)) {
        pdf += evalMicrofacetReflectionPdf_TR(wo, wi, normalize(wo + wi), um.roughness);
        ++numBxDFs;
    }
    if (numBxDFs > 1) pdf /= numBxDFs;
    return pdf;
}

// bxdfs.cl:892-1053 sampleUberBSDF
HD V3 sampleUberBSDF(const UberProps& um, const RTInteraction* si, V2 u, int mode, int type, V3 woWorld, V3* wiWorld,
                     float* pdf, int* numNonDeltaTypes, int* sampledType) {
    V3 t = v3(1.0f) - um.opacity;
    int numBxDFs = 0;
    V3 kd = um.Kd * um.opacity, ks = um.Ks * um.opacity, kt = um.Kt * um.opacity, kr = um.Kr * um.opacity;
    V3 wo = dirToShadingSpace(woWorld, si);
    V3 wi = v3(0.0f);   // E7: defined as zero until a lobe writes it
    bool isPerfectSpecularTransmission = um.KtW < 0.5f;
    *sampledType = BSDF_NONE;
    *numNonDeltaTypes = 0;

    const bool hasT = isNotBlack(t) && BxDFMatchesFlags(BSDF_SPECULAR_TRANSMISSION, type);
    const bool hasKd = isNotBlack(kd) && BxDFMatchesFlags(BSDF_LAMBERTIAN_REFLECTION, type);
    const bool hasKs = isNotBlack(ks) && BxDFMatchesFlags(BSDF_MICROFACET_REFLECTION, type);
    const bool hasKr = isNotBlack(kr) && BxDFMatchesFlags(BSDF_SPECULAR_REFLECTION, type);
    const bool hasKt = isNotBlack(kt);
    if (hasT) ++numBxDFs;
    if (hasKd) { (*numNonDeltaTypes)++; ++numBxDFs; }
    if (hasKs) { (*numNonDeltaTypes)++; ++numBxDFs; }
    if (hasKr) ++numBxDFs;
    if (hasKt) {
        if (isPerfectSpecularTransmission) {
            if (BxDFMatchesFlags(BSDF_SPECULAR_TRANSMISSION, type)) ++numBxDFs;
        } else if (BxDFMatchesFlags(BSDF_MICROFACET_TRANSMISSION, type)) {
            (*numNonDeltaTypes)++;
            ++numBxDFs;
        }
    }
    if (numBxDFs == 0) return v3(0.0f);

    int chosenBxDFIdx = min((int)floorf(u.x * numBxDFs), numBxDFs - 1);
    u.x = u.x * numBxDFs - chosenBxDFIdx;

    V3 f = v3(0.0f);
    *pdf = 0.0f;
    bool isSamplingSpecular = false;

    if (hasT) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleSpecularTransmission(t, 1.0f, 1.0f, mode, wo, &wi, pdf);
            *sampledType |= BSDF_SPECULAR_TRANSMISSION;
            isSamplingSpecular = true;
        }
    }
    if (isPerfectSpecularTransmission && hasKt && BxDFMatchesFlags(BSDF_SPECULAR_TRANSMISSION, type)) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleSpecularTransmission(kt, 1.0f, um.eta, mode, wo, &wi, pdf);
            *sampledType |= BSDF_SPECULAR_TRANSMISSION;
            isSamplingSpecular = true;
        }
    }
    if (hasKr) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleSpecularReflection_Dielectric(kr, 1.0f, um.eta, wo, &wi, pdf);
            *sampledType |= BSDF_SPECULAR_REFLECTION;
            isSamplingSpecular = true;
        }
    }
    // NB (bxdfs.cl:999): this test does not look at Kt.w, exactly like the reference
    if (hasKt && BxDFMatchesFlags(BSDF_MICROFACET_TRANSMISSION, type)) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleMicrofacetTransmission_TR(u, kt, mode, um.roughness, 1.0f, um.eta, wo, &wi, pdf);
            *sampledType |= BSDF_MICROFACET_TRANSMISSION;
        }
    }
    bool isLambertianEvaluationRequested = false;
    if (hasKd) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleLambertianReflection(kd, u, wo, &wi, pdf);
            *sampledType |= BSDF_LAMBERTIAN_REFLECTION;
        } else if (!isSamplingSpecular && isSameHemisphere(wi, wo)) {
            isLambertianEvaluationRequested = true;
        }
    }
    if (hasKs) {
        if (chosenBxDFIdx-- == 0) {
            f += sampleMicrofacetReflection_TR(u, ks, um.roughness, 1.0f, um.eta, wo, &wi, pdf);
            *sampledType |= BSDF_MICROFACET_REFLECTION;
        } else if (!isSamplingSpecular && isSameHemisphere(wi, wo)) {
            f += evaluateMicrofacetReflection_TR(ks, um.roughness, 1.0f, um.eta, wo, wi);
            *pdf += evalMicrofacetReflectionPdf_TR(wo, wi, normalize(wo + wi), um.roughness);
        }
    }
    if (isLambertianEvaluationRequested) {
        f += evaluateLambertianReflection(kd);
        *pdf += evaluateLambertianReflectionPdf(wo, wi);
    }
    *pdf = fdiv(*pdf, (float)numBxDFs);
    *wiWorld = dirFromShadingToWorld(wi, si);
    return f;
}

// materials.cl:161-179 hasMaterialNonDeltaComponents (BDPT)
HD bool hasMaterialNonDeltaComponents(const Scene& sc, int materialIdx, const RTInteraction* si) {
    const RTMaterial* m = sc.materials + materialIdx;
    V3 Kd = v3(m->kd), Ks = v3(m->ks), op = v3(m->opacity);
    if (m->diffuseTexId != RT_INVALID_ID) Kd = v3(readTexture2Df(sc, m->diffuseTexId, si->uv));
    if (m->glossyTexId != RT_INVALID_ID) Ks = v3(readTexture2Df(sc, m->glossyTexId, si->uv));
    if (m->opacityTexId != RT_INVALID_ID) op = v3(readTexture2Df(sc, m->opacityTexId, si->uv));
    V3 kd = Kd * op, ks = Ks * op;
    return (!isBlack(kd) || !isBlack(ks));
}

// ------------------------------------------------------------------ lights (lights.cl)
struct ShadowRay {
    V3 o, d;
    float tmax;
    bool written;
};

// lights.cl:29-39
HD V3 evalLightLe(const RTLight* light, V3 gn, V3 w) {
    switch (light->type) {
        case RT_DISK_AREA_LIGHT:
        case RT_TRIANGLE_MESH_AREA_LIGHT:
            return dot(gn, w) > 0.0f ? v3(light->intensity) : v3(0.0f);
        default:
            return v3(0.0f);
    }
}

// lights.cl:45-146 sampleLightLi.  `sr->written` tells whether the reference's setRay() ran (when it
// did not, the reference's shadow ray keeps stale contents and the contribution is zero anyway).
HD V3 sampleLightLi(int lightIdx, const Scene& sc, const RTInteraction* it, V2 u, V3* lightPosition, V3* lightNormal,
                    V3* wi, float* pdf, ShadowRay* sr) {
    const RTLight* light = sc.lights + lightIdx;
    sr->written = false;
    switch (light->type) {
        case RT_DIRECTIONAL_LIGHT: {
            *wi = -v3(light->d);
            *lightNormal = v3(0.0f);
            *lightPosition = it->p + *wi * light->radius * 2.0f;
            *pdf = 1.0f;
            sr->o = it->p + it->gn * it->traceErrorOffset;
            sr->tmax = 1000.0f;
            sr->d = *wi;
            sr->written = true;
            return v3(light->intensity);
        }
        case RT_POINT_LIGHT: {
            *wi = v3(light->p) - it->p;
            float distSq = dot(*wi, *wi);
            if (isNearZero(distSq)) return v3(0.0f);
            float dist = sqrtf(distSq);
            *wi /= dist;
            *lightNormal = v3(0.0f);
            *pdf = 1.0f;
            V3 rayOrigin = it->p + it->gn * it->traceErrorOffset;
            *lightPosition = v3(light->p);
            sr->o = rayOrigin;
            sr->tmax = dist;
            sr->d = *wi;
            sr->written = true;
            return v3(light->intensity) / distSq;
        }
        case RT_DISK_AREA_LIGHT: {
            V3 sp, sn;
            sampleDisk(v3(light->p), v3(light->d), light->radius, u, &sp, &sn, pdf);
            *lightNormal = sn;
            *lightPosition = sp;
            V3 rayOrigin = it->p + it->gn * it->traceErrorOffset;
            V3 rayTarget = sp + sn * RT_TRACE_OFFSET;
            *wi = normalize(rayTarget - rayOrigin);
            float distSq = distanceSquared(sp, it->p);
            float c = absDot(sn, -*wi);
            if (isNearZero(c)) {
                *pdf = 0.0f;
                return v3(0.0f);
            }
            *pdf *= distSq / c;
            sr->o = rayOrigin;
            sr->tmax = distance(rayOrigin, rayTarget);
            sr->d = *wi;
            sr->written = true;
            return dot(sn, -*wi) > 0.0f ? v3(light->intensity) : v3(0.0f);
        }
        case RT_TRIANGLE_MESH_AREA_LIGHT: {
            const RTShape* shape = sc.shapes + light->shapeId;
            const int nT = (int)shape->numTriangles;
            int triangleIdx = ((int)floorf(u.x * nT)) % nT;
            u.x = u.x * nT - triangleIdx;
            uint32_t i0, i1, i2;
            V3 p0, p1, p2;
            loadTrianglePositions(sc, shape, triangleIdx, &p0, &p1, &p2, &i0, &i1, &i2);
            V3 sp, sn;
            sampleTriangle(p0, p1, p2, u, &sp, &sn, pdf);
            *lightNormal = sn;
            *lightPosition = sp;
            *pdf = 1.0f / light->area;     // E10: uniform triangle pick, pdf of the whole area
            V3 rayOrigin = it->p + it->gn * it->traceErrorOffset;
            V3 rayTarget = sp + sn * RT_TRACE_OFFSET;
            *wi = normalize(sp - it->p);
            float distSq = distanceSquared(sp, it->p);
            float c = absDot(sn, -*wi);
            if (isNearZero(c)) {
                *pdf = 0.0f;
                return v3(0.0f);
            }
            *pdf *= distSq / c;
            sr->o = rayOrigin;
            sr->tmax = distance(rayOrigin, rayTarget);
            sr->d = *wi;
            sr->written = true;
            return dot(sn, -*wi) > 0.0f ? v3(light->intensity) : v3(0.0f);
        }
        default:
            return v3(0.0f);
    }
}

}  // namespace mcrt
