// pt_bxdf.cuh -- the uber BSDF of the shading stages (evaluate / sample / pdf).
//
// Restates (own code, same arithmetic and evaluation order) assets/kernels/bxdfs.cl: shading
// frame helpers :22-112, Fresnel :159-190, reflect/refract :228-245, specular lobes :259-307,
// Lambert :317-347, Trowbridge-Reitz D/Lambda/G :385-474, microfacet reflection / transmission
// :481-500,563-588,647-762, uber evaluate / pdf / sample :804-1053.
// Shading frame: x = tangent (sdpdu), y = normal (sn), z = binormal (sdpdv) -- y-up.
// Documented quirks kept on purpose: E7 `wi` is zero until a lobe writes it,
// E8 evalSinPhi always divides, E9 the `wh.z < 0` flip in microfacet transmission.
#pragma once
#include "pt_scene.cuh"

namespace mcrt {

enum BxDFType {
    BSDF_NONE = 0,
    BSDF_REFLECTION = 1 << 0,
    BSDF_TRANSMISSION = 1 << 1,
    BSDF_DIFFUSE = 1 << 2,
    BSDF_GLOSSY = 1 << 3,
    BSDF_SPECULAR = 1 << 4,
    BSDF_SPECULAR_REFLECTION = BSDF_REFLECTION | BSDF_SPECULAR,
    BSDF_SPECULAR_TRANSMISSION = BSDF_TRANSMISSION | BSDF_SPECULAR,
    BSDF_LAMBERTIAN_REFLECTION = BSDF_REFLECTION | BSDF_DIFFUSE,
    BSDF_MICROFACET_REFLECTION = BSDF_REFLECTION | BSDF_GLOSSY,
    BSDF_MICROFACET_TRANSMISSION = BSDF_TRANSMISSION | BSDF_GLOSSY,
    BSDF_ALL = BSDF_DIFFUSE | BSDF_GLOSSY | BSDF_SPECULAR | BSDF_REFLECTION | BSDF_TRANSMISSION
};
enum TransportMode { TRANSPORT_MODE_RADIANCE = 0, TRANSPORT_MODE_IMPORTANCE = 1 };

HD float evalCosTheta(V3 w) { return w.y; }
HD float evalCosSqTheta(V3 w) { return w.y * w.y; }
HD float evalAbsCosTheta(V3 w) { return fabsf(w.y); }
HD float evalSinSqTheta(V3 w) { return fmaxf(0.0f, 1.0f - evalCosSqTheta(w)); }
HD float evalSinTheta(V3 w) { return sqrtf(evalSinSqTheta(w)); }
HD float evalTanTheta(V3 w) { return evalSinTheta(w) / evalCosTheta(w); }
HD float evalTanSqTheta(V3 w) { return evalSinSqTheta(w) / evalCosSqTheta(w); }
HD float evalCosPhi(V3 w) {
    float st = evalSinTheta(w);
    return st == 0 ? 1.0f : clampf(fdiv(w.x, st), -1.0f, 1.0f);
}
// bxdfs.cl:35-39 compares a function pointer with 0, so the division always happens (E8);
// clamp(NaN) is defined as fmin(fmax(x,-1),1) = -1 here and in the oracle.
HD float evalSinPhi(V3 w) {
    float st = evalSinTheta(w);
    return clampf(fdiv(w.z, st), -1.0f, 1.0f);
}
HD float evalCosSqPhi(V3 w) { return evalCosPhi(w) * evalCosPhi(w); }
HD float evalSinSqPhi(V3 w) { return evalSinPhi(w) * evalSinPhi(w); }
HD bool isSameHemisphere(V3 a, V3 b) { return a.y * b.y > 0.0f; }

#define BLACK_EPS 0.000001f
HD bool isBlack(V3 c) { return c.x < BLACK_EPS && c.y < BLACK_EPS && c.z < BLACK_EPS; }
HD bool isNotBlack(V3 c) { return c.x > BLACK_EPS || c.y > BLACK_EPS || c.z > BLACK_EPS; }

HD V3 dirToShadingSpace(V3 w, const RTInteraction* si) { return v3(dot(si->sdpdu, w), dot(si->sn, w), dot(si->sdpdv, w)); }
HD V3 dirFromShadingToWorld(V3 w, const RTInteraction* si) {
    const V3 t = si->sdpdu, n = si->sn, bn = si->sdpdv;
    return v3(t.x * w.x + n.x * w.y + bn.x * w.z, t.y * w.x + n.y * w.y + bn.y * w.z, t.z * w.x + n.z * w.y + bn.z * w.z);
}
HD bool isReflection(V3 woWorld, V3 wiWorld, const RTInteraction* si) {
    return dot(si->gn, woWorld) * dot(si->gn, wiWorld) > 0.0f;
}
HD bool BxDFMatchesFlags(int bxdfType, int flags) { return (bxdfType & flags) == bxdfType; }

// bxdfs.cl:159-190
HD_BXDF float evaluateFresnelDielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = clampf(cosThetaI, -1.0f, 1.0f);
    if (cosThetaI <= 0.0f) {
        float h = etaI;
        etaI = etaT;
        etaT = h;
        cosThetaI = fabsf(cosThetaI);
    }
    float sinThetaI = sqrtf(fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1.0f) return 1.0f;
    float cosThetaT = sqrtf(fmaxf(0.0f, 1.0f - sinThetaT * sinThetaT));
    float rparl = fdiv((etaT * cosThetaI) - (etaI * cosThetaT), (etaT * cosThetaI) + (etaI * cosThetaT));
    float rperp = fdiv((etaI * cosThetaI) - (etaT * cosThetaT), (etaI * cosThetaI) + (etaT * cosThetaT));
    return (rparl * rparl + rperp * rperp) * 0.5f;
}

// bxdfs.cl:228-245
HD V3 reflectOut(V3 wo, V3 n) { return -wo + 2.0f * dot(n, wo) * n; }
HD bool refractDir(V3 wi, V3 n, float eta, V3* wt) {
    float cosThetaI = dot(n, wi);
    float sin2ThetaI = fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI);
    float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1.0f) return false;
    float cosThetaT = sqrtf(1.0f - sin2ThetaT);
    *wt = -eta * wi + (eta * cosThetaI - cosThetaT) * n;
    return true;
}

// bxdfs.cl:259-268
HD V3 sampleSpecularReflection_Dielectric(V3 R, float etaI, float etaT, V3 wo, V3* wi, float* pdf) {
    *wi = v3(-wo.x, wo.y, -wo.z);
    *pdf = 1.0f;
    float F = evaluateFresnelDielectric(evalCosTheta(*wi), etaI, etaT);
    return F * R / evalAbsCosTheta(*wi);
}

// bxdfs.cl:288-307
HD V3 sampleSpecularTransmission(V3 T, float etaA, float etaB, int mode, V3 wo, V3* wi, float* pdf) {
    bool isEntering = evalCosTheta(wo) > 0.0f;
    float etaI = isEntering ? etaA : etaB;
    float etaT = isEntering ? etaB : etaA;
    V3 n = v3(0.0f, 1.0f, 0.0f) * signf(wo.y);
    if (!refractDir(wo, n, etaI / etaT, wi)) return v3(0.0f);
    *pdf = 1.0f;
    V3 ft = T * (1.0f - evaluateFresnelDielectric(evalCosTheta(*wi), etaA, etaB));
    if (mode == TRANSPORT_MODE_RADIANCE) ft *= (etaI * etaI) / (etaT * etaT);
    return ft / evalAbsCosTheta(*wi);
}

// bxdfs.cl:317-347
HD void sampleCosineHemisphere(V2 u, V3 wo, V3* wi, float* pdf) {
    *wi = cosineSampleHemisphere(u);
    if (wo.y < 0.0f) wi->y *= -1.0f;
    *pdf = evalAbsCosTheta(*wi) * PT_PI_INV;
}
HD float evaluateLambertianReflectionPdf(V3 wo, V3 wi) { return isSameHemisphere(wo, wi) ? evalAbsCosTheta(wi) * PT_PI_INV : 0.0f; }
HD V3 evaluateLambertianReflection(V3 R) { return R * PT_PI_INV; }
HD V3 sampleLambertianReflection(V3 R, V2 u, V3 wo, V3* wi, float* pdf) {
    sampleCosineHemisphere(u, wo, wi, pdf);
    return R * PT_PI_INV;
}

// bxdfs.cl:385-390
HD float roughnessToAlpha(float roughness) {
    roughness = fmaxf(roughness, 1e-3f);
    float x = log_ref(roughness);
    return 1.62142f + 0.819955f * x + 0.1734f * x * x + 0.0171201f * x * x * x + 0.000640711f * x * x * x * x;
}

// bxdfs.cl:406-415
HD_BXDF float computeTrowbridgeReitzDistribution(V3 wh, V2 alpha) {
    float tan2Theta = evalTanSqTheta(wh);
    if (isinf(tan2Theta)) return 0.0f;
    const float cos4Theta = evalCosSqTheta(wh) * evalCosSqTheta(wh);
    float e = (evalCosSqPhi(wh) / (alpha.x * alpha.x) + evalSinSqPhi(wh) / (alpha.y * alpha.y)) * tan2Theta;
    return 1.0f / (PT_PI * alpha.x * alpha.y * cos4Theta * (1.0f + e) * (1.0f + e));
}
// bxdfs.cl:435-445
HD_BXDF float computeTrowbridgeReitzDistributionLambda(V3 w, V2 alpha) {
    float absTanTheta = fabsf(evalTanTheta(w));
    if (isinf(absTanTheta)) return 0.0f;
    float alphaW = sqrtf(evalCosSqPhi(w) * alpha.x * alpha.x + evalSinSqPhi(w) * alpha.y * alpha.y);
    float alpha2Tan2Theta = (alphaW * absTanTheta) * (alphaW * absTanTheta);
    return (-1.0f + sqrtf(1.f + alpha2Tan2Theta)) / 2.0f;
}
// bxdfs.cl:462-474
HD float computeTrowbridgeReitzDistributionG(V3 wo, V3 wi, V2 alpha) {
    return 1.0f / (1.0f + computeTrowbridgeReitzDistributionLambda(wo, alpha) + computeTrowbridgeReitzDistributionLambda(wi, alpha));
}

// bxdfs.cl:481-500
HD_BXDF V3 evaluateMicrofacetReflection_TR(V3 R, V2 alpha, float etaI, float etaT, V3 wo, V3 wi) {
    float cosThetaO = evalAbsCosTheta(wo);
    float cosThetaI = evalAbsCosTheta(wi);
    V3 wh = wi + wo;
    if (cosThetaI == 0.0f || cosThetaO == 0.0f) return v3(0.0f);
    if (wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f) return v3(0.0f);
    wh = normalize(wh);
    float F = evaluateFresnelDielectric(dot(wi, wh), etaI, etaT);
    return R * computeTrowbridgeReitzDistribution(wh, alpha) * computeTrowbridgeReitzDistributionG(wo, wi, alpha) * F /
           (4.0f * cosThetaI * cosThetaO);
}

// bxdfs.cl:563-588
HD_BXDF V3 evaluateMicrofacetTransmission_TR(V3 T, int mode, V2 alpha, float etaI, float etaT, V3 wo, V3 wi) {
    if (isSameHemisphere(wo, wi)) return v3(0.0f);
    float cosThetaO = evalCosTheta(wo);
    float cosThetaI = evalCosTheta(wi);
    if (cosThetaI == 0.0f || cosThetaO == 0.0f) return v3(0.0f);
    float eta = evalCosTheta(wo) > 0.0f ? (etaT / etaI) : (etaI / etaT);
    V3 wh = normalize(wo + wi * eta);
    if (wh.z < 0) wh = -wh;   // E9: z although the normal is y
    float F = evaluateFresnelDielectric(dot(wo, wh), etaI, etaT);
    float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    float factor = (mode == TRANSPORT_MODE_RADIANCE) ? (1.0f / eta) : 1.0f;
    return (v3(1.0f) - v3(F)) * T *
           fabsf(computeTrowbridgeReitzDistribution(wh, alpha) * computeTrowbridgeReitzDistributionG(wo, wi, alpha) * eta * eta *
                 absDot(wi, wh) * absDot(wo, wh) * factor * factor / (cosThetaI * cosThetaO * sqrtDenom * sqrtDenom));
}

// bxdfs.cl:647-675
HD_BXDF V3 sampleTrowbridgeReitzDistribution_wh(V2 u, V3 wo, V2 alpha) {
    float cosTheta = 0.0f;
    float phi = (2.0f * PT_PI) * u.y;
    if (alpha.x == alpha.y) {
        float tanTheta2 = alpha.x * alpha.x * u.x / (1.0f - u.x);
        cosTheta = 1.0f / sqrtf(1.0f + tanTheta2);
    } else {
        phi = atanf(alpha.y / alpha.x * tanf(2.0f * PT_PI * u.y + 0.5f * PT_PI));
        if (u.y > .5f) phi += PT_PI;
        float sinPhi = sin_ref(phi);
        float cosPhi = cos_ref(phi);
        const float alphax2 = alpha.x * alpha.x, alphay2 = alpha.y * alpha.y;
        const float alpha2 = 1.0f / (cosPhi * cosPhi / alphax2 + sinPhi * sinPhi / alphay2);
        float tanTheta2 = alpha2 * u.x / (1.0f - u.x);
        cosTheta = 1.0f / sqrtf(1.0f + tanTheta2);
    }
    float sinTheta = sqrtf(fmaxf(0.0f, 1.0f - cosTheta * cosTheta));
    V3 wh = v3(sinTheta * cos_ref(phi), cosTheta, sinTheta * sin_ref(phi));
    if (!isSameHemisphere(wo, wh)) wh = -wh;
    return wh;
}

HD float evalTrowbridgeReitzPdf_wh(V3 wh, V2 alpha) { return computeTrowbridgeReitzDistribution(wh, alpha) * evalAbsCosTheta(wh); }
// bxdfs.cl:695-701
HD float evalMicrofacetReflectionPdf_TR(V3 wo, V3 wi, V3 wh, V2 alpha) {
    if (!isSameHemisphere(wo, wi)) return 0.0f;
    return evalTrowbridgeReitzPdf_wh(wh, alpha) / (4.0f * dot(wo, wh));
}
// bxdfs.cl:717-729
HD float evalMicrofacetTransmissionPdf_TR(V3 wo, V3 wi, V2 alpha, float etaA, float etaB) {
    if (isSameHemisphere(wo, wi)) return 0.0f;
    float eta = evalCosTheta(wo) > 0.0f ? (etaB / etaA) : (etaA / etaB);
    V3 wh = normalize(wo + wi * eta);
    float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    float dwh_dwi = fabsf((eta * eta * dot(wi, wh)) / (sqrtDenom * sqrtDenom));
    return evalTrowbridgeReitzPdf_wh(wh, alpha) * dwh_dwi;
}

// bxdfs.cl:731-749
HD V3 sampleMicrofacetReflection_TR(V2 u, V3 R, V2 alpha, float etaI, float etaT, V3 wo, V3* wi, float* pdf) {
    V3 wh = sampleTrowbridgeReitzDistribution_wh(u, wo, alpha);
    *wi = reflectOut(wo, wh);
    if (!isSameHemisphere(wo, *wi)) return v3(0.0f);
    *pdf = evalMicrofacetReflectionPdf_TR(wo, *wi, wh, alpha);
    return evaluateMicrofacetReflection_TR(R, alpha, etaI, etaT, wo, *wi);
}
// bxdfs.cl:751-762
HD V3 sampleMicrofacetTransmission_TR(V2 u, V3 T, int mode, V2 alpha, float etaA, float etaB, V3 wo, V3* wi, float* pdf) {
    V3 wh = sampleTrowbridgeReitzDistribution_wh(u, wo, alpha);
    float eta = evalCosTheta(wo) > 0.0f ? (etaA / etaB) : (etaB / etaA);
    if (!refractDir(wo, wh, eta, wi)) return v3(0.0f);
    *pdf = evalMicrofacetTransmissionPdf_TR(wo, *wi, alpha, etaA, etaB);
    return evaluateMicrofacetTransmission_TR(T, mode, alpha, etaA, etaB, wo, *wi);
}

}  // namespace mcrt
