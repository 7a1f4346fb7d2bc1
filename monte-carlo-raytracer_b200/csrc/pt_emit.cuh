// pt_emit.cuh -- emission sampling of the four light types and the pinhole camera's importance
// functions, used only by the bidirectional path tracer.
//   sampleLightLe / evalLightPdfLe            assets/kernels/lights.cl:148-253
//   evalPinholeCameraWe / PdfWe / sampleWi    assets/kernels/cameras.cl:8-71
// Same expression order as the reference (strict fp32, no contraction: see pt_math.cuh).
#pragma once
#include "pt_material.cuh"

namespace mcrt {

HD float uniformSpherePdf() { return PT_PI4_INV; }

// lights.cl:148-224.  The directional light's p / radius are set by the host to
// worldCentre - d * worldRadius and the world radius (scene/RTScene.cpp), so the disk covers the scene.
HD V3 sampleLightLe(int lightIdx, const Scene& sc, V2 u1, V2 u2, V3* rayOrigin, V3* rayDirection, V3* lightNormal,
                    float* pdfPos, float* pdfDir) {
    const RTLight* light = sc.lights + lightIdx;
    switch (light->type) {
        case RT_DIRECTIONAL_LIGHT: {
            V3 sp, sn;
            sampleDisk(v3(light->p), v3(light->d), light->radius, u1, &sp, &sn, pdfPos);
            *lightNormal = v3(light->d);
            *pdfDir = 1.0f;
            *rayOrigin = sp;
            *rayDirection = v3(light->d);
            return v3(light->intensity);
        }
        case RT_POINT_LIGHT: {
            *rayDirection = uniformSampleSphere(u1);
            *rayOrigin = v3(light->p);
            *lightNormal = *rayDirection;
            *pdfPos = 1.0f;
            *pdfDir = uniformSpherePdf();
            return v3(light->intensity);
        }
        case RT_DISK_AREA_LIGHT: {
            V3 sp, sn;
            sampleDisk(v3(light->p), v3(light->d), light->radius, u1, &sp, &sn, pdfPos);
            *lightNormal = sn;
            V3 w = cosineSampleHemisphere(u2);
            *pdfDir = cosineHemispherePdf(w.y);
            V3 v0 = computeOrthogonalVector(sn);
            V3 v1 = cross(v0, sn);
            *rayDirection = w.x * v0 + w.y * sn + w.z * v1;
            *rayOrigin = sp + sn * RT_TRACE_OFFSET;
            return v3(light->intensity);
        }
        case RT_TRIANGLE_MESH_AREA_LIGHT: {
            const RTShape* shape = sc.shapes + light->shapeId;
            const int nT = (int)shape->numTriangles;
            int triangleIdx = ((int)floorf(u1.x * nT)) % nT;
            u1.x = u1.x * nT - triangleIdx;
            uint32_t i0, i1, i2;
            V3 p0, p1, p2;
            loadTrianglePositions(sc, shape, triangleIdx, &p0, &p1, &p2, &i0, &i1, &i2);
            V3 sp, sn;
            sampleTriangle(p0, p1, p2, u1, &sp, &sn, pdfPos);
            *pdfPos = 1.0f / light->area;      // E10
            *lightNormal = sn;
            V3 w = cosineSampleHemisphere(u2);
            *pdfDir = cosineHemispherePdf(w.y);
            V3 v0 = computeOrthogonalVector(sn);
            V3 v1 = cross(v0, sn);
            *rayDirection = w.x * v0 + w.y * sn + w.z * v1;
            *rayOrigin = sp + sn * RT_TRACE_OFFSET;
            return v3(light->intensity);
        }
        default:
            return v3(0.0f);
    }
}

// lights.cl:226-253.  An unknown light type leaves both outputs untouched in the reference; here they
// are defined as zero.
HD void evalLightPdfLe(int lightIdx, const Scene& sc, V3 rayDirection, V3 lightNormal, float* pdfPos, float* pdfDir) {
    const RTLight* light = sc.lights + lightIdx;
    *pdfPos = 0.0f;
    *pdfDir = 0.0f;
    switch (light->type) {
        case RT_DIRECTIONAL_LIGHT:
            *pdfPos = 1.0f / light->area;
            *pdfDir = 0.0f;
            break;
        case RT_POINT_LIGHT:
            *pdfPos = 0.0f;
            *pdfDir = uniformSpherePdf();
            break;
        case RT_DISK_AREA_LIGHT:
        case RT_TRIANGLE_MESH_AREA_LIGHT:
            *pdfPos = 1.0f / light->area;
            *pdfDir = cosineHemispherePdf(dot(lightNormal, rayDirection));
            break;
        default:
            break;
    }
}

// cameras.cl:8-32
HD V3 evalPinholeCameraWe(const RTPinholeCamera* camera, V3 rayOrigin, V3 rayDirection, V2* normalizedImagePos) {
    float cosTheta = dot(rayDirection, v3(camera->direction));
    if (cosTheta <= 0.0f) return v3(0.0f);
    V3 pf = rayOrigin + rayDirection / cosTheta;
    float4 imgClip = transformVector4(camera->worldToClip, make_float4(pf.x, pf.y, pf.z, 1.0f));
    if (imgClip.x < -imgClip.w || imgClip.x > imgClip.w || imgClip.y < -imgClip.w || imgClip.y > imgClip.w)
        return v3(0.0f);
    if (normalizedImagePos) {
        normalizedImagePos->x = (imgClip.x / imgClip.w + 1.0f) * 0.5f;
        normalizedImagePos->y = (imgClip.y / imgClip.w + 1.0f) * 0.5f;
    }
    return v3(1.0f / (camera->area * cosTheta * cosTheta * cosTheta));
}

// cameras.cl:34-61
HD void evalPinholeCameraPdfWe(const RTPinholeCamera* camera, V3 rayOrigin, V3 rayDirection, float* pdfPos, float* pdfDir) {
    float cosTheta = dot(rayDirection, v3(camera->direction));
    *pdfPos = 0.0f;
    *pdfDir = 0.0f;
    if (cosTheta <= 0.0f) return;
    V3 pf = rayOrigin + rayDirection * (1.0f / cosTheta);
    float4 imgClip = transformVector4(camera->worldToClip, make_float4(pf.x, pf.y, pf.z, 1.0f));
    if (imgClip.x < -imgClip.w || imgClip.x > imgClip.w || imgClip.y < -imgClip.w || imgClip.y > imgClip.w) return;
    *pdfPos = 1.0f;
    *pdfDir = 1.0f / (camera->area * cosTheta * cosTheta * cosTheta);
}

// cameras.cl:63-71
HD V3 samplePinholeCameraWi(const RTPinholeCamera* camera, V3 p, V3* wi, float* pdf, V2* normalizedImagePos) {
    *wi = v3(camera->pos) - p;
    float dist = length(*wi);
    *wi /= dist;
    *pdf = (dist * dist) / absDot(v3(camera->direction), *wi);
    return evalPinholeCameraWe(camera, v3(camera->pos), -*wi, normalizedImagePos);
}

}  // namespace mcrt
