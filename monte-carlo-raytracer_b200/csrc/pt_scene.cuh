// pt_scene.cuh -- device-side scene view, surface interaction and textures of the shading stages.
//
// Data stays in the reference's device layouts (assets/kernels/kernel_data.h:36-152,317-326) so the
// host uploads RTHostScene arrays unchanged: RTShape 160 B, RTMaterial 128 B, RTLight 80 B,
// TextureDesc2D 16 B, float3 pools with a 16-byte stride.  tangents / binormals / colors are never
// read on this path and are not carried.
// Restates geometry.cl:9-28,177-215, textures.cl:70-125,204-241, materials.cl:14-30.
#pragma once
#include "pt_math.cuh"

namespace mcrt {

// Compile-time switch of the per-triangle shading records (Scene::triShade).  Measured on C3 (profiles/r02_shade.md):
// neutral for the default shade kernel (1.991 vs 1.998 ms per step), +6 % for the sorted one, and the second code
// path costs 1.5 K SASS instructions and registers in every kernel that shades -- so it is compiled out by default
// (build with -DPT_TRI_RECORDS=1 and run with MCRT_TRI_RECORDS=1 to use it).
#ifndef PT_TRI_RECORDS
#define PT_TRI_RECORDS 0
#endif

struct RTShape {            // kernel_data.h:36-52
    float4 toWorld[4];
    float4 toWorldInvT[4];
    uint32_t startIdx, startVertex, numTriangles;
    int materialId, lightID;
    float area;
    int pad[2];
};
struct RTMaterial {         // kernel_data.h:87-113
    float4 kd, ks, kr, kt, opacity;
    float2 roughness;
    float eta;
    int type;
    int normalMapId, diffuseTexId, glossyTexId, specReflectionTexId, transmissionTexId, opacityTexId,
        roughnessTexId, iorTexId;
};
struct RTLight {            // kernel_data.h:137-152
    float4 d, p, intensity;
    float radius, area, choicePdf;
    int shapeId, type, flags;
    int pad[2];
};
struct TextureDesc2D {      // kernel_data.h:317-326
    unsigned short width, height, numMipLevels, format, wrap, pad;
    uint32_t memOffset;
};
struct RTPinholeCamera {    // kernel_data.h:246-264
    float4 worldToClip[4];
    float4 r00, r10, r11, r01, pos, direction;
    uint32_t width, height;
    float area;
    int padding;
};
static_assert(sizeof(RTShape) == 160 && sizeof(RTMaterial) == 128 && sizeof(RTLight) == 80, "layout");
static_assert(sizeof(TextureDesc2D) == 16 && sizeof(RTPinholeCamera) == 176, "layout");

enum { RT_DIRECTIONAL_LIGHT = 0, RT_POINT_LIGHT = 1, RT_DISK_AREA_LIGHT = 2, RT_TRIANGLE_MESH_AREA_LIGHT = 3 };
enum { TEX_WRAP_REPEAT = 0, TEX_WRAP_MIRRORED_REPEAT = 1, TEX_WRAP_CLAMP_TO_EDGE = 2, TEX_WRAP_CLAMP_TO_BORDER = 3 };
enum {
    RT_LIGHT_FLAG_DELTA_POSITION = 1 << 1, RT_LIGHT_FLAG_DELTA_DIRECTION = 1 << 2, RT_LIGHT_FLAG_AREA = 1 << 3,
    RT_LIGHT_FLAG_INFINITE = 1 << 4,
    RT_LIGHT_FLAG_DELTA_LIGHT = RT_LIGHT_FLAG_DELTA_POSITION | RT_LIGHT_FLAG_DELTA_DIRECTION
};

struct Scene {              // kernel_data.h:353-370 (device pointers)
    const RTShape* __restrict__ shapes;
    const uint32_t* __restrict__ indices;
    const float4* __restrict__ positions;
    const float2* __restrict__ uvs;
    const float4* __restrict__ normals;
    const TextureDesc2D* __restrict__ textures2D;
    const unsigned char* __restrict__ texData2D;
    const RTLight* __restrict__ lights;
    const RTMaterial* __restrict__ materials;
    const RTPinholeCamera* __restrict__ camera;
    const uint32_t* __restrict__ sobolMatrices;   // NULL = random sampler
    // roughnessToAlpha(material.roughness) of every material, computed once on the device by the same function
    // (k_material_alpha); read instead of re-evaluating two logf polynomials per material lookup when the
    // material has no roughness texture.  NULL = always evaluate.
    const float2* __restrict__ materialAlpha;
    // Per-triangle shading records (k_build_tri_shade), 9 float4 per scene triangle in index-array order
    // (record = startIdx / 3 + prim), or NULL: world-space positions, transformed normals, uvs, the geometric
    // normal and the uv partial derivatives -- everything computeSurfaceInteraction derives from the triangle alone,
    // computed once by the same expressions.  A vertex then needs ONE round of independent 16-byte loads instead
    // of the dependent chain shape -> indices -> 3 x (position, normal, uv) and two matrix products per vertex.
    const float4* __restrict__ triShade;
    int numLights;
    int staged;   // 1 = shapes / materials / lights / textures2D / materialAlpha point into SHARED memory (no __ldg)
};

struct RTInteraction {      // kernel_data.h:162-200, the fields this path uses
    V3 wo, p;
    V2 uv;
    float traceErrorOffset;
    int shapeIdx;
    V3 gn, sn, dpdu, dpdv, sdpdu, sdpdv;
};

// geometry.cl:9-28
HD void computeTrianglePartialDerivates(V2 uv0, V2 uv1, V2 uv2, V3 p0, V3 p1, V3 p2, V3 normal, V3* dpdu, V3* dpdv) {
    V2 duv02 = uv0 - uv2;
    V2 duv12 = uv1 - uv2;
    V3 dp02 = p0 - p2;
    V3 dp12 = p1 - p2;
    float det = duv02.x * duv12.y - duv02.y * duv12.x;
    if (isNotNearZero(det)) {
        float invdet = 1.0f / det;
        *dpdu = (duv12.y * dp02 - duv02.y * dp12) * invdet;
        *dpdv = -(-duv12.x * dp02 + duv02.x * dp12) * invdet;
    } else {
        *dpdu = normalize(computeOrthogonalVector(normal));
        *dpdv = normalize(cross(normal, *dpdu));
    }
}

HD void loadTrianglePositions(const Scene& sc, const RTShape* shape, int primIdx, V3* p0, V3* p1, V3* p2,
                              uint32_t* i0, uint32_t* i1, uint32_t* i2) {
    if (PT_TRI_RECORDS && sc.triShade) {      // world-space vertices straight from the per-triangle record (the indices are not needed)
        const float4* r = sc.triShade + 9 * (size_t)(shape->startIdx / 3u + (uint32_t)primIdx);
        *p0 = v3(__ldg(r + 0));
        *p1 = v3(__ldg(r + 1));
        *p2 = v3(__ldg(r + 2));
        *i0 = *i1 = *i2 = 0;
        return;
    }
    const uint32_t base = shape->startIdx + 3u * (uint32_t)primIdx;
    *i0 = sc.indices[base];
    *i1 = sc.indices[base + 1];
    *i2 = sc.indices[base + 2];
    *p0 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + shape->startVertex + *i0)));
    *p1 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + shape->startVertex + *i1)));
    *p2 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + shape->startVertex + *i2)));
}

// The triangle-only part of computeSurfaceInteraction (geometry.cl:177-215), evaluated once per triangle.
struct TriShade {
    V3 p0, p1, p2, n0, n1, n2, gn, dpdu, dpdv;
    V2 uv0, uv1, uv2;
    bool hasDerivatives;    // the uv determinant is not near zero: dpdu / dpdv do not depend on the shading normal
};
HD TriShade computeTriShade(const Scene& sc, int shapeId, int primIdx) {
    TriShade t;
    const RTShape* shape = sc.shapes + shapeId;
    const uint32_t base = shape->startIdx + 3u * (uint32_t)primIdx;
    const uint32_t i0 = sc.indices[base], i1 = sc.indices[base + 1], i2 = sc.indices[base + 2];
    const uint32_t sv = shape->startVertex;
    t.p0 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + sv + i0)));
    t.p1 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + sv + i1)));
    t.p2 = transformPoint3(shape->toWorld, v3(__ldg(sc.positions + sv + i2)));
    const float2 a = __ldg(sc.uvs + sv + i0), b = __ldg(sc.uvs + sv + i1), c = __ldg(sc.uvs + sv + i2);
    t.uv0 = v2(a.x, a.y);
    t.uv1 = v2(b.x, b.y);
    t.uv2 = v2(c.x, c.y);
    t.n0 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i0)));
    t.n1 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i1)));
    t.n2 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i2)));
    t.gn = normalize(cross(t.p0 - t.p2, t.p1 - t.p2));
    const V2 duv02 = t.uv0 - t.uv2, duv12 = t.uv1 - t.uv2;
    const float det = duv02.x * duv12.y - duv02.y * duv12.x;
    t.hasDerivatives = isNotNearZero(det);
    t.dpdu = t.dpdv = v3(0.0f);
    if (t.hasDerivatives) computeTrianglePartialDerivates(t.uv0, t.uv1, t.uv2, t.p0, t.p1, t.p2, v3(0.0f), &t.dpdu, &t.dpdv);
    return t;
}
HD void storeTriShade(float4* __restrict__ r, const TriShade& t) {
    r[0] = make_float4(t.p0.x, t.p0.y, t.p0.z, t.uv0.x);
    r[1] = make_float4(t.p1.x, t.p1.y, t.p1.z, t.uv0.y);
    r[2] = make_float4(t.p2.x, t.p2.y, t.p2.z, t.uv1.x);
    r[3] = make_float4(t.n0.x, t.n0.y, t.n0.z, t.uv1.y);
    r[4] = make_float4(t.n1.x, t.n1.y, t.n1.z, t.uv2.x);
    r[5] = make_float4(t.n2.x, t.n2.y, t.n2.z, t.uv2.y);
    r[6] = make_float4(t.gn.x, t.gn.y, t.gn.z, t.hasDerivatives ? 1.0f : 0.0f);
    r[7] = make_float4(t.dpdu.x, t.dpdu.y, t.dpdu.z, 0.0f);
    r[8] = make_float4(t.dpdv.x, t.dpdv.y, t.dpdv.z, 0.0f);
}
HD TriShade loadTriShade(const float4* __restrict__ r) {
    const float4 a = __ldg(r + 0), b = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3), e = __ldg(r + 4), f = __ldg(r + 5),
                 g = __ldg(r + 6), h = __ldg(r + 7), i = __ldg(r + 8);
    TriShade t;
    t.p0 = v3(a);
    t.p1 = v3(b);
    t.p2 = v3(c);
    t.n0 = v3(d);
    t.n1 = v3(e);
    t.n2 = v3(f);
    t.uv0 = v2(a.w, b.w);
    t.uv1 = v2(c.w, d.w);
    t.uv2 = v2(e.w, f.w);
    t.gn = v3(g);
    t.hasDerivatives = g.w != 0.0f;
    t.dpdu = v3(h);
    t.dpdv = v3(i);
    return t;
}

// geometry.cl:177-215 computeSurfaceInteraction (wo / traceErrorOffset are set by the caller)
HD RTInteraction computeSurfaceInteraction(const Scene& sc, int shapeId, int primIdx, float bu, float bv) {
    RTInteraction si;
    if (PT_TRI_RECORDS && sc.triShade) {      // the same arithmetic, its triangle-only half read back from the per-triangle record
        const RTShape* shp = sc.shapes + shapeId;
        const TriShade t = loadTriShade(sc.triShade + 9 * (size_t)(shp->startIdx / 3u + (uint32_t)primIdx));
        const float w = 1.0f - bu - bv;
        si.p = t.p0 * w + t.p1 * bu + t.p2 * bv;
        si.uv = t.uv0 * w + t.uv1 * bu + t.uv2 * bv;
        si.gn = t.gn;
        si.sn = normalize(t.n0 * w + t.n1 * bu + t.n2 * bv);
        if (t.hasDerivatives) {
            si.dpdu = t.dpdu;
            si.dpdv = t.dpdv;
        } else {
            si.dpdu = normalize(computeOrthogonalVector(si.sn));
            si.dpdv = normalize(cross(si.sn, si.dpdu));
        }
        si.sdpdu = normalize(si.dpdu - dot(si.sn, si.dpdu) * si.sn);
        si.sdpdv = normalize(si.dpdv - dot(si.sn, si.dpdv) * si.sn - dot(si.sdpdu, si.dpdv) * si.sdpdu);
        si.shapeIdx = shapeId;
        si.wo = v3(0.f);
        si.traceErrorOffset = 0.f;
        return si;
    }
    const RTShape* shape = sc.shapes + shapeId;
    uint32_t i0, i1, i2;
    V3 p0, p1, p2;
    loadTrianglePositions(sc, shape, primIdx, &p0, &p1, &p2, &i0, &i1, &i2);
    const uint32_t sv = shape->startVertex;
    float2 t0 = __ldg(sc.uvs + sv + i0), t1 = __ldg(sc.uvs + sv + i1), t2 = __ldg(sc.uvs + sv + i2);
    V2 uv0 = v2(t0.x, t0.y), uv1 = v2(t1.x, t1.y), uv2 = v2(t2.x, t2.y);
    V3 n0 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i0)));
    V3 n1 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i1)));
    V3 n2 = transformVector3(shape->toWorldInvT, v3(__ldg(sc.normals + sv + i2)));
    const float w = 1.0f - bu - bv;
    si.p = p0 * w + p1 * bu + p2 * bv;
    si.uv = uv0 * w + uv1 * bu + uv2 * bv;
    si.gn = normalize(cross(p0 - p2, p1 - p2));
    si.sn = normalize(n0 * w + n1 * bu + n2 * bv);
    computeTrianglePartialDerivates(uv0, uv1, uv2, p0, p1, p2, si.sn, &si.dpdu, &si.dpdv);
    si.sdpdu = normalize(si.dpdu - dot(si.sn, si.dpdu) * si.sn);
    si.sdpdv = normalize(si.dpdv - dot(si.sn, si.dpdv) * si.sn - dot(si.sdpdu, si.dpdv) * si.sdpdu);
    si.shapeIdx = shapeId;
    si.wo = v3(0.f);
    si.traceErrorOffset = 0.f;
    return si;
}

// `i % n` of the reference (textures.cl:104-107) for the operands that occur: after wrapping, uv is in
// [0, 1], so floor(uv * n) is in [0, n] and the +1 neighbour in [1, n + 1].  One conditional subtract
// gives the same integer as the modulo without the integer-division sequence; anything outside
// [0, 2n) takes the general path.
static __device__ __noinline__ int wrap_index_general(int i, int n) { return i % n; }
HD int wrap_index(int i, int n) {
    if ((unsigned)i < (unsigned)n) return i;
    if ((unsigned)i < 2u * (unsigned)n) return i - n;
    return wrap_index_general(i, n);
}
// convert_float4(uchar4): exact for 0..255.  The magic-number form keeps the conversion on the FP32
// pipe instead of 16 I2F per fetch on the quarter-rate XU pipe; the value is identical.
HD float byte_to_float(unsigned b) { return __uint_as_float(0x4B000000u | b) - 8388608.0f; }
HD float4 texel_to_float4(uchar4 c) {
    return make_float4(byte_to_float(c.x), byte_to_float(c.y), byte_to_float(c.z), byte_to_float(c.w));
}

// textures.cl:70-125 readTexture2Df_linear (the only filter reached: textures.cl:204-209)
HD_TEX float4 readTexture2Df(const Scene& sc, int texId, V2 uv) {
    const TextureDesc2D tex = sc.textures2D[texId];
    const int tw = tex.width, th = tex.height;
    uv.x -= 1.0f / tw * 0.5f;
    uv.y -= 1.0f / th * 0.5f;
    switch (tex.wrap) {
        case TEX_WRAP_REPEAT:
            uv.x -= floorf(uv.x);
            uv.y -= floorf(uv.y);
            break;
        case TEX_WRAP_MIRRORED_REPEAT:
            if (uv.x > 1.0f || uv.x < 0.0f) uv.x = 1.0f - (uv.x - floorf(uv.x));
            if (uv.y > 1.0f || uv.y < 0.0f) uv.y = 1.0f - (uv.y - floorf(uv.y));
            break;
        case TEX_WRAP_CLAMP_TO_EDGE:
            uv = v2(clampf(uv.x, 0.0f, 1.0f), clampf(uv.y, 0.0f, 1.0f));
            break;
        case TEX_WRAP_CLAMP_TO_BORDER:
            if (uv.x > 1.0f || uv.x < 0.0f || uv.y > 1.0f || uv.y < 0.0f) return make_float4(0.f, 0.f, 0.f, 0.f);
            break;
    }
    const float fx = uv.x * tw, fy = uv.y * th;
    int x0 = wrap_index((int)floorf(fx), tw);
    int y0 = wrap_index((int)floorf(fy), th);
    int x1 = wrap_index(x0 + 1, tw);
    int y1 = wrap_index(y0 + 1, th);
    x0 = min(max(x0, 0), tw - 1);
    y0 = min(max(y0, 0), th - 1);
    x1 = min(max(x1, 0), tw - 1);
    y1 = min(max(y1, 0), th - 1);
    const float tx = fx - floorf(fx), ty = fy - floorf(fy);
    const uchar4* texD = (const uchar4*)(sc.texData2D + tex.memOffset);
    const uchar4 c00 = __ldg(texD + x0 + y0 * tw), c10 = __ldg(texD + x1 + y0 * tw);
    const uchar4 c01 = __ldg(texD + x0 + y1 * tw), c11 = __ldg(texD + x1 + y1 * tw);
    const float4 v00 = texel_to_float4(c00), v10 = texel_to_float4(c10);
    const float4 v01 = texel_to_float4(c01), v11 = texel_to_float4(c11);
    const float4 r = mix4(mix4(v00, v10, tx), mix4(v01, v11, tx), ty);
    const float k = 1.0f / 255.0f;
    return make_float4(r.x * k, r.y * k, r.z * k, r.w * k);
}

// materials.cl:14-30
HD void applyNormalMapping(const Scene& sc, int materialIdx, RTInteraction* si) {
    if (materialIdx == RT_INVALID_ID) return;
    const int texId = sc.materials[materialIdx].normalMapId;
    if (texId == RT_INVALID_ID) return;
    const float4 t = readTexture2Df(sc, texId, si->uv);
    const V3 nm = 2.0f * v3(t) - v3(1.0f);
    si->sn = normalize(si->sdpdu * nm.x + si->sdpdv * nm.y + si->sn * nm.z);
    si->sdpdu = normalize(cross(si->sn, si->sdpdv));
    si->sdpdv = normalize(cross(si->sdpdu, si->sn));
}

}  // namespace mcrt
