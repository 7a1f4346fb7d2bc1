// pt_math.cuh -- fp32 vector helpers, RNG and sampling routines of the shading stages.
//
// Restates (own code, same arithmetic) assets/kernels/math.cl, matrix.cl, rng.cl and the sampling
// half of samplers.cl.  Floating-point contract: compiled with --fmad=false
// --prec-div=true --prec-sqrt=true, so `a*b+c` is two roundings, dot() sums left to right,
// normalize(v) = v / sqrt(dot(v,v)) component-wise, mix(a,b,t) = a + (b-a)*t -- the definitions
// the strict host build of the reference's source uses (oracle/ref_build/clshim.hpp).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pt_libm.h"

namespace mcrt {

// sin / cos / log of the shading path: bit-identical to the host C library the oracle links (pt_libm.h), so that a
// sampled direction is the SAME float on both sides.  PT_LIBM_GLIBC = 0 falls back to CUDA's libm (1-2 ulp away).
#ifndef PT_LIBM_GLIBC
#define PT_LIBM_GLIBC 1
#endif
#if PT_LIBM_GLIBC
__device__ __forceinline__ float sin_ref(float x) { return libm::sinf_glibc(x); }
__device__ __forceinline__ float cos_ref(float x) { return libm::cosf_glibc(x); }
__device__ __forceinline__ float log_ref(float x) { return libm::logf_glibc(x); }
#else
__device__ __forceinline__ float sin_ref(float x) { return sinf(x); }
__device__ __forceinline__ float cos_ref(float x) { return cosf(x); }
__device__ __forceinline__ float log_ref(float x) { return logf(x); }
#endif

#define PT_PI 3.14159265359f
#define PT_PI_INV 0.31830988618f
#define PT_PI2_INV 0.15915494309f
#define PT_PI4_INV 0.07957747154f
#define PT_PI2 6.28318530718f
#define PT_PI_DIV_4 0.78539816339f
#define PT_PI_DIV_2 1.57079632679f

#define RT_TRACE_OFFSET 0.00001f          // kernel_data.h:10
#define RT_MAX_TRACE_DISTANCE 1000.0f     // kernel_data.h:12
#define RT_INVALID_ID (-1)

struct V2 {
    float x, y;
};
struct V3 {
    float x, y, z;
};

#define HD __device__ __forceinline__
// Large helpers with several call sites can be kept out of line to shrink the shade kernels (their
// SASS otherwise exceeds the instruction cache); results do not depend on it (no contraction).
#ifndef PT_TEX_NOINLINE
#define PT_TEX_NOINLINE 0
#endif
#ifndef PT_BXDF_NOINLINE
#define PT_BXDF_NOINLINE 0
#endif
#if PT_TEX_NOINLINE
#define HD_TEX static __device__ __noinline__
#else
#define HD_TEX __device__ __forceinline__
#endif
#if PT_BXDF_NOINLINE
#define HD_BXDF static __device__ __noinline__
#else
#define HD_BXDF __device__ __forceinline__
#endif

HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
HD V3 v3(float a) { return v3(a, a, a); }
HD V3 v3(const float4& f) { return v3(f.x, f.y, f.z); }
HD V2 v2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }

// IEEE a / b.  The hardware division sequence range-checks its operands (FCHK) and sends a ZERO numerator down a
// ~25-instruction subroutine; components of axis-aligned normals, tangents and colours are zero all the time
// (ncu r01: every warp of k_shade took that subroutine at 21 call sites, 10 % of its instructions).
// PT_FDIV_MODE 1 (default) -- branch-free shortcut: a zero numerator is replaced by 2^-40, whose quotient t cannot
// overflow for any finite b (>= 2^-149), and the result is a * t: a signed zero with sign(a) ^ sign(b) for finite
// non-zero b and for b = inf (t = 0), NaN for b = 0 (t = inf) and b = NaN -- exactly what 0 / b gives.  Non-zero
// numerators are untouched.
// PT_FDIV_MODE 2 -- the quotient formed in DOUBLE precision, a * (1 / b) with the reciprocal from MUFU.RCP64H + two
// Newton steps, rounded once to float.  It is exact for every operand (a quotient of two 24-bit numbers cannot lie
// within 2^-49, relatively, of a float midpoint; the double result is within 2^-51: tests/test_math_cpu.py, and
// ptb_selftest_fdiv on the device) and needs a third of the instructions (ncu r02: with --prec-div the `V3 / float`
// sites are 40 % of the warp instructions k_shade executes) -- and it is SLOWER on B200: the F2F conversions and
// DFMA / DMUL run on narrow pipes, k_shade 1.94 -> 2.76 ms per step (profiles/r02_shade.md).  Kept as a switch.
// PT_FDIV_MODE 0 -- plain `/`.
#ifndef PT_FDIV_MODE
#define PT_FDIV_MODE 1
#endif
HD double rcp_f64(float b) {           // 1 / b for finite non-zero b
    const double bd = (double)b;
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(bd));
    r = fma(fma(-bd, r, 1.0), r, r);
    r = fma(fma(-bd, r, 1.0), r, r);
    return r;
}
HD bool fdiv_plain_needed(float b) { return !(fabsf(b) > 0.0f && fabsf(b) < __int_as_float(0x7f800000)); }
HD float fdiv(float a, float b) {
#if PT_FDIV_MODE == 2
    if (fdiv_plain_needed(b)) return a / b;
    return (float)((double)a * rcp_f64(b));
#elif PT_FDIV_MODE == 1
    const bool zero = a == 0.0f;
    const float q = (zero ? 9.094947017729282e-13f : a) / b;
    const float z = a * q;
    return zero ? z : q;
#else
    return a / b;
#endif
}

HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
HD V3 operator/(V3 a, V3 b) { return v3(fdiv(a.x, b.x), fdiv(a.y, b.y), fdiv(a.z, b.z)); }
HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
HD V3 operator/(V3 a, float s) {
#if PT_FDIV_MODE == 2
    if (fdiv_plain_needed(s)) return v3(a.x / s, a.y / s, a.z / s);
    const double r = rcp_f64(s);
    return v3((float)((double)a.x * r), (float)((double)a.y * r), (float)((double)a.z * r));
#else
    return v3(fdiv(a.x, s), fdiv(a.y, s), fdiv(a.z, s));
#endif
}
HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
HD V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
HD V3& operator*=(V3& a, V3 b) { a = a * b; return a; }
HD V3& operator*=(V3& a, float s) { a = a * s; return a; }
HD V3& operator/=(V3& a, float s) { a = a / s; return a; }

HD V2 operator+(V2 a, V2 b) { return v2(a.x + b.x, a.y + b.y); }
HD V2 operator-(V2 a, V2 b) { return v2(a.x - b.x, a.y - b.y); }
HD V2 operator*(V2 a, float s) { return v2(a.x * s, a.y * s); }
HD V2 operator*(float s, V2 a) { return v2(s * a.x, s * a.y); }

HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
HD float length(V3 a) { return sqrtf(dot(a, a)); }
HD V3 normalize(V3 a) { return a / length(a); }
HD float distance(V3 a, V3 b) { return length(a - b); }
HD V3 mix(V3 a, V3 b, float t) { return a + (b - a) * t; }
HD float4 mix4(float4 a, float4 b, float t) {
    return make_float4(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t, a.w + (b.w - a.w) * t);
}
HD float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
HD float signf(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }
HD float absDot(V3 a, V3 b) { return fabsf(dot(a, b)); }
HD bool isNearZero(float v) { return fabsf(v) < 1e-8f; }
HD bool isNotNearZero(float v) { return fabsf(v) > 1e-8f; }
HD float distanceSquared(V3 a, V3 b) { return dot(a - b, a - b); }

// math.cl:61-75
HD V3 computeOrthogonalVector(V3 n) {
    if (fabsf(n.z) > 0.0f) {
        float d = sqrtf(n.z * n.z + n.x * n.x);
        return v3(-n.z / d, 0.0f, n.x / d);
    } else {
        float d = sqrtf(n.y * n.y + n.x * n.x);
        return v3(n.y / d, -n.x / d, 0.0f);
    }
}

// math.cl:88-91
HD V3 lerpDirection(V3 d0, V3 d1, V3 d2, V3 d3, float t0, float t1) {
    return normalize(mix(mix(d0, d1, t0), mix(d3, d2, t0), t1));
}

// matrix.cl:44-60; m points at 4 row-major float4 rows
HD V3 transformVector3(const float4* m, V3 v) {
    return v3(dot(v3(m[0]), v), dot(v3(m[1]), v), dot(v3(m[2]), v));
}
HD V3 transformPoint3(const float4* m, V3 v) {
    return v3(dot(v3(m[0]), v) + m[0].w, dot(v3(m[1]), v) + m[1].w, dot(v3(m[2]), v) + m[2].w);
}
HD float4 transformVector4(const float4* m, float4 v) {
    float4 r;
    r.x = m[0].x * v.x + m[0].y * v.y + m[0].z * v.z + m[0].w * v.w;
    r.y = m[1].x * v.x + m[1].y * v.y + m[1].z * v.z + m[1].w * v.w;
    r.z = m[2].x * v.x + m[2].y * v.y + m[2].z * v.z + m[2].w * v.w;
    r.w = m[3].x * v.x + m[3].y * v.y + m[3].z * v.z + m[3].w * v.w;
    return r;
}

// ------------------------------------------------------------------ RNG (rng.cl:48-66,104-107)
HD uint32_t wang_hash(uint32_t seed) {
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}
HD uint32_t randUInt(uint32_t* seed) {
    *seed += 2463534242u;
    *seed ^= (*seed << 13);
    *seed ^= (*seed >> 17);
    *seed ^= (*seed << 5);
    return *seed;
}
// (float)u / 0xffffffff: the divisor converts to 2^32 in fp32, so 1.0f can be returned (E6)
HD float randFloat(uint32_t* seed) { return ((float)randUInt(seed)) / 4294967296.0f; }

// Sampler (samplers.cl:19-25,75-128).  The reference picks the variant at compile time
// (RT_SAMPLER, random by default); here it is a run-time choice: `sobol` == NULL selects the random
// variant.  All index arithmetic wraps in 32 bits (E14); `bufferIdx`, `width` and `height` are the
// GLOBAL pixel index and image size on every rank.
#define SOBOL_MATRIX_SIZE 52
struct Sampler {
    uint32_t idx;
    uint32_t dimension;
    uint32_t scramble;
    const uint32_t* sobol;
};
// random:  idx = wang_hash(bufferIdx + (frame+1) * W * H * (offset+1))                       (:83-84)
// Sobol:   idx = bufferIdx + frame * W * H, dimension 0, scramble = xorshift(wang_hash((frame+1)*(offset+1)))  (:75-80)
HD Sampler makeSampler(const uint32_t* sobol, int bufferIdx, int frame, int width, int height, int offset) {
    Sampler s;
    s.sobol = sobol;
    s.dimension = 0;
    s.scramble = 0;
    if (sobol) {
        s.idx = (uint32_t)bufferIdx + (uint32_t)frame * (uint32_t)width * (uint32_t)height;
        uint32_t seed = wang_hash((uint32_t)(frame + 1) * (uint32_t)(offset + 1));
        s.scramble = randUInt(&seed);
    } else {
        s.idx = wang_hash((uint32_t)bufferIdx + (uint32_t)(frame + 1) * (uint32_t)width * (uint32_t)height * (uint32_t)(offset + 1));
    }
    return s;
}
// samplers.cl:64-72: v * 0x1p-32f
HD float sobolSample(uint32_t idx, uint32_t dimension, uint32_t scramble, const uint32_t* __restrict__ m) {
    uint32_t v = scramble;
    for (uint32_t i = dimension * SOBOL_MATRIX_SIZE; idx != 0; idx >>= 1, ++i)
        if (idx & 1u) v ^= __ldg(m + i);
    return (float)v * 2.3283064365386963e-10f;
}
HD float getSample1D(Sampler* s) {
    if (s->sobol) return sobolSample(s->idx, s->dimension++, s->scramble, s->sobol);
    return randFloat(&s->idx);
}
HD V2 getSample2D(Sampler* s) {
    float a = getSample1D(s);
    float b = getSample1D(s);
    return v2(a, b);
}
// advance past n one-dimensional draws without using them
HD void skipSamples(Sampler* s, int n) {
    if (s->sobol) {
        s->dimension += (uint32_t)n;
    } else {
        for (int i = 0; i < n; ++i) randUInt(&s->idx);
    }
}

// ------------------------------------------------------------------ sampling (samplers.cl:129-285)
HD V3 uniformSampleSphere(V2 u) {
    float y = 1.0f - 2.0f * u.x;
    float r = sqrtf(fmaxf(0.0f, 1.0f - y * y));
    float phi = 2.0f * 3.14159265359f * u.y;
    return v3(r * cos_ref(phi), y, r * sin_ref(phi));
}

HD V2 concentricSampleDisc(V2 u) {
    V2 uOffset = 2.0f * u - v2(1.0f, 1.0f);
    if (u.x < 1e-8f && u.y < 1e-8f) return v2(0.0f, 0.0f);
    float theta, r;
    if (fabsf(uOffset.x) > fabsf(uOffset.y)) {
        r = uOffset.x;
        theta = PT_PI_DIV_4 * (uOffset.y / uOffset.x);
    } else {
        r = uOffset.y;
        theta = PT_PI_DIV_2 - PT_PI_DIV_4 * (uOffset.x / uOffset.y);
    }
    return r * v2(cos_ref(theta), sin_ref(theta));
}

HD V3 cosineSampleHemisphere(V2 u) {
    V2 d = concentricSampleDisc(u);
    float y = sqrtf(fmaxf(0.0f, 1.0f - d.x * d.x - d.y * d.y));
    return v3(d.x, y, d.y);
}
HD float cosineHemispherePdf(float cosTheta) { return cosTheta * PT_PI_INV; }

HD V2 uniformSampleTriangle(V2 u) {
    float su0 = sqrtf(u.x);
    return v2(1.0f - su0, u.y * su0);
}

// samplers.cl:259-269 sampleDisk -> point and normal on the disk, pdf = 1/area
HD void sampleDisk(V3 p, V3 n, float radius, V2 u, V3* outP, V3* outN, float* pdf) {
    V2 p2d = concentricSampleDisc(u);
    *outN = n;
    V3 t = computeOrthogonalVector(n);
    V3 b = normalize(cross(n, t));
    *outP = p + t * p2d.x * radius + b * p2d.y * radius;
    *pdf = 1.0f / (3.14159265359f * radius * radius);
}

// samplers.cl:275-285 sampleTriangle
HD void sampleTriangle(V3 p0, V3 p1, V3 p2, V2 u, V3* outP, V3* outN, float* pdf) {
    V2 b = uniformSampleTriangle(u);
    *outP = b.x * p0 + b.y * p1 + (1 - b.x - b.y) * p2;
    V3 c = cross(p1 - p0, p2 - p0);
    *outN = normalize(c);
    float area = length(c) * 0.5f;
    *pdf = 1.0f / area;
}

}  // namespace mcrt
