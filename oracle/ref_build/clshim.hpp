// clshim.hpp -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
//
// A minimal OpenCL-C 1.2 vocabulary for g++ so that the reference's own kernel
// sources (/root/reference/assets/kernels/*.cl and Radeon Rays'
// src/kernels/CL/{common,intersect_bvh2_lds}.cl) compile as plain C++ and run one
// work-item per function call.  Nothing of the reference is copied here: the
// sources are read from /root/reference at build time, pushed through the
// mechanical rewrite in cl2cpp.py (vector literals `(float3)(..)` ->
// `float3{..}`, evaluated left to right, the `__cplusplus` branch of kernel_data.h switched off) and
// only ever written to oracle/_ref/ (git-ignored).
//
// Floating-point contract of this build: strict IEEE fp32, no
// contraction (-ffp-contract=off), native_recip(x) := 1.0f/x, mad(a,b,c) := a*b+c
// rounded twice, dot() summed left to right, normalize(v) := v / sqrt(dot(v,v)),
// mix(a,b,t) := a + (b-a)*t, clamp(x,lo,hi) := fmin(fmax(x,lo),hi).
// Uninitialised automatic vector variables are zero.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stddef.h>

typedef unsigned int uint;
typedef unsigned short ushort;
typedef unsigned char uchar;

#define __kernel
#define __global
#define __local static thread_local
#define __constant const
#define __private
#define restrict __restrict__
#define APPLE 1 /* common.cl: skip its own make_floatN helpers, the shim has them */

// ---------------------------------------------------------------- work-item ids
namespace clshim {
extern thread_local size_t g_gid[3];
extern thread_local size_t g_lid[3];
}
static inline size_t get_global_id(int d) { return clshim::g_gid[d]; }
static inline size_t get_local_id(int d) { return clshim::g_lid[d]; }

// ---------------------------------------------------------------- vector types
struct float2;
struct float3;
struct float4;

// Swizzle proxies live inside the unions of the vector types and alias their storage.
template <typename V2, int A, int B, int N> struct Swz2 {
    float v[N];
    operator V2() const { return V2(v[A], v[B]); }
    Swz2& operator=(const V2& o) { float a = o.x, b = o.y; v[A] = a; v[B] = b; return *this; }
};
template <typename V3, int A, int B, int C> struct Swz3 {
    float v[4];
    operator V3() const { return V3(v[A], v[B], v[C]); }
    Swz3& operator=(const V3& o) { float a = o.x, b = o.y, c = o.z; v[A] = a; v[B] = b; v[C] = c; return *this; }
    Swz3& operator=(const Swz3& o) { return *this = (V3)o; }
    Swz3& operator+=(const V3& o) { v[A] += o.x; v[B] += o.y; v[C] += o.z; return *this; }
    Swz3& operator-=(const V3& o) { v[A] -= o.x; v[B] -= o.y; v[C] -= o.z; return *this; }
    Swz3& operator*=(const V3& o) { v[A] *= o.x; v[B] *= o.y; v[C] *= o.z; return *this; }
    Swz3& operator*=(float s) { v[A] *= s; v[B] *= s; v[C] *= s; return *this; }
    Swz3& operator/=(float s) { v[A] /= s; v[B] /= s; v[C] /= s; return *this; }
    V3 operator-() const { return V3(-v[A], -v[B], -v[C]); }
};

struct alignas(8) float2 {
    union {
        struct { float x, y; };
        float s[2];
        Swz2<float2, 0, 1, 2> xy;
    };
    float2() : x(0.f), y(0.f) {}
    template <typename A> explicit float2(A a) : x((float)a), y((float)a) {}
    template <typename A, typename B> float2(A a, B b) : x((float)a), y((float)b) {}
    float2(const float2& o) : x(o.x), y(o.y) {}
    float2& operator=(const float2& o) { x = o.x; y = o.y; return *this; }
    float2& operator=(float a) { x = a; y = a; return *this; } /* OpenCL scalar widening (matrix.cl:35) */
};

struct alignas(16) float3 {
    union {
        struct { float x, y, z, _w; };
        float s[4];
        Swz2<float2, 0, 1, 4> xy;
        Swz2<float2, 1, 2, 4> yz;
        Swz2<float2, 0, 2, 4> xz;
        Swz3<float3, 0, 1, 2> xyz;
    };
    float3() : x(0.f), y(0.f), z(0.f), _w(0.f) {}
    explicit float3(float a) : x(a), y(a), z(a), _w(0.f) {}
    explicit float3(double a) : x((float)a), y((float)a), z((float)a), _w(0.f) {}
    explicit float3(int a) : x((float)a), y((float)a), z((float)a), _w(0.f) {}
    template <typename A, typename B, typename C> float3(A a, B b, C c) : x((float)a), y((float)b), z((float)c), _w(0.f) {}
    float3(const float3& o) : x(o.x), y(o.y), z(o.z), _w(0.f) {}
    float3& operator=(const float3& o) { x = o.x; y = o.y; z = o.z; return *this; }
};

struct alignas(16) float4 {
    union {
        struct { float x, y, z, w; };
        float s[4];
        Swz2<float2, 0, 1, 4> xy;
        Swz3<float3, 0, 1, 2> xyz;
    };
    float4() : x(0.f), y(0.f), z(0.f), w(0.f) {}
    explicit float4(float a) : x(a), y(a), z(a), w(a) {}
    template <typename A, typename B, typename C, typename D> float4(A a, B b, C c, D d) : x((float)a), y((float)b), z((float)c), w((float)d) {}
    float4(const float3& v, float d);
    float4(const float4& o) : x(o.x), y(o.y), z(o.z), w(o.w) {}
    float4& operator=(const float4& o) { x = o.x; y = o.y; z = o.z; w = o.w; return *this; }
};

struct alignas(8) int2 {
    int x, y;
    int2() : x(0), y(0) {}
    template <typename A> explicit int2(A a) : x((int)a), y((int)a) {}
    template <typename A, typename B> int2(A a, B b) : x((int)a), y((int)b) {}
};
struct alignas(8) uint2 {
    uint x, y;
    uint2() : x(0), y(0) {}
    template <typename A, typename B> uint2(A a, B b) : x((uint)a), y((uint)b) {}
};
struct alignas(16) int3 {
    int x, y, z, _w;
    int3() : x(0), y(0), z(0), _w(0) {}
    template <typename A, typename B, typename C> int3(A a, B b, C c) : x((int)a), y((int)b), z((int)c), _w(0) {}
};
struct alignas(4) uchar4 {
    uchar x, y, z, w;
};

inline float4::float4(const float3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}

static_assert(sizeof(float2) == 8 && sizeof(float3) == 16 && sizeof(float4) == 16, "OpenCL vector sizes");
static_assert(sizeof(int2) == 8 && sizeof(uchar4) == 4, "OpenCL vector sizes");

// ---------------------------------------------------------------- literals
static inline float2 make_float2(float a) { return float2(a, a); }
static inline float2 make_float2(float a, float b) { return float2(a, b); }
static inline float3 make_float3(float a) { return float3(a, a, a); }
static inline float3 make_float3(float a, float b, float c) { return float3(a, b, c); }
static inline float4 make_float4(float a) { return float4(a, a, a, a); }
static inline float4 make_float4(float a, float b, float c, float d) { return float4(a, b, c, d); }
static inline float4 make_float4(const float3& v, float d) { return float4(v.x, v.y, v.z, d); }
static inline int2 make_int2(long long a) { return int2((int)a, (int)a); }
static inline int2 make_int2(long long a, long long b) { return int2((int)a, (int)b); }
static inline uint2 make_uint2(uint a, uint b) { return uint2(a, b); }
static inline int3 make_int3(int a, int b, int c) { return int3(a, b, c); }

// ---------------------------------------------------------------- operators
#define CLSHIM_BINOP2(op)                                                                         \
    static inline float2 operator op(const float2& a, const float2& b) { return float2(a.x op b.x, a.y op b.y); } \
    static inline float2 operator op(const float2& a, float b) { return float2(a.x op b, a.y op b); }             \
    static inline float2 operator op(float a, const float2& b) { return float2(a op b.x, a op b.y); }
#define CLSHIM_BINOP3(op)                                                                         \
    static inline float3 operator op(const float3& a, const float3& b) { return float3(a.x op b.x, a.y op b.y, a.z op b.z); } \
    static inline float3 operator op(const float3& a, float b) { return float3(a.x op b, a.y op b, a.z op b); }               \
    static inline float3 operator op(float a, const float3& b) { return float3(a op b.x, a op b.y, a op b.z); }
#define CLSHIM_BINOP4(op)                                                                         \
    static inline float4 operator op(const float4& a, const float4& b) { return float4(a.x op b.x, a.y op b.y, a.z op b.z, a.w op b.w); } \
    static inline float4 operator op(const float4& a, float b) { return float4(a.x op b, a.y op b, a.z op b, a.w op b); }                 \
    static inline float4 operator op(float a, const float4& b) { return float4(a op b.x, a op b.y, a op b.z, a op b.w); }
CLSHIM_BINOP2(+) CLSHIM_BINOP2(-) CLSHIM_BINOP2(*) CLSHIM_BINOP2(/)
CLSHIM_BINOP3(+) CLSHIM_BINOP3(-) CLSHIM_BINOP3(*) CLSHIM_BINOP3(/)
CLSHIM_BINOP4(+) CLSHIM_BINOP4(-) CLSHIM_BINOP4(*) CLSHIM_BINOP4(/)

// int / double scalars mixed with vectors (e.g. `2 * PI`-style literals, `1 - b.x`)
static inline float3 operator*(const float3& a, int b) { return a * (float)b; }
static inline float3 operator*(int a, const float3& b) { return (float)a * b; }
static inline float3 operator*(const float3& a, double b) { return a * (float)b; }
static inline float3 operator*(double a, const float3& b) { return (float)a * b; }
static inline float3 operator/(const float3& a, int b) { return a / (float)b; }

static inline float2 operator-(const float2& a) { return float2(-a.x, -a.y); }
static inline float3 operator-(const float3& a) { return float3(-a.x, -a.y, -a.z); }
static inline float4 operator-(const float4& a) { return float4(-a.x, -a.y, -a.z, -a.w); }

#define CLSHIM_ASSIGNOP(T, op)                                                         \
    static inline T& operator op##=(T& a, const T& b) { a = a op b; return a; }        \
    static inline T& operator op##=(T& a, float b) { a = a op b; return a; }
CLSHIM_ASSIGNOP(float2, +) CLSHIM_ASSIGNOP(float2, -) CLSHIM_ASSIGNOP(float2, *) CLSHIM_ASSIGNOP(float2, /)
CLSHIM_ASSIGNOP(float3, +) CLSHIM_ASSIGNOP(float3, -) CLSHIM_ASSIGNOP(float3, *) CLSHIM_ASSIGNOP(float3, /)
CLSHIM_ASSIGNOP(float4, +) CLSHIM_ASSIGNOP(float4, -) CLSHIM_ASSIGNOP(float4, *) CLSHIM_ASSIGNOP(float4, /)

// ---------------------------------------------------------------- scalar builtins
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline uint min(uint a, uint b) { return a < b ? a : b; }
static inline uint max(uint a, uint b) { return a > b ? a : b; }
static inline float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static inline int clamp(int x, int lo, int hi) { return min(max(x, lo), hi); }
static inline float mix(float a, float b, float t) { return a + (b - a) * t; }
static inline float mad(float a, float b, float c) { return a * b + c; }
static inline float sign(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }
static inline float native_recip(float x) { return 1.0f / x; }
static inline float pow(float a, int b) { return powf(a, (float)b); }
static inline uint as_uint(float f) { uint u; memcpy(&u, &f, 4); return u; }
static inline int convert_int(uint u) { return (int)u; }
static inline uint clz(uint v) { return v == 0 ? 32u : (uint)__builtin_clz(v); }

// ---------------------------------------------------------------- vector builtins
static inline float dot(const float2& a, const float2& b) { return a.x * b.x + a.y * b.y; }
static inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
static inline float3 cross(const float3& a, const float3& b) {
    return float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(const float3& a) { return sqrtf(dot(a, a)); }
static inline float length(const float2& a) { return sqrtf(dot(a, a)); }
static inline float distance(const float3& a, const float3& b) { return length(a - b); }
static inline float3 normalize(const float3& a) { return a / length(a); }
static inline float2 normalize(const float2& a) { return a / length(a); }
static inline float3 mix(const float3& a, const float3& b, float t) { return a + (b - a) * t; }
static inline float4 mix(const float4& a, const float4& b, float t) { return a + (b - a) * t; }
static inline float3 mad(const float3& a, const float3& b, const float3& c) { return a * b + c; }
static inline float3 min(const float3& a, const float3& b) { return float3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline float3 max(const float3& a, const float3& b) { return float3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline float4 min(const float4& a, const float4& b) { return float4(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), fminf(a.w, b.w)); }
static inline float4 max(const float4& a, const float4& b) { return float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w)); }
static inline float4 clamp(const float4& a, float lo, float hi) { return float4(clamp(a.x, lo, hi), clamp(a.y, lo, hi), clamp(a.z, lo, hi), clamp(a.w, lo, hi)); }
static inline float3 clamp(const float3& a, float lo, float hi) { return float3(clamp(a.x, lo, hi), clamp(a.y, lo, hi), clamp(a.z, lo, hi)); }
static inline float2 floor(const float2& a) { return float2(floorf(a.x), floorf(a.y)); }
static inline float3 sqrt(const float3& a) { return float3(sqrtf(a.x), sqrtf(a.y), sqrtf(a.z)); }
static inline float3 fabs(const float3& a) { return float3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
static inline float4 convert_float4(const uchar4& c) { return float4((float)c.x, (float)c.y, (float)c.z, (float)c.w); }

// ---------------------------------------------------------------- atomics (BDPT splat)
static inline uint atomic_cmpxchg(volatile uint* p, uint cmp, uint val) {
    __atomic_compare_exchange_n((uint*)p, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;
}

// component-wise log / exp on float4 (ToneMapping.cl's log-average kernels)
static inline float4 log(float4 v) { return float4{logf(v.x), logf(v.y), logf(v.z), logf(v.w)}; }
static inline float4 exp(float4 v) { return float4{expf(v.x), expf(v.y), expf(v.z), expf(v.w)}; }

// ---------------------------------------------------------------- images
// The reference's post-processing kernels (ToneMapping.cl, Denoise.cl) read and write GL-interop
// image2d_t objects.  Here an image is a plain RGBA32F buffer; the only sampler those kernels use is
// non-normalised / nearest / clamp-to-edge (image_samplers.cl).
struct clshim_image {
    float4* data;
    int width, height;
};
typedef clshim_image* image2d_t;
typedef int sampler_t;
#define write_only
#define read_only
#define CLK_NORMALIZED_COORDS_FALSE 0
#define CLK_ADDRESS_CLAMP_TO_EDGE 0
#define CLK_ADDRESS_CLAMP 0
#define CLK_FILTER_NEAREST 0
static inline int get_image_width(image2d_t im) { return im->width; }
static inline int get_image_height(image2d_t im) { return im->height; }
static inline void write_imagef(image2d_t im, int2 p, float4 v) {
    if (im && p.x >= 0 && p.y >= 0 && p.x < im->width && p.y < im->height) im->data[p.y * im->width + p.x] = v;
}
static inline float4 read_imagef(image2d_t im, sampler_t, int2 p) {
    int x = p.x < 0 ? 0 : (p.x >= im->width ? im->width - 1 : p.x);
    int y = p.y < 0 ? 0 : (p.y >= im->height ? im->height - 1 : p.y);
    return im->data[y * im->width + x];
}
static inline float4 read_imagef(image2d_t im, int2 p) { return read_imagef(im, 0, p); }
