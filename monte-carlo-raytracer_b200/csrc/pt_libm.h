// pt_libm.h -- sinf / cosf / logf whose results are bit-identical to the host C library's (glibc 2.39, x86-64,
// the FMA variants its ifunc resolvers pick on every AVX2 machine), for host and device code.
//
// Why: with the identical per-pixel random stream, the only arithmetic in the shading path that CUDA and the strict
// host build of the reference's kernels (oracle/_ref) do NOT round identically are the C library's transcendentals
// (CUDA's sinf / cosf / logf are within 1-2 ulp of glibc's, not bit-equal).  A sampled direction that differs in its
// last bit is harmless for the radiance (1e-6 relative) until the next ray lands on the other side of a geometric
// edge or a magnified texel boundary -- then the path is a different sample (r01: 0.05 % of the Cornell pixels,
// 0.8 % of the textured 1080p frame).  Evaluating the same polynomials in the same order closes that gap: per-sample
// radiance becomes bit-identical to the oracle's, which turns the parity tests from "fraction within 1e-4" into
// equality.
//
// Algorithms: glibc's float sin / cos / log come from Arm's optimized-routines (MIT licence; glibc
// sysdeps/ieee754/flt-32/s_sincosf.h, s_sinf.c, s_cosf.c, e_logf.c): the argument is widened to double, reduced
// with one FMA against pi/2 (|x| < 120) or with a 192-bit integer multiply against 4/pi (larger), and a degree-7 / 8
// polynomial is evaluated in double and rounded once to float.  The operation order below -- in particular where
// the x86-64 FMA build contracts a multiply-add -- was read off the disassembly of the installed libm.so.6, the
// coefficient tables from its .rodata, and the result is checked against the installed sinf / cosf / logf for every
// float in tests/cpp/libm_check.cpp (CPU suite) and on the device through ptb_selftest_libm (GPU suite).
// Double-precision add / mul / fma are IEEE on both sides, so host and device agree by construction.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define MCRT_LIBM_HD __host__ __device__ __forceinline__
#else
#define MCRT_LIBM_HD inline
#endif

namespace mcrt {
namespace libm {

// 4/pi as 24 overlapping 32-bit words (s_sincosf_data.c: __inv_pio4)
#define MCRT_INV_PIO4                                                                                         \
    {0xa2u,       0xa2f9u,     0xa2f983u,   0xa2f9836eu, 0xf9836e4eu, 0x836e4e44u, 0x6e4e4415u, 0x4e441529u, \
     0x441529fcu, 0x1529fc27u, 0x29fc2757u, 0xfc2757d1u, 0x2757d1f5u, 0x57d1f534u, 0xd1f534ddu, 0xf534ddc0u, \
     0x34ddc0dbu, 0xddc0db62u, 0xc0db6295u, 0xdb629599u, 0x6295993cu, 0x95993c43u, 0x993c4390u, 0x3c439041u}
// e_logf_data.c: {1 / c, log(c)} for 16 subintervals of [sqrt(2)/2, sqrt(2))
#define MCRT_LOGF_TAB                                                                                 \
    {{0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},    \
     {0x1.49539f0f010b0p+0, -0x1.01eae7f513a67p-2}, {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},    \
     {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8ea0p+0, -0x1.1aa2bc79c8100p-3},    \
     {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},    \
     {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1.0000000000000p+0, 0x0.0p+0},                 \
     {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aa0p-1, 0x1.c5e53aa362eb4p-4},     \
     {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d224770p-3},     \
     {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}}

static const uint32_t kInvPio4Host[24] = MCRT_INV_PIO4;
static const double kLogfTabHost[16][2] = MCRT_LOGF_TAB;
#ifdef __CUDACC__
static __device__ __constant__ uint32_t kInvPio4Dev[24] = MCRT_INV_PIO4;
static __device__ __constant__ double kLogfTabDev[16][2] = MCRT_LOGF_TAB;
#endif
#ifdef __CUDA_ARCH__
#define MCRT_INV_PIO4_TAB kInvPio4Dev
#define MCRT_LOGF_TABLE kLogfTabDev
#else
#define MCRT_INV_PIO4_TAB kInvPio4Host
#define MCRT_LOGF_TABLE kLogfTabHost
#endif

MCRT_LIBM_HD uint32_t as_u32(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}
MCRT_LIBM_HD float as_f32(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// polynomial coefficients on [-pi/4, pi/4] (s_sincosf_data.c: __sincosf_table[0]); table[1] is the same with the
// cosine coefficients negated, i.e. the negated cosine polynomial -- negation commutes with every rounding
#define MCRT_C0 0x1p0
#define MCRT_C1 -0x1.ffffffd0c621cp-2
#define MCRT_C2 0x1.55553e1068f19p-5
#define MCRT_C3 -0x1.6c087e89a359dp-10
#define MCRT_C4 0x1.99343027bf8c3p-16
#define MCRT_S1 -0x1.555545995a603p-3
#define MCRT_S2 0x1.1107605230bc4p-7
#define MCRT_S3 -0x1.994eb3774cf24p-13
#define MCRT_HPI_INV 0x1.45F306DC9C883p+23   // 2 / pi * 2^24
#define MCRT_HPI 0x1.921FB54442D18p0         // pi / 2
#define MCRT_PI63 0x1.921FB54442D18p-62      // pi / 2^63

// sinf_poly (s_sincosf.h): sine of x for even n, cosine for odd n; x2 = x * x
MCRT_LIBM_HD double sin_poly(double x, double x2) {
    const double x3 = x * x2;
    const double s1 = fma(x2, MCRT_S3, MCRT_S2);
    const double x7 = x3 * x2;
    const double s = fma(x3, MCRT_S1, x);
    return fma(s1, x7, s);
}
MCRT_LIBM_HD double cos_poly(double x2) {
    const double x4 = x2 * x2;
    const double c1 = fma(x2, MCRT_C1, MCRT_C0);
    const double c2 = fma(x2, MCRT_C4, MCRT_C3);
    const double x6 = x2 * x4;
    const double c = fma(x4, MCRT_C2, c1);
    return fma(c2, x6, c);
}

// reduce_fast: x - n * pi/2 with n = round(x * 2/pi), for |x| < 120
MCRT_LIBM_HD double reduce_fast(double x, int* np) {
    const double r = x * MCRT_HPI_INV;
    const int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return fma(-(double)n, MCRT_HPI, x);
}
// reduce_large: |x| >= 120, 192-bit fixed-point product with 4/pi; returns the reduced argument in [-pi/4, pi/4]
MCRT_LIBM_HD double reduce_large(uint32_t xi, int* np) {
    const uint32_t* arr = &MCRT_INV_PIO4_TAB[(xi >> 26) & 15];
    const int shift = (xi >> 23) & 7;
    xi = (xi & 0xffffff) | 0x800000;
    xi <<= shift;
    uint64_t res0 = (uint64_t)(uint32_t)(xi * arr[0]);
    const uint64_t res1 = (uint64_t)xi * arr[4];
    const uint64_t res2 = (uint64_t)xi * arr[8];
    res0 = (res2 >> 32) | (res0 << 32);
    res0 += res1;
    const uint64_t n = (res0 + (1ULL << 61)) >> 62;
    res0 -= n << 62;
    *np = (int)n;
    return (double)(int64_t)res0 * MCRT_PI63;
}

// shared tail of sinf / cosf: `cosine` selects which of the two the caller wants
MCRT_LIBM_HD float sincos_core(float y, bool cosine) {
    const uint32_t iy = as_u32(y);
    const uint32_t top = (iy >> 20) & 0x7ff;
    double x = (double)y;
    if (top <= 0x3f3) {                    // |y| < pi/4
        const double x2 = x * x;
        if (top <= 0x397) return cosine ? 1.0f : y;      // |y| < 2^-12
        return (float)(cosine ? cos_poly(x2) : sin_poly(x, x2));
    }
    int n;
    int sign_index;
    if (top <= 0x42e) {                    // |y| < 120
        x = reduce_fast(x, &n);
        sign_index = n;
    } else if (top <= 0x7f7) {             // finite
        x = reduce_large(iy, &n);
        sign_index = n + (int)(iy >> 31);
    } else {
        return y - y;                      // inf / NaN -> NaN
    }
    const double s = ((sign_index & 3) == 1 || (sign_index & 3) == 2) ? -1.0 : 1.0;    // sign of sine per quadrant
    const bool negate_cos = (sign_index & 2) != 0;
    const bool odd = ((n & 1) != 0) != cosine;
    if (!odd) return (float)sin_poly(x * s, x * x);
    const double c = cos_poly(x * x);
    return (float)(negate_cos ? -c : c);
}

MCRT_LIBM_HD float sinf_glibc(float y) { return sincos_core(y, false); }
MCRT_LIBM_HD float cosf_glibc(float y) { return sincos_core(y, true); }

// e_logf.c
MCRT_LIBM_HD float logf_glibc(float xf) {
    uint32_t ix = as_u32(xf);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -INFINITY;                      // log(+-0) = -inf
        if (ix == 0x7f800000u) return xf;                        // log(inf) = inf
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return (xf - xf) / (xf - xf);   // negative or NaN
        ix = as_u32(xf * 0x1p23f);                               // subnormal: normalise
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) % 16;
    const int k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = MCRT_LOGF_TABLE[i][0], logc = MCRT_LOGF_TABLE[i][1];
    const double z = (double)as_f32(iz);
    const double r = fma(z, invc, -1.0);
    const double y0 = fma((double)k, 0x1.62e42fefa39efp-1, logc);
    const double r2 = r * r;
    double y = fma(r, 0x1.5575b0be00b6ap-2, -0x1.ffffef20a4123p-2);
    y = fma(r2, -0x1.00ea348b88334p-2, y);
    y = fma(y, r2, y0 + r);
    return (float)y;
}

}  // namespace libm
}  // namespace mcrt
