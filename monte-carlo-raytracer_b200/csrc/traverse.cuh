// traverse.cuh -- closest-hit / any-hit traversal of the compressed 8-wide BVH (bvh8.h) for sm_100a.
//
// Takes the role of Radeon Rays' intersect_main / occluded_main
// (RadeonRays/src/kernels/CL/intersect_bvh2_lds.cl:66-226,229-363).  The tree and the walk are
// different (8-wide quantised nodes, octant-ordered bit-mask stack entries, persistent warps with
// dynamic ray fetch); the *triangle test* is the reference's Moeller-Trumbore formulation,
// operation for operation (common.cl:178-218), evaluated in fp32 without contraction, so the
// accepted set and every t are bit-identical to the reference's strict build:
//   accept iff denom != 0, 0 <= b1 <= 1, b2 >= 0, b1 + b2 <= 1, 0 <= t <= closest, then t < closest.
// Ties (equal fp32 t) are broken deterministically by the smallest (shape id, prim id) instead of
// the reference's tree-order "first found" (SURVEY Appendix A, E1).
// Barycentrics are recomputed from the hit point with the reference's dot-product method
// (common.cl:250-277).  Ray mask rule: a leaf triangle is skipped iff ray.extra.x == shape id
// (intersect_bvh2_lds.cl:164, RR_RAY_MASK on).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "async_copy.cuh"

namespace mcrt {

#define MCRT_STACK_SIZE 40
#ifndef MCRT_DYNAMIC_FETCH_THRESHOLD
#define MCRT_DYNAMIC_FETCH_THRESHOLD 24
#endif
// Triangle postponing (Ylitie et al. 2017, section 4.4): a lane that reaches its leaf triangles while fewer than
// MCRT_POSTPONE_PERCENT % of the warp's traversing lanes do, pushes the triangle group back and goes on with
// nodes, so that triangle tests run with more lanes converged.  0 = off.  Measured on the 1M-triangle knot:
// incoherent closest hit 2.49 -> 3.05 Grays/s, any hit 3.48 -> 4.11 Grays/s.  Results do not depend on the order
// (closest hit has a deterministic tie rule, any hit reports a boolean).
#ifndef MCRT_POSTPONE_PERCENT
#define MCRT_POSTPONE_PERCENT 35      // profiles/r01_traversal_variants.md
#endif

// Latency hints (experiment knob, off).  Bit 0: while a triangle is tested, the next one of the group is prefetched.
// Bit 1: as soon as a node's hit mask is known, the child that will be visited first is prefetched.  Measured
// (profiles/r01_c5_tiles.md): neither shortens the tail of a launch, and the node prefetch costs 20 % of the mixed
// trace on C3 (issue slots), so both stay off.
#ifndef MCRT_TRAV_PREFETCH
#define MCRT_TRAV_PREFETCH 0
#endif
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

struct Ray {
    float ox, oy, oz, tmax;
    float dx, dy, dz;
    int mask;
};

struct Hit {
    float t;
    int tri;  // absolute triangle index in the triangle array, -1 = miss
};

__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x) {
    uint32_t r;
    asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(r) : "r"(x));
    return r;
}

__device__ __forceinline__ uint32_t extract_byte(uint32_t x, uint32_t i) { return (x >> (i * 8)) & 0xffu; }

// Quantised plane coordinate -> float without the quarter-rate I2F of the XU pipe (ncu r01: 48 I2F per node visit,
// XU the busiest pipe of the traversal kernels): one PRMT drops byte J of x into mantissa bits [15:8] of the
// constant c = 1.0f, giving exactly 1 + b * 2^-15; the node's scale and offset absorb the affine map
// (aix' = aix * 2^15, orx' = orx - aix').  Rounding orx' costs at most 2^-24 * 2^15 = 2^-9 quantisation step of
// accuracy per plane on top of the errors proportional to t that the relative widening of the far side already
// covers.  It is paid back twice over: the bytes of a node's HIGH planes are converted with c = 1.0f + 4 ulp, i.e.
// 2^-6 step further out, and the node offset is lowered by 2^-7 step (Traverser::run), so that LOW planes end up
// 2^-7 step below and HIGH planes 2^-7 step above what the node stores: the slab interval only widens, on the near
// side as well as on the far side, whatever the sign of the direction.
#ifndef MCRT_QUANT_PRMT
#define MCRT_QUANT_PRMT 1
#endif
#define MCRT_QUANT_C_LO 0x3F800000u
#define MCRT_QUANT_C_HI 0x3F800004u
template <int J>
__device__ __forceinline__ float quant_byte_to_float(uint32_t x, uint32_t c) {
#if MCRT_QUANT_PRMT
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(c), "n"(0x7604 | (J << 4)));
    return __uint_as_float(r);
#else
    return (float)extract_byte(x, J);
#endif
}

__device__ __forceinline__ float safe_inv(float d) {
    return __fdiv_rn(1.0f, fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d));
}

// fp32 helpers that the compiler may not contract into FMAs
__device__ __forceinline__ float mul_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float dot3_(float ax, float ay, float az, float bx, float by, float bz) {
    return add_(add_(mul_(ax, bx), mul_(ay, by)), mul_(az, bz));
}

// common.cl:178-218 fast_intersect_triangle; returns true and t when accepted against t_max.
__device__ __forceinline__ bool intersect_triangle_ref(const Ray& r, const float4 v1, const float4 v2, const float4 v3,
                                                       float t_max, float& t_out) {
    const float e1x = sub_(v2.x, v1.x), e1y = sub_(v2.y, v1.y), e1z = sub_(v2.z, v1.z);
    const float e2x = sub_(v3.x, v1.x), e2y = sub_(v3.y, v1.y), e2z = sub_(v3.z, v1.z);
    // s1 = cross(d, e2)
    const float s1x = sub_(mul_(r.dy, e2z), mul_(r.dz, e2y));
    const float s1y = sub_(mul_(r.dz, e2x), mul_(r.dx, e2z));
    const float s1z = sub_(mul_(r.dx, e2y), mul_(r.dy, e2x));
    const float denom = dot3_(s1x, s1y, s1z, e1x, e1y, e1z);
    if (denom == 0.f) return false;
    const float invd = __fdiv_rn(1.0f, denom);
    const float dx = sub_(r.ox, v1.x), dy = sub_(r.oy, v1.y), dz = sub_(r.oz, v1.z);
    const float b1 = mul_(dot3_(dx, dy, dz, s1x, s1y, s1z), invd);
    // s2 = cross(d, e1)
    const float s2x = sub_(mul_(dy, e1z), mul_(dz, e1y));
    const float s2y = sub_(mul_(dz, e1x), mul_(dx, e1z));
    const float s2z = sub_(mul_(dx, e1y), mul_(dy, e1x));
    const float b2 = mul_(dot3_(r.dx, r.dy, r.dz, s2x, s2y, s2z), invd);
    const float temp = mul_(dot3_(e2x, e2y, e2z, s2x, s2y, s2z), invd);
    if (b1 < 0.f || b1 > 1.f || b2 < 0.f || add_(b1, b2) > 1.f || temp < 0.f || temp > t_max) return false;
    t_out = temp;
    return true;
}

// common.cl:250-277 triangle_calculate_barycentrics
__device__ __forceinline__ float2 barycentrics_ref(float px, float py, float pz, const float4 v1, const float4 v2,
                                                   const float4 v3) {
    const float e1x = sub_(v2.x, v1.x), e1y = sub_(v2.y, v1.y), e1z = sub_(v2.z, v1.z);
    const float e2x = sub_(v3.x, v1.x), e2y = sub_(v3.y, v1.y), e2z = sub_(v3.z, v1.z);
    const float ex = sub_(px, v1.x), ey = sub_(py, v1.y), ez = sub_(pz, v1.z);
    const float d00 = dot3_(e1x, e1y, e1z, e1x, e1y, e1z);
    const float d01 = dot3_(e1x, e1y, e1z, e2x, e2y, e2z);
    const float d11 = dot3_(e2x, e2y, e2z, e2x, e2y, e2z);
    const float d20 = dot3_(ex, ey, ez, e1x, e1y, e1z);
    const float d21 = dot3_(ex, ey, ez, e2x, e2y, e2z);
    const float denom = sub_(mul_(d00, d11), mul_(d01, d01));
    if (denom == 0.f) return make_float2(0.f, 0.f);
    const float inv = __fdiv_rn(1.0f, denom);
    const float b1 = mul_(sub_(mul_(d11, d20), mul_(d01, d21)), inv);
    const float b2 = mul_(sub_(mul_(d00, d21), mul_(d01, d20)), inv);
    return make_float2(b1, b2);
}

// Per-ray traversal state, kept in registers across dynamic-fetch rounds.
struct TravState {
    uint2 ngroup;   // x: child node base index, y: [31:24] pending internal children, [7:0] imask
    uint2 tgroup;   // x: triangle base index,   y: [23:0] pending triangles
    int sp;
    float idx, idy, idz;
    uint32_t octinv4;
    float closest;
    int best_tri;
    int best_sid, best_pid;
    bool any_hit;   // read only by Traverser<*, *, MIXED = true>: this ray wants any hit, not the closest
    // two-level traversal only (Traverser<*, *, *, TWO_LEVEL = true>): the ray in the space of the tree being walked
    // (world at the top level, the instance's object space below), the stack depth at which the instance was entered
    // (-1 = at the top level), the instance being walked and the one that owns the best hit
    float cox, coy, coz, cdx, cdy, cdz;
    int inst_sp, cur_inst, cur_sid, best_inst;
    // run<SPLIT = true> only (the cooperative tail of persistent_trace): this lane walks the ray restricted to
    // tlo < t <= closest; tlo_box is tlo rounded down for the slab test.  Not touched by init().
    float tlo, tlo_box;
};

// ray.o / ray.d through rows 0..2 of a world -> object matrix: transform_point / transform_vector of the reference's
// two-level kernel (intersect_bvh2level_skiplinks.cl:103-131), same operation order; the direction is NOT
// re-normalised, so t means the same in both spaces
__device__ __forceinline__ void to_object_space(const float4 m0, const float4 m1, const float4 m2, const Ray& r, float& ox,
                                                float& oy, float& oz, float& dx, float& dy, float& dz) {
    ox = add_(add_(add_(mul_(m0.x, r.ox), mul_(m0.y, r.oy)), mul_(m0.z, r.oz)), m0.w);
    oy = add_(add_(add_(mul_(m1.x, r.ox), mul_(m1.y, r.oy)), mul_(m1.z, r.oz)), m1.w);
    oz = add_(add_(add_(mul_(m2.x, r.ox), mul_(m2.y, r.oy)), mul_(m2.z, r.oz)), m2.w);
    dx = add_(add_(mul_(m0.x, r.dx), mul_(m0.y, r.dy)), mul_(m0.z, r.dz));
    dy = add_(add_(mul_(m1.x, r.dx), mul_(m1.y, r.dy)), mul_(m1.z, r.dz));
    dz = add_(add_(mul_(m2.x, r.dx), mul_(m2.y, r.dy)), mul_(m2.z, r.dz));
}
// the ray the best hit was found with: the world ray, or that ray in the object space of the hit's instance
__device__ __forceinline__ Ray hit_space_ray(const float4* __restrict__ instances, const Ray& r, int best_inst) {
    if (!instances || best_inst < 0) return r;
    Ray o = r;
    to_object_space(__ldg(instances + 4 * (size_t)best_inst), __ldg(instances + 4 * (size_t)best_inst + 1),
                    __ldg(instances + 4 * (size_t)best_inst + 2), r, o.ox, o.oy, o.oz, o.dx, o.dy, o.dz);
    return o;
}

// MIXED: closest-hit and any-hit rays share one launch; the kind is per ray (TravState::any_hit).
// TWO_LEVEL: the node array starts with a top-level tree whose leaf entries are shapes (InstanceRecord table,
// bvh8.h) instead of triangles; entering one transforms the ray into the shape's object space and walks the
// bottom-level tree of its mesh (takes the role of intersect_bvh2level_skiplinks.cl:135-300,310-470).
template <bool ANY_HIT, bool COUNT, bool MIXED = false, bool TWO_LEVEL = false>
struct Traverser {
    const uint4* __restrict__ nodes;
    const float4* __restrict__ tris;
    const uint4* top = nullptr;     // shared-memory copy of nodes [0, top_count) (stage_top_nodes), or unused
    unsigned top_count = 0;
    const float4* __restrict__ instances = nullptr;   // TWO_LEVEL: InstanceRecord table, 4 float4 each

    __device__ __forceinline__ void init(TravState& s, const Ray& r) const {
        s.ngroup = make_uint2(0u, 0x80000000u);   // root node = child slot 7^octinv of a virtual parent
        s.tgroup = make_uint2(0u, 0u);
        s.sp = 0;
        s.idx = safe_inv(r.dx);
        s.idy = safe_inv(r.dy);
        s.idz = safe_inv(r.dz);
        // signs are taken from the inverse direction so that d = -0.0f (idir = -1e20) is handled as
        // "negative" consistently by the octant order and by the near/far plane selection below
        const uint32_t oct = (s.idx < 0.f ? 1u : 0u) | (s.idy < 0.f ? 2u : 0u) | (s.idz < 0.f ? 4u : 0u);
        s.octinv4 = (7u - oct) * 0x01010101u;
        s.closest = r.tmax;
        s.best_tri = -1;
        s.best_sid = 0x7fffffff;
        s.best_pid = 0x7fffffff;
        s.any_hit = ANY_HIT;
        if (TWO_LEVEL) {
            s.cox = r.ox; s.coy = r.oy; s.coz = r.oz;
            s.cdx = r.dx; s.cdy = r.dy; s.cdz = r.dz;
            s.inst_sp = -1;
            s.cur_inst = s.cur_sid = s.best_inst = -1;
        }
    }

    // Runs until the ray is finished (returns true) or the warp's active-lane count drops under
    // the dynamic-fetch threshold (returns false: caller refills idle lanes and calls again).
    // SPLIT: the lane is one of several that share ONE ray, each restricted to its own interval (s.tlo, s.closest] of
    // t -- same origin and direction, so every t is the number the unrestricted walk computes; boxes that end before
    // tlo are culled, triangles are accepted only for t > tlo.
    template <bool SPLIT = false>
    __device__ __forceinline__ bool run(TravState& s, const Ray& r, uint2* stack, int* counters) const {
        uint2 ngroup = s.ngroup, tgroup = s.tgroup;
        int sp = s.sp;
        uint32_t octinv4 = s.octinv4;
        uint32_t octinv = octinv4 & 7u;
        bool finished = false;
        // the ray in the space being walked
#define MCRT_RO(c) (TWO_LEVEL ? s.co##c : r.o##c)
#define MCRT_RD(c) (TWO_LEVEL ? s.cd##c : r.d##c)
        while (true) {
#if MCRT_POSTPONE_PERCENT > 0
            const int postpone_below = (__popc(__activemask()) * MCRT_POSTPONE_PERCENT) / 100;
#endif
            if (ngroup.y > 0x00FFFFFFu) {
                const uint32_t hits = ngroup.y;
                const uint32_t imask = ngroup.y;
                const uint32_t child_bit = 31u - __clz(hits);
                ngroup.y &= ~(1u << child_bit);
                if (ngroup.y > 0x00FFFFFFu) stack[sp++] = ngroup;
                const uint32_t slot = (child_bit - 24u) ^ octinv;
                const uint32_t rel = __popc(imask & ~(0xFFFFFFFFu << slot));
                const uint32_t ni = ngroup.x + rel;
                uint4 n0, n1, n2, n3, n4;
                if (ni < top_count) {      // a top-level node: from the CTA's shared-memory copy
                    const uint4* np = top + 5 * ni;
                    n0 = np[0];
                    n1 = np[1];
                    n2 = np[2];
                    n3 = np[3];
                    n4 = np[4];
                } else {
                    n0 = __ldg(nodes + 5 * (size_t)ni + 0);
                    n1 = __ldg(nodes + 5 * (size_t)ni + 1);
                    n2 = __ldg(nodes + 5 * (size_t)ni + 2);
                    n3 = __ldg(nodes + 5 * (size_t)ni + 3);
                    n4 = __ldg(nodes + 5 * (size_t)ni + 4);
                }
                if (COUNT) counters[0]++;

                ngroup.x = n1.x;
                tgroup.x = n1.y;
                tgroup.y = 0u;
                uint32_t hitmask = 0u;
#if MCRT_QUANT_PRMT
                // scale = 2^(e + 15): the exponent byte is >= 1 for every node the builder emits, and far below 240
                const float aix = mul_(__uint_as_float((extract_byte(n0.w, 0) + 15u) << 23), s.idx);
                const float aiy = mul_(__uint_as_float((extract_byte(n0.w, 1) + 15u) << 23), s.idy);
                const float aiz = mul_(__uint_as_float((extract_byte(n0.w, 2) + 15u) << 23), s.idz);
                // ... and 2^-7 quantisation step less: every plane, LOW or HIGH, is thereby moved by -2^-7 step.  Together
                // with the HIGH constant's +2^-6 step the LOW planes end up 2^-7 step below and the HIGH planes 2^-7 step
                // above their stored position -- outward on BOTH sides, twice the rounding error of these offsets
                // (2 x 2^-9 step: they are rounded at the magnitude of ai').  r01 biased the HIGH planes only, which
                // keeps a box's own interval open but not the test against the ray's tmax: a shadow ray that ends on
                // the emitter it was aimed at hits at t = tmax - a few ulp on the LOW face of the flat emitter's box,
                // and a near plane that came out 2^-9 step late culled it (r02 parity report: 2-5 such rays per
                // 512^2 Cornell frame disagreed with the brute-force oracle).
                const float orx = __fmaf_rn(aix, -2.384185791015625e-07f, sub_(mul_(sub_(__uint_as_float(n0.x), MCRT_RO(x)), s.idx), aix));
                const float ory = __fmaf_rn(aiy, -2.384185791015625e-07f, sub_(mul_(sub_(__uint_as_float(n0.y), MCRT_RO(y)), s.idy), aiy));
                const float orz = __fmaf_rn(aiz, -2.384185791015625e-07f, sub_(mul_(sub_(__uint_as_float(n0.z), MCRT_RO(z)), s.idz), aiz));
#else
                const float aix = mul_(__uint_as_float(extract_byte(n0.w, 0) << 23), s.idx);
                const float aiy = mul_(__uint_as_float(extract_byte(n0.w, 1) << 23), s.idy);
                const float aiz = mul_(__uint_as_float(extract_byte(n0.w, 2) << 23), s.idz);
                const float orx = mul_(sub_(__uint_as_float(n0.x), MCRT_RO(x)), s.idx);
                const float ory = mul_(sub_(__uint_as_float(n0.y), MCRT_RO(y)), s.idy);
                const float orz = mul_(sub_(__uint_as_float(n0.z), MCRT_RO(z)), s.idz);
#endif
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const uint32_t meta4 = half == 0 ? n1.z : n1.w;
                    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                    const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
                    const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
                    const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
                    const uint32_t qlx = half == 0 ? n2.x : n2.y, qhx = half == 0 ? n3.z : n3.w;
                    const uint32_t qly = half == 0 ? n2.z : n2.w, qhy = half == 0 ? n4.x : n4.y;
                    const uint32_t qlz = half == 0 ? n3.x : n3.y, qhz = half == 0 ? n4.z : n4.w;
                    const uint32_t xmin = s.idx < 0.f ? qhx : qlx, xmax = s.idx < 0.f ? qlx : qhx;
                    const uint32_t ymin = s.idy < 0.f ? qhy : qly, ymax = s.idy < 0.f ? qly : qhy;
                    const uint32_t zmin = s.idz < 0.f ? qhz : qlz, zmax = s.idz < 0.f ? qlz : qhz;
                    // conversion constants travel with the words: HIGH planes get the outward bias
                    const uint32_t cxmin = s.idx < 0.f ? MCRT_QUANT_C_HI : MCRT_QUANT_C_LO, cxmax = cxmin ^ 4u;
                    const uint32_t cymin = s.idy < 0.f ? MCRT_QUANT_C_HI : MCRT_QUANT_C_LO, cymax = cymin ^ 4u;
                    const uint32_t czmin = s.idz < 0.f ? MCRT_QUANT_C_HI : MCRT_QUANT_C_LO, czmax = czmin ^ 4u;
#pragma unroll
#define MCRT_CHILD_TEST(j)                                                                                      \
    {                                                                                                          \
        const float tminx = __fmaf_rn(quant_byte_to_float<j>(xmin, cxmin), aix, orx);                          \
        const float tminy = __fmaf_rn(quant_byte_to_float<j>(ymin, cymin), aiy, ory);                          \
        const float tminz = __fmaf_rn(quant_byte_to_float<j>(zmin, czmin), aiz, orz);                          \
        const float tmaxx = __fmaf_rn(quant_byte_to_float<j>(xmax, cxmax), aix, orx);                          \
        const float tmaxy = __fmaf_rn(quant_byte_to_float<j>(ymax, cymax), aiy, ory);                          \
        const float tmaxz = __fmaf_rn(quant_byte_to_float<j>(zmax, czmax), aiz, orz);                          \
        const float cmin = fmaxf(fmaxf(tminx, tminy), fmaxf(tminz, SPLIT ? s.tlo_box : 0.f));                  \
        /* 1 + 3 ulp widening keeps the quantised slab test conservative w.r.t. fp32 rounding */               \
        const float cmax = fminf(fminf(tmaxx, tmaxy), fminf(tmaxz, s.closest)) * 1.0000004f;                   \
        if (cmin <= cmax) hitmask |= extract_byte(child_bits4, j) << extract_byte(bit_index4, j);              \
    }
                    MCRT_CHILD_TEST(0)
                    MCRT_CHILD_TEST(1)
                    MCRT_CHILD_TEST(2)
                    MCRT_CHILD_TEST(3)
#undef MCRT_CHILD_TEST
                }
                ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
                tgroup.y = hitmask & 0x00FFFFFFu;
#if MCRT_TRAV_PREFETCH & 2
                if (ngroup.y > 0x00FFFFFFu) {      // the child the next iteration will pop
                    const uint32_t nslot = ((31u - __clz(ngroup.y)) - 24u) ^ octinv;
                    const uint4* nn = nodes + 5 * (size_t)(ngroup.x + __popc(ngroup.y & ~(0xFFFFFFFFu << nslot)));
                    prefetch_l1(nn);
                    prefetch_l1(nn + 4);           // an 80-byte node can straddle two 128-byte lines
                }
#endif
            } else {
                tgroup = ngroup;
                ngroup = make_uint2(0u, 0u);
            }

            while (tgroup.y != 0u) {
                if (TWO_LEVEL && s.inst_sp < 0) {
                    // top level: the group's entries are shapes.  Take one, park what is still pending up here, and
                    // go down into its mesh's tree with the ray in its object space.
                    const uint32_t ii = tgroup.x + (31u - __clz(tgroup.y));
                    tgroup.y &= ~(1u << (31u - __clz(tgroup.y)));
                    const float4 m0 = __ldg(instances + 4 * (size_t)ii + 0);
                    const float4 m1 = __ldg(instances + 4 * (size_t)ii + 1);
                    const float4 m2 = __ldg(instances + 4 * (size_t)ii + 2);
                    const float4 w3 = __ldg(instances + 4 * (size_t)ii + 3);
                    const int sid = __float_as_int(w3.y);
                    if (r.mask == sid) continue;      // masked vs this ray (intersect_bvh2level_skiplinks.cl:236-240)
                    if (tgroup.y != 0u) stack[sp++] = tgroup;
                    if (ngroup.y > 0x00FFFFFFu) stack[sp++] = ngroup;
                    s.inst_sp = sp;
                    s.cur_inst = (int)ii;
                    s.cur_sid = sid;
                    to_object_space(m0, m1, m2, r, s.cox, s.coy, s.coz, s.cdx, s.cdy, s.cdz);
                    s.idx = safe_inv(s.cdx);
                    s.idy = safe_inv(s.cdy);
                    s.idz = safe_inv(s.cdz);
                    octinv = 7u - ((s.idx < 0.f ? 1u : 0u) | (s.idy < 0.f ? 2u : 0u) | (s.idz < 0.f ? 4u : 0u));
                    octinv4 = octinv * 0x01010101u;
                    ngroup = make_uint2((uint32_t)__float_as_int(w3.x), 0x80000000u);
                    tgroup = make_uint2(0u, 0u);
                    break;
                }
#if MCRT_POSTPONE_PERCENT > 0
                if (__popc(__activemask()) < postpone_below) {
                    stack[sp++] = tgroup;
                    break;
                }
#endif
                const uint32_t ti = 31u - __clz(tgroup.y);
                tgroup.y &= ~(1u << ti);
                const size_t ta = (size_t)tgroup.x + ti;
#if MCRT_TRAV_PREFETCH & 1
                if (tgroup.y != 0u) {
                    const float4* nt = tris + 3 * ((size_t)tgroup.x + (31u - __clz(tgroup.y)));
                    prefetch_l1(nt);
                    prefetch_l1(nt + 2);
                }
#endif
                const float4 v0 = __ldg(tris + 3 * ta + 0);
                const float4 v1 = __ldg(tris + 3 * ta + 1);
                const float4 v2 = __ldg(tris + 3 * ta + 2);
                if (COUNT) counters[1]++;
                const int sid = TWO_LEVEL ? s.cur_sid : __float_as_int(v0.w);
                if (!TWO_LEVEL && r.mask == sid) continue;
                float t;
                Ray cr = r;
                if (TWO_LEVEL) {
                    cr.ox = s.cox; cr.oy = s.coy; cr.oz = s.coz;
                    cr.dx = s.cdx; cr.dy = s.cdy; cr.dz = s.cdz;
                }
                if (intersect_triangle_ref(cr, v0, v1, v2, s.closest, t) && (!SPLIT || t > s.tlo)) {
                    if (ANY_HIT || (MIXED && s.any_hit)) {
                        if (t < s.closest) {
                            s.best_tri = (int)ta;
                            if (TWO_LEVEL) s.best_inst = s.cur_inst;
                            finished = true;
                            break;
                        }
                    } else {
                        const int pid = __float_as_int(v1.w);
                        const bool closer = t < s.closest;
                        const bool tie = (t == s.closest) && s.best_tri >= 0 &&
                                         (sid < s.best_sid || (sid == s.best_sid && pid < s.best_pid));
                        if (closer || tie) {
                            s.closest = t;
                            s.best_tri = (int)ta;
                            s.best_sid = sid;
                            s.best_pid = pid;
                            if (TWO_LEVEL) s.best_inst = s.cur_inst;
                        }
                    }
                }
            }
            if (finished) break;

            if (ngroup.y <= 0x00FFFFFFu) {
                if (TWO_LEVEL && sp == s.inst_sp) {
                    // the shape's subtree is exhausted: back to world space and to what was parked at the top level
                    s.inst_sp = -1;
                    s.cox = r.ox; s.coy = r.oy; s.coz = r.oz;
                    s.cdx = r.dx; s.cdy = r.dy; s.cdz = r.dz;
                    s.idx = safe_inv(r.dx);
                    s.idy = safe_inv(r.dy);
                    s.idz = safe_inv(r.dz);
                    octinv = 7u - ((s.idx < 0.f ? 1u : 0u) | (s.idy < 0.f ? 2u : 0u) | (s.idz < 0.f ? 4u : 0u));
                    octinv4 = octinv * 0x01010101u;
                }
                if (sp == 0) {
                    finished = true;
                    break;
                }
                ngroup = stack[--sp];
            }
            // dynamic fetch: leave when too few lanes of the warp still traverse
            if (__popc(__activemask()) < MCRT_DYNAMIC_FETCH_THRESHOLD) break;
        }
#undef MCRT_RO
#undef MCRT_RD
        s.ngroup = ngroup;
        s.tgroup = tgroup;
        s.sp = sp;
        s.octinv4 = octinv4;
        return finished;
    }
};

}  // namespace mcrt
