// ref_rr_kernels.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// Compiles Radeon Rays' own intersect_bvh2_lds.cl (+ common.cl), read from /root/reference
// through cl2cpp.py, as C++ and runs one work-item per call.  Build defines mirror the
// reference's build: RR_RAY_MASK on (third_party/CMakeLists.txt:26-36, RR_ENABLE_RAYMASK ON),
// RR_BACKFACE_CULL off, USE_SAFE_MATH off (native_recip -> exact 1/x in this strict build).
#include "clshim.hpp"
#include <vector>
#include <omp.h>

namespace clshim {
thread_local size_t g_gid[3] = {0, 0, 0};
thread_local size_t g_lid[3] = {0, 0, 0};
}

#define RR_RAY_MASK 1

namespace refrr {
#include "rr/intersect_bvh2_lds.cl"
}

static const int kChunk = 16384;

template <typename HitT, typename Fn>
static void run_chunks(const void* nodes, const void* rays, long long numrays, HitT* hits, Fn fn) {
    long long nchunks = (numrays + kChunk - 1) / kChunk;
#pragma omp parallel
    {
        // The kernel addresses rays/hits/stack only through its global index, so a chunk with a
        // rebased pointer and index is identical to one big launch (intersector_lds.cpp:266-295).
        std::vector<uint> stack((size_t)kChunk * 48 + 64);
#pragma omp for schedule(dynamic, 1)
        for (long long c = 0; c < nchunks; ++c) {
            long long base = c * kChunk;
            int n = (int)((numrays - base) < kChunk ? (numrays - base) : kChunk);
            const refrr::ray* r = (const refrr::ray*)rays + base;
            for (int i = 0; i < n; ++i) {
                clshim::g_gid[0] = (size_t)i;
                clshim::g_lid[0] = (size_t)(i % 64);
                fn((const refrr::bvh_node*)nodes, r, &n, stack.data(), hits + base);
            }
        }
    }
}

extern "C" {

// intersect_bvh2_lds.cl:66 intersect_main over numrays work-items.
void ref_rr_intersect(const void* nodes, const void* rays, long long numrays, void* hits) {
    run_chunks(nodes, rays, numrays, (refrr::Intersection*)hits,
               [](const refrr::bvh_node* n, const refrr::ray* r, const int* cnt, uint* st, refrr::Intersection* h) {
                   refrr::intersect_main(n, r, cnt, st, h);
               });
}

// intersect_bvh2_lds.cl:229 occluded_main over numrays work-items.
void ref_rr_occluded(const void* nodes, const void* rays, long long numrays, int* hits) {
    run_chunks(nodes, rays, numrays, hits,
               [](const refrr::bvh_node* n, const refrr::ray* r, const int* cnt, uint* st, int* h) {
                   refrr::occluded_main(n, r, cnt, st, h);
               });
}

int ref_rr_sizeof_ray(void) { return (int)sizeof(refrr::ray); }
int ref_rr_sizeof_intersection(void) { return (int)sizeof(refrr::Intersection); }
int ref_rr_sizeof_node(void) { return (int)sizeof(refrr::bvh_node); }
}
