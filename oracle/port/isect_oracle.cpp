// isect_oracle.cpp -- TEST INFRASTRUCTURE (oracle, CPU restatement).  Never on the product path.
//
// Plain C++ restatement of Radeon Rays' ray/triangle arithmetic, as an O(rays x triangles)
// brute force -- the same role utils.cpp:28-163 plays for the reference's own conformance tests
// (UnitTest/radeon_rays_conformance_test_cl.h:532-542), but using the KERNEL's formulation so
// that every accepted t is bit-identical to the kernel's:
//   * fast_intersect_triangle         RadeonRays/src/kernels/CL/common.cl:178-218
//   * triangle_calculate_barycentrics RadeonRays/src/kernels/CL/common.cl:250-277
//   * closest-hit accept `t < closest_t`, mask rule, inactive rays not written
//                                     RadeonRays/src/kernels/CL/intersect_bvh2_lds.cl:84-223
//   * any-hit HIT_MARKER 1 / MISS_MARKER -1          intersect_bvh2_lds.cl:229-363
// Pinned against oracle/_ref (the compiled reference kernel) and the known-answer cases of
// UnitTest/radeon_rays_apitest_cl.h in tests/test_oracle.py.
//
// Build: g++ -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/Makefile).
// Strict IEEE fp32: native_recip(x) := 1.0f / x, no contraction, dot() summed left to right.
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct Ray {     // RadeonRays::ray, include/math/ray.h:103-107
    float o[4];  // xyz origin, w = max t
    float d[4];  // xyz direction, w = time
    int extra[2];  // x = mask, y = active
    int backface, pad;
};
struct Isect {   // RadeonRays::Intersection, radeon_rays.h:140-154
    int shapeid, primid, pad0, pad1;
    float uvwt[4];
};
struct Tri {     // world-space vertices + ids (what a Bvh2 leaf node holds, bvh2.h:346-369)
    float v0[3];
    int shape;
    float v1[3];
    int prim;
    float v2[3];
    int pad;
};

inline float dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline void cross3(const float* a, const float* b, float* r) {
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}

// common.cl:178-218
float intersect_triangle(const Ray& r, const float* v1, const float* v2, const float* v3, float t_max) {
    float e1[3] = {v2[0] - v1[0], v2[1] - v1[1], v2[2] - v1[2]};
    float e2[3] = {v3[0] - v1[0], v3[1] - v1[1], v3[2] - v1[2]};
    float s1[3];
    cross3(r.d, e2, s1);
    float denom = dot3(s1, e1);
    if (denom == 0.f) return t_max;
    float invd = 1.0f / denom;
    float d[3] = {r.o[0] - v1[0], r.o[1] - v1[1], r.o[2] - v1[2]};
    float b1 = dot3(d, s1) * invd;
    float s2[3];
    cross3(d, e1, s2);
    float b2 = dot3(r.d, s2) * invd;
    float temp = dot3(e2, s2) * invd;
    if (b1 < 0.f || b1 > 1.f || b2 < 0.f || b1 + b2 > 1.f || temp < 0.f || temp > t_max) return t_max;
    return temp;
}

// common.cl:250-277
void barycentrics(const float* p, const float* v1, const float* v2, const float* v3, float* uv) {
    float e1[3] = {v2[0] - v1[0], v2[1] - v1[1], v2[2] - v1[2]};
    float e2[3] = {v3[0] - v1[0], v3[1] - v1[1], v3[2] - v1[2]};
    float e[3] = {p[0] - v1[0], p[1] - v1[1], p[2] - v1[2]};
    float d00 = dot3(e1, e1), d01 = dot3(e1, e2), d11 = dot3(e2, e2), d20 = dot3(e, e1), d21 = dot3(e, e2);
    float denom = d00 * d11 - d01 * d01;
    if (denom == 0.f) {
        uv[0] = uv[1] = 0.f;
        return;
    }
    float inv = 1.0f / denom;
    uv[0] = (d11 * d20 - d01 * d21) * inv;
    uv[1] = (d00 * d21 - d01 * d20) * inv;
}

}  // namespace

extern "C" {

// World-space triangle records from one mesh: ((m00*x + m01*y) + m02*z) + m03, the order of
// transform_point (include/math/mathutils.h:111-118) as used by Bvh2::SetPrimitive.
void oracle_make_triangles(const float* verts, int vstride_bytes, const int* idx, int nfaces, const float* m16,
                           int shape_id, void* out_tris) {
    Tri* out = (Tri*)out_tris;
    const int vs = vstride_bytes ? vstride_bytes : 12;
    for (int f = 0; f < nfaces; ++f) {
        float* dst[3] = {out[f].v0, out[f].v1, out[f].v2};
        for (int k = 0; k < 3; ++k) {
            const float* v = (const float*)((const char*)verts + (size_t)idx[3 * f + k] * vs);
            for (int r = 0; r < 3; ++r) {
                const float* row = m16 + 4 * r;
                float s = row[0] * v[0];
                s = s + row[1] * v[1];
                s = s + row[2] * v[2];
                s = s + row[3];
                dst[k][r] = s;
            }
        }
        out[f].shape = shape_id;
        out[f].prim = f;
        out[f].pad = 0;
    }
}

// Closest hit over all triangles.  `order`: 0 = smallest (shape, prim) wins among equal t (the
// CUDA path's documented rule); the tied candidate set is reported through tie_count[i] (number
// of triangles accepted at exactly the winning t) so tests can classify mismatches.
void oracle_intersect(const void* tris_, long long ntris, const void* rays_, long long nrays, void* hits_,
                      int* tie_count) {
    const Tri* tris = (const Tri*)tris_;
    const Ray* rays = (const Ray*)rays_;
    Isect* hits = (Isect*)hits_;
#pragma omp parallel for schedule(dynamic, 64)
    for (long long i = 0; i < nrays; ++i) {
        const Ray& r = rays[i];
        if (r.extra[1] == 0) continue;  // inactive: nothing written
        float closest = r.o[3];
        long long best = -1;
        int ties = 0;
        for (long long k = 0; k < ntris; ++k) {
            const Tri& t = tris[k];
            if (r.extra[0] == t.shape) continue;  // RR_RAY_MASK rule
            float tt = intersect_triangle(r, t.v0, t.v1, t.v2, closest);
            if (tt < closest) {
                closest = tt;
                best = k;
                ties = 1;
            } else if (best >= 0 && tt == closest) {
                // fast_intersect_triangle returns t_max both for "rejected" and for "accepted at
                // exactly t_max": re-test against a larger bound to tell them apart
                float chk = intersect_triangle(r, t.v0, t.v1, t.v2, closest * 2.0f + 1.0f);
                if (chk == closest) {
                    ++ties;
                    const Tri& b = tris[best];
                    if (t.shape < b.shape || (t.shape == b.shape && t.prim < b.prim)) best = k;
                }
            }
        }
        if (best >= 0) {
            const Tri& t = tris[best];
            float p[3] = {r.o[0] + closest * r.d[0], r.o[1] + closest * r.d[1], r.o[2] + closest * r.d[2]};
            float uv[2];
            barycentrics(p, t.v0, t.v1, t.v2, uv);
            hits[i].shapeid = t.shape;
            hits[i].primid = t.prim;
            hits[i].uvwt[0] = uv[0];
            hits[i].uvwt[1] = uv[1];
            hits[i].uvwt[2] = 0.0f;
            hits[i].uvwt[3] = closest;
        } else {
            hits[i].shapeid = -1;
            hits[i].primid = -1;
        }
        if (tie_count) tie_count[i] = ties;
    }
}

// Is (shape, prim) one of the triangles accepted at exactly t for this ray?  Used to check that a
// tie was resolved inside the tied set.
int oracle_hit_in_tied_set(const void* tris_, long long ntris, const void* ray_, float t, int shape, int prim) {
    const Tri* tris = (const Tri*)tris_;
    const Ray& r = *(const Ray*)ray_;
    for (long long k = 0; k < ntris; ++k) {
        const Tri& tr = tris[k];
        if (tr.shape != shape || tr.prim != prim) continue;
        float chk = intersect_triangle(r, tr.v0, tr.v1, tr.v2, t * 2.0f + 1.0f);
        return chk == t ? 1 : 0;
    }
    return 0;
}

void oracle_occluded(const void* tris_, long long ntris, const void* rays_, long long nrays, int* out) {
    const Tri* tris = (const Tri*)tris_;
    const Ray* rays = (const Ray*)rays_;
#pragma omp parallel for schedule(dynamic, 64)
    for (long long i = 0; i < nrays; ++i) {
        const Ray& r = rays[i];
        if (r.extra[1] == 0) continue;
        const float closest = r.o[3];
        int res = -1;
        for (long long k = 0; k < ntris; ++k) {
            const Tri& t = tris[k];
            if (r.extra[0] == t.shape) continue;
            if (intersect_triangle(r, t.v0, t.v1, t.v2, closest) < closest) {
                res = 1;
                break;
            }
        }
        out[i] = res;
    }
}

// ---- two-level worlds (instances): intersect_bvh2level_skiplinks.cl
// Every attached shape is visited; the ray is taken into the shape's object space with transform_point /
// transform_vector on the rows of its world -> object matrix (:103-131, the direction is not re-normalised, so t is
// shared by both spaces), tested there against the OBJECT-space triangles of its mesh with the same triangle test
// (:205-222), the hit point and barycentrics taken in object space, and the shape's id reported (:213).  Brute
// force over shapes x triangles; ties at equal t go to the smallest (shape id, prim), the CUDA path's rule.
struct Shape2L {     // one attached shape
    float minv[12];  // rows 0..2 of the world -> object matrix
    int first_tri;   // its mesh's object-space triangles in `tris`
    int num_tris;
    int id;
    int pad;
};
static inline Ray to_object(const Ray& r, const float* m) {
    Ray o = r;
    for (int k = 0; k < 3; ++k) {
        const float* row = m + 4 * k;
        o.o[k] = row[0] * r.o[0] + row[1] * r.o[1] + row[2] * r.o[2] + row[3];
        o.d[k] = row[0] * r.d[0] + row[1] * r.d[1] + row[2] * r.d[2];
    }
    return o;
}
void oracle_intersect_2level(const void* tris_, const void* shapes_, int nshapes, const void* rays_, long long nrays,
                             void* hits_) {
    const Tri* tris = (const Tri*)tris_;
    const Shape2L* shapes = (const Shape2L*)shapes_;
    const Ray* rays = (const Ray*)rays_;
    Isect* hits = (Isect*)hits_;
#pragma omp parallel for schedule(dynamic, 64)
    for (long long i = 0; i < nrays; ++i) {
        const Ray& r = rays[i];
        if (r.extra[1] == 0) continue;
        float closest = r.o[3];
        int best_shape = -1, best_id = 0, best_prim = 0;
        long long best_tri = -1;
        for (int s = 0; s < nshapes; ++s) {
            const Shape2L& sh = shapes[s];
            if (r.extra[0] == sh.id) continue;
            const Ray o = to_object(r, sh.minv);
            for (int k = 0; k < sh.num_tris; ++k) {
                const Tri& t = tris[sh.first_tri + k];
                float tt = intersect_triangle(o, t.v0, t.v1, t.v2, closest);
                bool take = tt < closest;
                if (!take && best_tri >= 0 && tt == closest &&
                    intersect_triangle(o, t.v0, t.v1, t.v2, closest * 2.0f + 1.0f) == closest)
                    take = sh.id < best_id || (sh.id == best_id && t.prim < best_prim);
                if (take) {
                    closest = tt;
                    best_shape = s;
                    best_id = sh.id;
                    best_prim = t.prim;
                    best_tri = sh.first_tri + k;
                }
            }
        }
        if (best_tri >= 0) {
            const Ray o = to_object(r, shapes[best_shape].minv);
            const Tri& t = tris[best_tri];
            float p[3] = {o.o[0] + closest * o.d[0], o.o[1] + closest * o.d[1], o.o[2] + closest * o.d[2]};
            float uv[2];
            barycentrics(p, t.v0, t.v1, t.v2, uv);
            hits[i].shapeid = best_id;
            hits[i].primid = best_prim;
            hits[i].uvwt[0] = uv[0];
            hits[i].uvwt[1] = uv[1];
            hits[i].uvwt[2] = 0.0f;
            hits[i].uvwt[3] = closest;
        } else {
            hits[i].shapeid = -1;
            hits[i].primid = -1;
        }
    }
}
void oracle_occluded_2level(const void* tris_, const void* shapes_, int nshapes, const void* rays_, long long nrays,
                            int* out) {
    const Tri* tris = (const Tri*)tris_;
    const Shape2L* shapes = (const Shape2L*)shapes_;
    const Ray* rays = (const Ray*)rays_;
#pragma omp parallel for schedule(dynamic, 64)
    for (long long i = 0; i < nrays; ++i) {
        const Ray& r = rays[i];
        if (r.extra[1] == 0) continue;
        const float closest = r.o[3];
        int res = -1;
        for (int s = 0; s < nshapes && res < 0; ++s) {
            const Shape2L& sh = shapes[s];
            if (r.extra[0] == sh.id) continue;
            const Ray o = to_object(r, sh.minv);
            for (int k = 0; k < sh.num_tris; ++k) {
                const Tri& t = tris[sh.first_tri + k];
                if (intersect_triangle(o, t.v0, t.v1, t.v2, closest) < closest) {
                    res = 1;
                    break;
                }
            }
        }
        out[i] = res;
    }
}

// ---- RNG / sampler restatement (assets/kernels/rng.cl:48-66,104-107; samplers.cl:83-84)
uint32_t oracle_wang_hash(uint32_t seed) {
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}
uint32_t oracle_xorshift(uint32_t* state) {
    *state += 2463534242u;
    *state ^= (*state << 13);
    *state ^= (*state >> 17);
    *state ^= (*state << 5);
    return *state;
}
float oracle_rand_float(uint32_t* state) { return ((float)oracle_xorshift(state)) / 4294967296.0f; }
uint32_t oracle_sampler_seed(uint32_t buffer_idx, int frame, int width, int height, int offset) {
    return oracle_wang_hash(buffer_idx + (uint32_t)(frame + 1) * (uint32_t)width * (uint32_t)height * (uint32_t)(offset + 1));
}
}
