// lbvh_build.cu -- device-side build of the compressed 8-wide BVH ("rebuild on GPU").
//
// Radeon Rays selects its GPU-built hierarchy with the option acc.type = "hlbvh"
// (RadeonRays/src/intersector/intersector_hlbvh.cpp, src/accelerator/hlbvh.cpp: Morton codes, radix sort,
// Karras' radix tree, bottom-up refit -- all as OpenCL kernels); the engine exposes the option in
// RTScene.cpp:203-238.  The same option selects this builder.  What is built is not RR's binary HLBVH but the
// 8-wide compressed tree of bvh8.h, so that the one traversal kernel serves it:
//   1. k_prim_bounds      world-space triangle boxes (the builder's / RR's fp32 transform order) + centroid bounds
//   2. k_morton           63-bit Morton code of the box centre; cub radix sort of (code, triangle)
//   3. k_radix_tree       Karras 2012: one thread per internal node finds its range and split (ties broken by index)
//   4. k_fit_binary       bottom-up boxes of the binary nodes (second arrival at a node computes it)
//   5. k_collapse_level   one thread per wide node of a level: opens the child with the largest surface area until 8
//                         children (subtrees of <= 3 triangles are leaf children; they are opened last, if slots
//                         remain), assigns octant slots like the host builder, reserves its children's node indices
//                         (contiguous, slot order) and its triangles' positions with one atomic each
//   6. k_emit_triangles   triangle -> emitted position; then refit_bvh8 (refit_kernels.cu) writes the triangle records
//                         and quantises every node bottom-up, exactly as for a host-built tree.
// Hit results do not depend on the tree (closest hit: deterministic tie rule; any hit: boolean), so the parity tests
// of the host-built tree apply unchanged.
#include <cub/device/device_radix_sort.cuh>

#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "lbvh_build.h"

namespace mcrt {

namespace {

__device__ __forceinline__ float xform_row(const float* __restrict__ r, float x, float y, float z) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r[0], x), __fmul_rn(r[1], y)), __fmul_rn(r[2], z)), r[3]);
}
// order-preserving float <-> uint map for atomicMin / atomicMax
__device__ __forceinline__ unsigned f2ord(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned o) {
    const unsigned u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

// bounds[0..2] = min centroid (ordered uint), bounds[3..5] = max centroid
__global__ void __launch_bounds__(256)
k_prim_bounds(const float* __restrict__ vertices, const int* __restrict__ indices, const RefitMesh* __restrict__ meshes,
              const int* __restrict__ face_prefix, int num_meshes, int n, float4* __restrict__ plo,
              float4* __restrict__ phi, unsigned* __restrict__ bounds) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float c[3] = {0.f, 0.f, 0.f};
    const bool live = t < n;
    if (live) {
        int lo = 0, hi = num_meshes;   // last mesh with face_prefix[m] <= t
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (face_prefix[mid] <= t) lo = mid;
            else hi = mid;
        }
        const RefitMesh m = meshes[lo];
        const int face = t - face_prefix[lo];
        const int* id = indices + m.index_base + 3 * (size_t)face;
        float bl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, bh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float* v = vertices + 3 * ((size_t)m.vertex_base + id[k]);
            const float x = v[0], y = v[1], z = v[2];
            const float w[3] = {xform_row(m.m + 0, x, y, z), xform_row(m.m + 4, x, y, z), xform_row(m.m + 8, x, y, z)};
            for (int a = 0; a < 3; ++a) {
                bl[a] = fminf(bl[a], w[a]);
                bh[a] = fmaxf(bh[a], w[a]);
            }
        }
        plo[t] = make_float4(bl[0], bl[1], bl[2], __int_as_float(lo));
        phi[t] = make_float4(bh[0], bh[1], bh[2], __int_as_float(face));
        for (int a = 0; a < 3; ++a) c[a] = 0.5f * (bl[a] + bh[a]);
    }
    for (int a = 0; a < 3; ++a) {
        unsigned mn = live ? f2ord(c[a]) : 0xffffffffu, mx = live ? f2ord(c[a]) : 0u;
        for (int o = 16; o > 0; o >>= 1) {
            mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds + a, mn);
            atomicMax(bounds + 3 + a, mx);
        }
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned v) {   // 21 bits -> every third bit
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(256)
k_morton(const float4* __restrict__ plo, const float4* __restrict__ phi, const unsigned* __restrict__ bounds, int n,
         unsigned long long* __restrict__ keys, unsigned* __restrict__ vals) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const float4 l = plo[t], h = phi[t];
    const float c[3] = {0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)};
    unsigned q[3];
    for (int a = 0; a < 3; ++a) {
        const float mn = ord2f(bounds[a]), mx = ord2f(bounds[3 + a]);
        const float ext = mx - mn;
        float u = ext > 0.f ? (c[a] - mn) / ext : 0.f;
        u = fminf(fmaxf(u, 0.f), 1.f);
        q[a] = min(2097151u, (unsigned)(u * 2097152.0f));
    }
    keys[t] = spread21(q[0]) | (spread21(q[1]) << 1) | (spread21(q[2]) << 2);
    vals[t] = (unsigned)t;
}

// length of the common prefix of the (key, index) pairs i and j; -1 outside the array
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long x = keys[i] ^ keys[j];
    return x ? __clzll((long long)x) : 64 + __clz(i ^ j);
}

// Binary references: r >= 0 internal node r, r < 0 leaf at sorted position ~r.
__global__ void __launch_bounds__(256)
k_radix_tree(const unsigned long long* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
             int* __restrict__ first, int* __restrict__ size, int* __restrict__ parent, int* __restrict__ leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lref = lo == gamma ? ~gamma : gamma;
    const int rref = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    left[i] = lref;
    right[i] = rref;
    first[i] = lo;
    size[i] = hi - lo + 1;
    if (lref < 0) leaf_parent[~lref] = i;
    else parent[lref] = i;
    if (rref < 0) leaf_parent[~rref] = i;
    else parent[rref] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ void ref_box(int r, const unsigned* __restrict__ vals, const float4* __restrict__ plo,
                                        const float4* __restrict__ phi, const float4* ilo, const float4* ihi,
                                        float4& lo, float4& hi) {
    if (r < 0) {
        const unsigned t = vals[~r];
        lo = plo[t];
        hi = phi[t];
    } else {
        lo = __ldcg(ilo + r);
        hi = __ldcg(ihi + r);
    }
}

__global__ void __launch_bounds__(256)
k_fit_binary(const unsigned* __restrict__ vals, const float4* __restrict__ plo, const float4* __restrict__ phi,
             const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
             const int* __restrict__ leaf_parent, int n, float4* ilo, float4* ihi, unsigned* flags) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int cur = leaf_parent[p];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(flags + cur, 1u) == 0u) return;   // the sibling subtree is not finished: its thread goes on
        float4 al, ah, bl, bh;
        ref_box(left[cur], vals, plo, phi, ilo, ihi, al, ah);
        ref_box(right[cur], vals, plo, phi, ilo, ihi, bl, bh);
        __stcg(ilo + cur, make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.f));
        __stcg(ihi + cur, make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.f));
        cur = parent[cur];
    }
}

struct CollapseArgs {
    const unsigned* vals;
    const float4 *plo, *phi, *ilo, *ihi;
    const int *left, *right, *first, *size;
    const int* cur_q;
    int* next_q;
    int count, level_first, next_first, node_cap;
    uint4* nodes;
    unsigned* emit_src;
    unsigned* counters;   // [0] next-level count, [1] triangle cursor, [2] overflow flag
};

__device__ __forceinline__ float half_area(const float4& lo, const float4& hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(128)
k_collapse_level(CollapseArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.count) return;
    const int b = a.cur_q[i];
    int refs[8];
    float area[8];
    int sz[8];
    int n = 2;
    refs[0] = a.left[b];
    refs[1] = a.right[b];
    auto describe = [&](int k) {
        float4 lo, hi;
        ref_box(refs[k], a.vals, a.plo, a.phi, a.ilo, a.ihi, lo, hi);
        area[k] = half_area(lo, hi);
        sz[k] = refs[k] < 0 ? 1 : a.size[refs[k]];
    };
    describe(0);
    describe(1);
    // phase 0 opens subtrees of more than 3 triangles (they would become wide nodes of their own), phase 1 spends
    // the slots that remain on splitting 2- and 3-triangle leaves into tighter ones
    for (int phase = 0; phase < 2; ++phase) {
        while (n < 8) {
            int best = -1;
            float best_area = -1.f;
            for (int k = 0; k < n; ++k) {
                const bool openable = refs[k] >= 0 && (phase == 0 ? sz[k] > 3 : true);
                if (openable && area[k] > best_area) {
                    best_area = area[k];
                    best = k;
                }
            }
            if (best < 0) break;
            const int r = refs[best];
            refs[best] = a.left[r];
            refs[n] = a.right[r];
            describe(best);
            describe(n);
            ++n;
        }
    }
    // octant slots (bvh8_build.cpp): slot s should hold the child that is nearest along direction signs s
    float4 nlo, nhi;
    ref_box(b, a.vals, a.plo, a.phi, a.ilo, a.ihi, nlo, nhi);
    const float pc[3] = {0.5f * (nlo.x + nhi.x), 0.5f * (nlo.y + nhi.y), 0.5f * (nlo.z + nhi.z)};
    float dcen[8][3];
    for (int k = 0; k < n; ++k) {
        float4 lo, hi;
        ref_box(refs[k], a.vals, a.plo, a.phi, a.ilo, a.ihi, lo, hi);
        dcen[k][0] = 0.5f * (lo.x + hi.x) - pc[0];
        dcen[k][1] = 0.5f * (lo.y + hi.y) - pc[1];
        dcen[k][2] = 0.5f * (lo.z + hi.z) - pc[2];
    }
    int child_in_slot[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
    unsigned assigned = 0, used = 0;
    for (int round = 0; round < n; ++round) {
        int bi = -1, bs = -1;
        float bv = -FLT_MAX;
        for (int k = 0; k < n; ++k) {
            if ((assigned >> k) & 1u) continue;
            for (int s = 0; s < 8; ++s) {
                if ((used >> s) & 1u) continue;
                const float proj = ((s & 1) ? -dcen[k][0] : dcen[k][0]) + ((s & 2) ? -dcen[k][1] : dcen[k][1]) +
                                   ((s & 4) ? -dcen[k][2] : dcen[k][2]);
                if (bi < 0 || -proj > bv) {
                    bv = -proj;
                    bi = k;
                    bs = s;
                }
            }
        }
        assigned |= 1u << bi;
        used |= 1u << bs;
        child_in_slot[bs] = bi;
    }
    int n_inner = 0, n_tris = 0;
    for (int k = 0; k < n; ++k) {
        if (refs[k] >= 0 && sz[k] > 3) ++n_inner;
        else n_tris += sz[k];
    }
    const unsigned cbase = n_inner ? atomicAdd(a.counters + 0, (unsigned)n_inner) : 0u;
    const unsigned tbase = n_tris ? atomicAdd(a.counters + 1, (unsigned)n_tris) : 0u;
    if (a.next_first + (long long)cbase + n_inner > (long long)a.node_cap) {
        a.counters[2] = 1u;
        return;
    }
    unsigned imask = 0, meta_lo = 0, meta_hi = 0;
    int inner_rank = 0, tri_off = 0;
    for (int s = 0; s < 8; ++s) {
        const int k = child_in_slot[s];
        if (k < 0) continue;
        unsigned meta;
        if (refs[k] >= 0 && sz[k] > 3) {
            imask |= 1u << s;
            meta = (1u << 5) | (24u + (unsigned)s);
            a.next_q[cbase + inner_rank++] = refs[k];
        } else {
            const unsigned unary = sz[k] == 1 ? 1u : (sz[k] == 2 ? 3u : 7u);
            meta = (unary << 5) | (unsigned)tri_off;
            const int f = refs[k] < 0 ? ~refs[k] : a.first[refs[k]];
            for (int q = 0; q < sz[k]; ++q) a.emit_src[tbase + tri_off + q] = (unsigned)(f + q);
            tri_off += sz[k];
        }
        if (s < 4) meta_lo |= meta << (8 * s);
        else meta_hi |= meta << (8 * (s - 4));
    }
    uint4* node = a.nodes + 5 * (size_t)(a.level_first + i);
    node[0] = make_uint4(0u, 0u, 0u, imask << 24);
    node[1] = make_uint4((unsigned)a.next_first + cbase, tbase, meta_lo, meta_hi);
    node[2] = node[3] = node[4] = make_uint4(0u, 0u, 0u, 0u);
}

// emitted position j holds the triangle at sorted position emit_src[j]
__global__ void __launch_bounds__(256)
k_emit_triangles(const unsigned* __restrict__ emit_src, const unsigned* __restrict__ vals,
                 const float4* __restrict__ plo, const float4* __restrict__ phi, int n, int* __restrict__ tri_mesh,
                 float4* __restrict__ tris) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const unsigned t = vals[emit_src[j]];
    tri_mesh[j] = __float_as_int(plo[t].w);
    tris[3 * (size_t)j + 1].w = phi[t].w;     // face index: what k_refit_triangles reads
}

// one device allocation carved into the builder's temporaries
struct Scratch {
    struct Req {
        void** p;
        size_t bytes;
    };
    std::vector<Req> reqs;
    void* base = nullptr;
    ~Scratch() {
        if (base) cudaFree(base);
    }
    template <typename T>
    void want(T** p, size_t count) {
        reqs.push_back({(void**)p, ((count ? count : 1) * sizeof(T) + 255) & ~(size_t)255});
    }
    cudaError_t commit() {
        size_t total = 0;
        for (const Req& r : reqs) total += r.bytes;
        cudaError_t e = cudaMalloc(&base, total);
        if (e != cudaSuccess) return e;
        size_t off = 0;
        for (const Req& r : reqs) {
            *r.p = (char*)base + off;
            off += r.bytes;
        }
        return cudaSuccess;
    }
};

}  // namespace

#define LB_CUDA(call)                                                        \
    do {                                                                     \
        cudaError_t e_ = (call);                                             \
        if (e_ != cudaSuccess) {                                             \
            *err = std::string(#call) + ": " + cudaGetErrorString(e_);       \
            return false;                                                    \
        }                                                                    \
    } while (0)

bool build_bvh8_device(const float* d_vertices, const int* d_indices, const RefitMesh* d_meshes,
                       const std::vector<int>& face_prefix, DeviceBvh8* out, cudaStream_t stream, std::string* err) {
    const int num_meshes = (int)face_prefix.size() - 1;
    const long long total = num_meshes > 0 ? face_prefix.back() : 0;
    if (total < 4) {
        *err = "fewer than 4 triangles";      // a single leaf: the host builder's degenerate case
        return false;
    }
    if (total > 0x3fffffffLL) {
        *err = "too many triangles";
        return false;
    }
    const int n = (int)total;
    const bool trace = getenv("MCRT_BUILD_TRACE") != nullptr;
    cudaEvent_t ev[6];
    if (trace) for (cudaEvent_t& e : ev) cudaEventCreate(&e);
    Scratch sc;
    int* d_prefix;
    float4 *plo, *phi, *ilo, *ihi;
    unsigned *bounds, *vals, *vals2, *flags, *emit_src, *counters;
    unsigned long long *keys, *keys2;
    int *left, *right, *first, *size, *parent, *leaf_parent, *q0, *q1;
    // every wide node but the root stands for a subtree of >= 4 triangles, disjoint per level and halving upwards
    const int node_cap = n / 2 + 64;
    uint4* nodes_tmp;
    unsigned char* sort_temp;
    size_t sort_temp_bytes = 0;
    {
        cub::DoubleBuffer<unsigned long long> dk(nullptr, nullptr);
        cub::DoubleBuffer<unsigned> dv(nullptr, nullptr);
        LB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_temp_bytes, dk, dv, n, 0, 63, stream));
    }
    sc.want(&d_prefix, face_prefix.size());
    sc.want(&plo, n);
    sc.want(&phi, n);
    sc.want(&ilo, n);
    sc.want(&ihi, n);
    sc.want(&bounds, 8);
    sc.want(&keys, n);
    sc.want(&keys2, n);
    sc.want(&vals, n);
    sc.want(&vals2, n);
    sc.want(&flags, n);
    sc.want(&emit_src, n);
    sc.want(&counters, 4);
    sc.want(&left, n);
    sc.want(&right, n);
    sc.want(&first, n);
    sc.want(&size, n);
    sc.want(&parent, n);
    sc.want(&leaf_parent, n);
    sc.want(&q0, node_cap);
    sc.want(&q1, node_cap);
    sc.want(&nodes_tmp, 5 * (size_t)node_cap);
    sc.want(&sort_temp, sort_temp_bytes);
    LB_CUDA(sc.commit());

    if (trace) cudaEventRecord(ev[0], stream);
    LB_CUDA(cudaMemcpyAsync(d_prefix, face_prefix.data(), face_prefix.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
    const unsigned init_bounds[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    LB_CUDA(cudaMemcpyAsync(bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, stream));
    LB_CUDA(cudaMemsetAsync(flags, 0, (size_t)n * sizeof(unsigned), stream));
    const int blocks = (n + 255) / 256;
    k_prim_bounds<<<blocks, 256, 0, stream>>>(d_vertices, d_indices, d_meshes, d_prefix, num_meshes, n, plo, phi, bounds);
    k_morton<<<blocks, 256, 0, stream>>>(plo, phi, bounds, n, keys, vals);
    if (trace) cudaEventRecord(ev[1], stream);
    {
        cub::DoubleBuffer<unsigned long long> dk(keys, keys2);
        cub::DoubleBuffer<unsigned> dv(vals, vals2);
        LB_CUDA(cub::DeviceRadixSort::SortPairs(sort_temp, sort_temp_bytes, dk, dv, n, 0, 63, stream));
        keys = dk.Current();
        vals = dv.Current();
    }
    if (trace) cudaEventRecord(ev[2], stream);
    k_radix_tree<<<(n - 1 + 255) / 256, 256, 0, stream>>>(keys, n, left, right, first, size, parent, leaf_parent);
    k_fit_binary<<<blocks, 256, 0, stream>>>(vals, plo, phi, left, right, parent, leaf_parent, n, ilo, ihi, flags);
    LB_CUDA(cudaGetLastError());
    if (trace) cudaEventRecord(ev[3], stream);

    // collapse, breadth-first; the level sizes come back to the host (the refit launches need them anyway)
    const unsigned zero4[4] = {0u, 0u, 0u, 0u};
    LB_CUDA(cudaMemcpyAsync(counters, zero4, sizeof(zero4), cudaMemcpyHostToDevice, stream));
    const int root = 0;
    LB_CUDA(cudaMemcpyAsync(q0, &root, sizeof(int), cudaMemcpyHostToDevice, stream));
    std::vector<int> level_offsets(1, 0);
    int count = 1, level_first = 0;
    int *cur_q = q0, *next_q = q1;
    while (count > 0) {
        if ((int)level_offsets.size() > DeviceBvh8::kMaxLevels) {
            *err = "tree deeper than the traversal stack allows";
            return false;
        }
        CollapseArgs a;
        a.vals = vals;
        a.plo = plo;
        a.phi = phi;
        a.ilo = ilo;
        a.ihi = ihi;
        a.left = left;
        a.right = right;
        a.first = first;
        a.size = size;
        a.cur_q = cur_q;
        a.next_q = next_q;
        a.count = count;
        a.level_first = level_first;
        a.next_first = level_first + count;
        a.node_cap = node_cap;
        a.nodes = nodes_tmp;
        a.emit_src = emit_src;
        a.counters = counters;
        k_collapse_level<<<(count + 127) / 128, 128, 0, stream>>>(a);
        unsigned h[4];
        LB_CUDA(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, stream));
        LB_CUDA(cudaStreamSynchronize(stream));
        if (h[2]) {
            *err = "node capacity exceeded";
            return false;
        }
        LB_CUDA(cudaMemsetAsync(counters, 0, sizeof(unsigned), stream));   // next-level count only
        level_first += count;
        level_offsets.push_back(level_first);
        count = (int)h[0];
        std::swap(cur_q, next_q);
        if (count == 0 && h[1] != (unsigned)n) {
            *err = "internal error: emitted " + std::to_string(h[1]) + " of " + std::to_string(n) + " triangles";
            return false;
        }
    }
    const int num_nodes = level_first;
    if (trace) cudaEventRecord(ev[4], stream);

    // results: exact-size node array, triangle records, per-triangle mesh index, per-node boxes
    void *nodes = nullptr, *tris = nullptr;
    int* tri_mesh = nullptr;
    float* boxes = nullptr;
    auto fail_free = [&]() {
        cudaFree(nodes);
        cudaFree(tris);
        cudaFree(tri_mesh);
        cudaFree(boxes);
    };
#define LB_CUDA2(call)                                                       \
    do {                                                                     \
        cudaError_t e_ = (call);                                             \
        if (e_ != cudaSuccess) {                                             \
            *err = std::string(#call) + ": " + cudaGetErrorString(e_);       \
            fail_free();                                                     \
            return false;                                                    \
        }                                                                    \
    } while (0)
    LB_CUDA2(cudaMalloc(&nodes, (size_t)num_nodes * 80));
    LB_CUDA2(cudaMalloc(&tris, (size_t)n * 48));
    LB_CUDA2(cudaMalloc((void**)&tri_mesh, (size_t)n * sizeof(int)));
    LB_CUDA2(cudaMalloc((void**)&boxes, (size_t)num_nodes * 6 * sizeof(float)));
    LB_CUDA2(cudaMemcpyAsync(nodes, nodes_tmp, (size_t)num_nodes * 80, cudaMemcpyDeviceToDevice, stream));
    k_emit_triangles<<<blocks, 256, 0, stream>>>(emit_src, vals, plo, phi, n, tri_mesh, (float4*)tris);
    LB_CUDA2(cudaGetLastError());
    LB_CUDA2(refit_bvh8(nodes, tris, boxes, tri_mesh, d_meshes, d_vertices, d_indices, n, level_offsets.data(),
                        (int)level_offsets.size() - 1, stream));
    if (trace) cudaEventRecord(ev[5], stream);
    LB_CUDA2(cudaStreamSynchronize(stream));
    if (trace) {
        const char* names[5] = {"bounds + morton", "radix sort", "radix tree + fit", "wide collapse", "emit + quantise"};
        for (int k = 0; k < 5; ++k) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, ev[k], ev[k + 1]);
            fprintf(stderr, "[lbvh] %-18s %8.3f ms\n", names[k], ms);
        }
        for (cudaEvent_t& e : ev) cudaEventDestroy(e);
    }
#undef LB_CUDA2
    out->nodes = nodes;
    out->tris = tris;
    out->tri_mesh = tri_mesh;
    out->boxes = boxes;
    out->num_nodes = num_nodes;
    out->num_tris = n;
    out->level_offsets = level_offsets;
    out->launches = 6 + 2 * ((int)level_offsets.size() - 1) + 2;
    return true;
}

}  // namespace mcrt
