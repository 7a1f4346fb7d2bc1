// refit_kernels.cu -- device-side refit of the compressed 8-wide BVH after shape transforms changed
// (SURVEY 8f row 3).  The reference rebuilds its whole BVH on the CPU for any transform change
// (intersector_lds.cpp:166 -> Bvh2::Build); here the topology of the last host build is kept and
//   1. every triangle record is recomputed from the object-space vertices with the shape's new
//      transform -- the same fp32 expression order as the host builder / Radeon Rays
//      (((m0*x + m1*y) + m2*z) + m3, bvh2.h:346-369), so hit results equal those of a rebuild;
//   2. the nodes are re-fitted bottom-up, one launch per tree level (nodes are stored breadth-first):
//      exact child boxes from the triangles or from the child nodes' boxes, a new origin / exponent
//      frame, and child planes re-quantised outwards exactly like bvh8_build.cpp does.
// Octant slot order and tree shape are those of the build; only the quality of the order can degrade.
#include "rrb_kernels.h"

#include <stdint.h>

#include <cfloat>

namespace mcrt {

__device__ __forceinline__ float xform(const float* __restrict__ r, float x, float y, float z) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r[0], x), __fmul_rn(r[1], y)), __fmul_rn(r[2], z)), r[3]);
}

// one thread per emitted triangle
__global__ void __launch_bounds__(256)
k_refit_triangles(float4* __restrict__ tris, const int* __restrict__ tri_mesh, const RefitMesh* __restrict__ meshes,
                  const float* __restrict__ vertices, const int* __restrict__ indices, int num_tris) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tris) return;
    const RefitMesh m = meshes[tri_mesh[t]];
    const int face = __float_as_int(tris[3 * (size_t)t + 1].w);
    const int* id = indices + m.index_base + 3 * (size_t)face;
    float4 out[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float* v = vertices + 3 * ((size_t)m.vertex_base + id[k]);
        const float x = v[0], y = v[1], z = v[2];
        out[k].x = xform(m.m + 0, x, y, z);
        out[k].y = xform(m.m + 4, x, y, z);
        out[k].z = xform(m.m + 8, x, y, z);
    }
    out[0].w = __int_as_float(m.shape_id);
    out[1].w = __int_as_float(face);
    out[2].w = __int_as_float(0);
    tris[3 * (size_t)t + 0] = out[0];
    tris[3 * (size_t)t + 1] = out[1];
    tris[3 * (size_t)t + 2] = out[2];
}

// one thread per node of one level; boxes[] holds the exact bounds of every node's subtree (lo xyz, hi xyz)
__global__ void __launch_bounds__(128)
k_refit_level(uint4* __restrict__ nodes, const float4* __restrict__ tris, float* __restrict__ boxes, int first, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int ni = first + i;
    uint4* n = nodes + 5 * (size_t)ni;
    const uint4 n0 = n[0], n1 = n[1];
    const uint32_t imask = n0.w >> 24;
    const uint32_t child_base = n1.x, tri_base = n1.y;
    float clo[8][3], chi[8][3];
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    uint32_t present = 0;
    for (int s = 0; s < 8; ++s) {
        const uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xFFu;
        if (meta == 0) continue;
        present |= 1u << s;
        float bl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, bh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        if ((imask >> s) & 1u) {
            const uint32_t ci = child_base + __popc(imask & ((1u << s) - 1u));
            const float* b = boxes + 6 * (size_t)ci;
            for (int a = 0; a < 3; ++a) {
                bl[a] = b[a];
                bh[a] = b[3 + a];
            }
        } else {
            const int cnt = __popc(meta >> 5);
            const uint32_t t0 = tri_base + (meta & 31u);
            for (int k = 0; k < cnt; ++k)
                for (int v = 0; v < 3; ++v) {
                    const float4 p = tris[3 * (size_t)(t0 + k) + v];
                    bl[0] = fminf(bl[0], p.x); bl[1] = fminf(bl[1], p.y); bl[2] = fminf(bl[2], p.z);
                    bh[0] = fmaxf(bh[0], p.x); bh[1] = fmaxf(bh[1], p.y); bh[2] = fmaxf(bh[2], p.z);
                }
        }
        for (int a = 0; a < 3; ++a) {
            clo[s][a] = bl[a];
            chi[s][a] = bh[a];
            lo[a] = fminf(lo[a], bl[a]);
            hi[a] = fmaxf(hi[a], bh[a]);
        }
    }
    float* nb = boxes + 6 * (size_t)ni;
    for (int a = 0; a < 3; ++a) {
        nb[a] = lo[a];
        nb[3 + a] = hi[a];
    }
    // quantisation frame (bvh8_build.cpp): scale = 2^ex, the smallest power of two with ceil(extent / scale) <= 255
    uint32_t ebits = 0;
    float scale[3];
    for (int a = 0; a < 3; ++a) {
        const double ext = (double)hi[a] - (double)lo[a];
        int ex = -100;
        if (ext > 0.0) {
            ex = (int)ceil(log2(ext / 255.0));
            while (ceil(ext / ldexp(1.0, ex)) > 255.0) ++ex;
        }
        ex = max(-126, min(127, ex));
        scale[a] = ldexpf(1.0f, ex);
        ebits |= (uint32_t)((ex + 127) & 0xFF) << (8 * a);
    }
    uint32_t q[6][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};
    for (int s = 0; s < 8; ++s) {
        if (!((present >> s) & 1u)) continue;
        for (int a = 0; a < 3; ++a) {
            const double l = floor(((double)clo[s][a] - (double)lo[a]) / (double)scale[a]);
            const double h = ceil(((double)chi[s][a] - (double)lo[a]) / (double)scale[a]);
            int qlo = (int)fmax(0.0, fmin(255.0, l));
            int qhi = (int)fmax(0.0, fmin(255.0, h));
            while (qlo > 0 && __fadd_rn(lo[a], __fmul_rn((float)qlo, scale[a])) > clo[s][a]) --qlo;
            while (qhi < 255 && __fadd_rn(lo[a], __fmul_rn((float)qhi, scale[a])) < chi[s][a]) ++qhi;
            q[a][s >> 2] |= (uint32_t)qlo << (8 * (s & 3));
            q[3 + a][s >> 2] |= (uint32_t)qhi << (8 * (s & 3));
        }
    }
    n[0] = make_uint4(__float_as_uint(lo[0]), __float_as_uint(lo[1]), __float_as_uint(lo[2]), ebits | (imask << 24));
    n[2] = make_uint4(q[0][0], q[0][1], q[1][0], q[1][1]);      // qlo_x | qlo_y
    n[3] = make_uint4(q[2][0], q[2][1], q[3][0], q[3][1]);      // qlo_z | qhi_x
    n[4] = make_uint4(q[4][0], q[4][1], q[5][0], q[5][1]);      // qhi_y | qhi_z
}

cudaError_t refit_bvh8(void* nodes, void* tris, float* boxes, const int* tri_mesh, const RefitMesh* meshes,
                       const float* vertices, const int* indices, int num_tris, const int* level_offsets_host,
                       int num_levels, cudaStream_t stream) {
    k_refit_triangles<<<(num_tris + 255) / 256, 256, 0, stream>>>((float4*)tris, tri_mesh, meshes, vertices, indices,
                                                                  num_tris);
    for (int l = num_levels - 1; l >= 0; --l) {
        const int first = level_offsets_host[l], count = level_offsets_host[l + 1] - first;
        if (count <= 0) continue;
        k_refit_level<<<(count + 127) / 128, 128, 0, stream>>>((uint4*)nodes, (const float4*)tris, boxes, first, count);
    }
    return cudaGetLastError();
}

}  // namespace mcrt
