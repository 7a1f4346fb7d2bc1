// bvh8.h -- compressed 8-wide BVH: host-side build interface and the HBM layout shared with the
// traversal kernels.
//
// Replaces Radeon Rays' Bvh2 (RadeonRays/src/accelerator/bvh2.{h,cpp}: binary tree, 64-byte
// nodes, one triangle per leaf stored inside the node) with an 8-wide tree of 80-byte nodes
// whose child boxes are quantised to 8 bits per plane relative to the node's origin
// (Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide
// BVHs", HPG 2017), plus a separate array of 48-byte triangle records.  The tree shape is not
// a parity requirement; the hit results are.
//
// HBM layout
//   nodes:     uint4[5 * num_nodes]           80 B per node
//     [0] p.x p.y p.z (fp32 origin) | e.x e.y e.z imask (4 x u8: per-axis exponent, internal mask)
//     [1] child_base_idx | tri_base_idx | meta[0..3] | meta[4..7]
//     [2] qlo_x[0..7] | qlo_y[0..7]     (u8 each)
//     [3] qlo_z[0..7] | qhi_x[0..7]
//     [4] qhi_y[0..7] | qhi_z[0..7]
//     meta[i]: empty 0; internal child 001_11sss (sss = slot); leaf ccc_ooooo with ccc the
//     unary triangle count (001,011,111) and ooooo < 24 the first triangle relative to tri_base.
//   triangles: float4[3 * num_tris]           48 B per triangle
//     [0] v0.xyz | shape id (int bits)   [1] v1.xyz | prim id (int bits)   [2] v2.xyz | 0
//     World-space vertices computed exactly as Bvh2::SetPrimitive does (bvh2.h:346-369,
//     math/mathutils.h:111-118): ((m00*x + m01*y) + m02*z) + m03, fp32, no contraction.
#pragma once
#include <cstdint>
#include <vector>

namespace mcrt {

struct BuildInputMesh {
    const float* vertices;   // object space, xyz at `vstride` bytes
    int vnum;
    int vstride;             // bytes
    const int* indices;      // 3 per face at `istride` bytes
    int istride;             // bytes
    int numfaces;
    int shape_id;
    float m[16];             // row-major object -> world
};

struct Bvh8 {
    std::vector<uint32_t> nodes;       // 20 words per node
    std::vector<float> triangles;      // 12 floats per triangle
    int64_t num_nodes = 0;
    int64_t num_tris = 0;
    // for the device refit (rrb_kernels.cu: refit_bvh8): which input mesh each emitted triangle came from,
    // and where each tree level starts in the node array (nodes are emitted breadth-first)
    std::vector<int32_t> tri_mesh;        // [num_tris] index into the build's mesh list
    std::vector<int32_t> level_offsets;   // [num_levels + 1]
    float scene_min[3] = {0, 0, 0};
    float scene_max[3] = {0, 0, 0};
};

// One attached shape of the two-level structure (the role of the `Shape` record of Radeon Rays' two-level kernel,
// intersect_bvh2level_skiplinks.cl:75-93): rows 0..2 of the world -> object matrix, the node index of the root of its
// mesh's bottom-level tree, and the id hits report.  64 bytes, in the order of the top-level tree's leaves.
struct InstanceRecord {
    float minv[12];
    int blas_root;
    int shape_id;
    int pad[2];
};

struct BuildOptions {
    bool use_sah = true;          // bvh.builder = sah | median
    int num_bins = 16;            // bins per axis for the binned SAH sweep
    float traversal_cost = 1.2f;  // relative to one triangle test, for the BVH2 leaf decision
    int num_threads = 0;          // 0 = hardware concurrency
    bool depth_guard = true;      // rebuild with balanced splits when the tree is deeper than the traversal stack allows
    // Collapse of the binary tree into 8-wide nodes: SAH-optimal forest per node by dynamic programming (Ylitie et al.
    // 2017, section 3.1), which also forms the leaves (at most 3 triangles); false = greedy largest-area-first.
    // MCRT_BVH_COLLAPSE=greedy|optimal, MCRT_BVH_COST=node:prim override (development switches).
    bool optimal_collapse = true;
    bool optimal_slots = true;         // child -> slot assignment with the largest total score instead of greedy (MCRT_BVH_SLOTS)
    float collapse_node_cost = 1.0f;   // cost of fetching + testing one 80-byte node ...
    float collapse_prim_cost = 0.6f;   // ... relative to one 48-byte triangle
    int max_levels = 19;          // ... i.e. than this many levels (two stack entries per level, 40-entry stack)
};

// Builds the compressed wide BVH over all faces of all meshes.  Throws std::runtime_error.
void build_bvh8(const std::vector<BuildInputMesh>& meshes, const BuildOptions& opt, Bvh8& out);
// The same tree over `n` axis-aligned boxes (6 floats each: lo.xyz, hi.xyz) -- the top level of the two-level
// structure.  Leaf record i (12 floats, in out.triangles) carries the index of its box in the int bits of word 3.
void build_bvh8_boxes(const float* boxes, int n, const BuildOptions& opt, Bvh8& out);

}  // namespace mcrt
