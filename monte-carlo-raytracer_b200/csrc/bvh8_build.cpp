// bvh8_build.cpp -- host builder of the compressed 8-wide BVH (see bvh8.h).
//
// Takes the role of Bvh2::Build / BuildImpl / HandleRequest / FindSahSplit
// (RadeonRays/src/accelerator/bvh2.h:206-315, bvh2.cpp:144-712) behind IntersectionApi::Commit,
// with a different algorithm: (1) multi-threaded binned-SAH binary build over all three axes with
// leaves of up to 3 triangles, (2) greedy surface-area collapse to 8-wide nodes, (3) octant-aware
// child slot assignment, (4) breadth-first emission into quantised 80-byte nodes and 48-byte
// triangle records.
#include "bvh8.h"

#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstring>
#include <future>
#include <limits>
#include <stdexcept>
#include <thread>

namespace mcrt {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() {
        for (int a = 0; a < 3; ++a) {
            lo[a] = std::numeric_limits<float>::infinity();
            hi[a] = -std::numeric_limits<float>::infinity();
        }
    }
    void grow(const Box& b) {
        for (int a = 0; a < 3; ++a) {
            lo[a] = std::min(lo[a], b.lo[a]);
            hi[a] = std::max(hi[a], b.hi[a]);
        }
    }
    void grow(const float* p) {
        for (int a = 0; a < 3; ++a) {
            lo[a] = std::min(lo[a], p[a]);
            hi[a] = std::max(hi[a], p[a]);
        }
    }
    float half_area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0.f)) return 0.f;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Node2 {
    Box box;
    int left = -1, right = -1;  // children (internal)
    int first = 0, count = 0;   // primitive range (leaf when count > 0)
};

struct Builder {
    const BuildOptions& opt;
    std::vector<Box> pbox;
    std::vector<float> pcent;  // 3 per primitive
    std::vector<int> prim;
    std::vector<Node2> nodes;
    std::atomic<int> next_node{0};
    std::atomic<int> live_threads{1};
    int max_threads = 1;
    // Binary depth from which splits are forced to equal counts (balanced, so the remaining subtree collapses three
    // binary levels per wide level).  INT_MAX = SAH everywhere; build_bvh8 lowers it only for a tree whose wide
    // depth would overflow the traversal stack (kMaxWideLevels).
    int sah_depth_cap = std::numeric_limits<int>::max();

    explicit Builder(const BuildOptions& o) : opt(o) {}

    int alloc_node() { return next_node.fetch_add(1); }

    void build_range(int ni, int first, int count, int depth) {
        Node2& n = nodes[(size_t)ni];
        Box b, cb;
        b.reset();
        cb.reset();
        for (int i = first; i < first + count; ++i) {
            b.grow(pbox[(size_t)prim[(size_t)i]]);
            cb.grow(&pcent[3 * (size_t)prim[(size_t)i]]);
        }
        n.box = b;
        if (count == 1) {
            n.first = first;
            n.count = count;
            return;
        }

        int best_axis = -1, best_bin = -1;
        float best_cost = std::numeric_limits<float>::infinity();
        const int NB = std::max(4, std::min(opt.num_bins, 64));
        const bool forced = depth >= sah_depth_cap;
        if (forced) {
            if (count <= 3) {
                n.first = first;
                n.count = count;
                return;
            }
        } else if (opt.use_sah) {
            for (int axis = 0; axis < 3; ++axis) {
                float cmin = cb.lo[axis], cext = cb.hi[axis] - cb.lo[axis];
                if (!(cext > 0.f)) continue;
                Box bins[64];
                int cnt[64];
                for (int k = 0; k < NB; ++k) {
                    bins[k].reset();
                    cnt[k] = 0;
                }
                float scale = (float)NB / cext;
                for (int i = first; i < first + count; ++i) {
                    int p = prim[(size_t)i];
                    int k = std::min(NB - 1, std::max(0, (int)((pcent[3 * (size_t)p + axis] - cmin) * scale)));
                    bins[k].grow(pbox[(size_t)p]);
                    cnt[k]++;
                }
                float right_area[64];
                int right_cnt[64];
                Box acc;
                acc.reset();
                int c = 0;
                for (int k = NB - 1; k > 0; --k) {
                    acc.grow(bins[k]);
                    c += cnt[k];
                    right_area[k] = acc.half_area();
                    right_cnt[k] = c;
                }
                acc.reset();
                c = 0;
                for (int k = 0; k < NB - 1; ++k) {
                    acc.grow(bins[k]);
                    c += cnt[k];
                    if (c == 0 || right_cnt[k + 1] == 0) continue;
                    float cost = acc.half_area() * (float)c + right_area[k + 1] * (float)right_cnt[k + 1];
                    if (cost < best_cost) {
                        best_cost = cost;
                        best_axis = axis;
                        best_bin = k;
                    }
                }
            }
            // Leaf decision (leaves hold at most 3 triangles, the wide node's per-child maximum).
            if (count <= 3) {
                float leaf_cost = (float)count * b.half_area();
                float split_cost = opt.traversal_cost * b.half_area() + best_cost;
                if (best_axis < 0 || leaf_cost <= split_cost) {
                    n.first = first;
                    n.count = count;
                    return;
                }
            }
        } else if (count <= 2) {
            n.first = first;
            n.count = count;
            return;
        }

        int mid;
        if (forced) {
            int axis = 0;
            float ext[3] = {cb.hi[0] - cb.lo[0], cb.hi[1] - cb.lo[1], cb.hi[2] - cb.lo[2]};
            if (ext[1] > ext[axis]) axis = 1;
            if (ext[2] > ext[axis]) axis = 2;
            int* beg = prim.data() + first;
            mid = count / 2;
            std::nth_element(beg, beg + mid, beg + count, [&](int a, int b) {
                return pcent[3 * (size_t)a + axis] < pcent[3 * (size_t)b + axis];
            });
        } else if (best_axis >= 0) {
            float cmin = cb.lo[best_axis], cext = cb.hi[best_axis] - cb.lo[best_axis];
            float scale = (float)NB / cext;
            int* beg = prim.data() + first;
            int* m = std::partition(beg, beg + count, [&](int p) {
                int k = std::min(NB - 1, std::max(0, (int)((pcent[3 * (size_t)p + best_axis] - cmin) * scale)));
                return k <= best_bin;
            });
            mid = (int)(m - beg);
        } else {
            // spatial median on the widest centroid axis; falls back to an equal-count split
            int axis = 0;
            float ext[3] = {cb.hi[0] - cb.lo[0], cb.hi[1] - cb.lo[1], cb.hi[2] - cb.lo[2]};
            if (ext[1] > ext[axis]) axis = 1;
            if (ext[2] > ext[axis]) axis = 2;
            float split = 0.5f * (cb.lo[axis] + cb.hi[axis]);
            int* beg = prim.data() + first;
            int* m = std::partition(beg, beg + count, [&](int p) { return pcent[3 * (size_t)p + axis] < split; });
            mid = (int)(m - beg);
        }
        if (mid <= 0 || mid >= count) {
            if (count <= 3) {
                n.first = first;
                n.count = count;
                return;
            }
            mid = count / 2;
        }

        int l = alloc_node(), r = alloc_node();
        nodes[(size_t)ni].left = l;
        nodes[(size_t)ni].right = r;
        bool spawn = count > 32768 && live_threads.load() < max_threads;
        if (spawn) {
            live_threads.fetch_add(1);
            auto fut = std::async(std::launch::async, [this, l, first, mid, depth]() {
                build_range(l, first, mid, depth + 1);
                live_threads.fetch_sub(1);
            });
            build_range(r, first + mid, count - mid, depth + 1);
            fut.get();
        } else {
            build_range(l, first, mid, depth + 1);
            build_range(r, first + mid, count - mid, depth + 1);
        }
    }
};

inline float transform_coord(const float* m_row, float x, float y, float z) {
    // ((m0*x + m1*y) + m2*z) + m3, fp32, the order of RadeonRays' transform_point (mathutils.h:111-118)
    volatile float a = m_row[0] * x;
    volatile float b = m_row[1] * y;
    volatile float c = m_row[2] * z;
    volatile float s = a + b;
    s = s + c;
    s = s + m_row[3];
    return s;
}

struct WideNode {
    Box box;
    int child[8];   // >=0: Node2 index (internal or leaf); -1: empty
    int nchild = 0;
};

}  // namespace

static void build_core(Builder& B, const std::vector<float>& tri, const std::vector<int32_t>& tri_mesh_in, size_t N,
                       const BuildOptions& opt, Bvh8& out);

void build_bvh8(const std::vector<BuildInputMesh>& meshes, const BuildOptions& opt, Bvh8& out) {
    // ---- 1. gather world-space triangles
    int64_t total = 0;
    for (const auto& m : meshes) total += m.numfaces;
    if (total <= 0) throw std::runtime_error("Scene is empty.");
    if (total > (int64_t)0x3fffffff) throw std::runtime_error("too many triangles");
    const size_t N = (size_t)total;
    std::vector<float> tri(12 * N);
    std::vector<int32_t> tri_mesh_in(N);
    Builder B(opt);
    B.pbox.resize(N);
    B.pcent.resize(3 * N);
    B.prim.resize(N);
    size_t t = 0;
    int32_t mesh_index = -1;
    for (const auto& m : meshes) {
        ++mesh_index;
        const int vstride = m.vstride ? m.vstride : 12;
        const int istride = m.istride ? m.istride : 12;
        for (int f = 0; f < m.numfaces; ++f, ++t) {
            const int* id = (const int*)((const char*)m.indices + (size_t)f * istride);
            float* o = &tri[12 * t];
            Box b;
            b.reset();
            for (int k = 0; k < 3; ++k) {
                if (id[k] < 0 || id[k] >= m.vnum) throw std::runtime_error("vertex index out of range");
                const float* v = (const float*)((const char*)m.vertices + (size_t)id[k] * vstride);
                o[4 * k + 0] = transform_coord(m.m + 0, v[0], v[1], v[2]);
                o[4 * k + 1] = transform_coord(m.m + 4, v[0], v[1], v[2]);
                o[4 * k + 2] = transform_coord(m.m + 8, v[0], v[1], v[2]);
                if (!std::isfinite(o[4 * k + 0]) || !std::isfinite(o[4 * k + 1]) || !std::isfinite(o[4 * k + 2]))
                    throw std::runtime_error("non-finite vertex position");
                b.grow(o + 4 * k);
            }
            int sid = m.shape_id, pid = f, zero = 0;
            std::memcpy(o + 3, &sid, 4);
            std::memcpy(o + 7, &pid, 4);
            std::memcpy(o + 11, &zero, 4);
            B.pbox[t] = b;
            for (int a = 0; a < 3; ++a) B.pcent[3 * t + a] = 0.5f * (b.lo[a] + b.hi[a]);
            B.prim[t] = (int)t;
            tri_mesh_in[t] = mesh_index;
        }
    }

    build_core(B, tri, tri_mesh_in, N, opt, out);
}

// Top level of the two-level structure: a tree over `n` boxes (lo.xyz, hi.xyz each).  The leaf "triangle" records of
// the result carry, in the int bits of word 3, the index of the box they stand for, in emission order: the caller
// lays its instance table out in that order, so a leaf's record index IS the instance index.
void build_bvh8_boxes(const float* boxes, int n, const BuildOptions& opt, Bvh8& out) {
    if (n <= 0) throw std::runtime_error("Scene is empty.");
    const size_t N = (size_t)n;
    std::vector<float> rec(12 * N, 0.0f);
    std::vector<int32_t> owner(N);
    Builder B(opt);
    B.pbox.resize(N);
    B.pcent.resize(3 * N);
    B.prim.resize(N);
    for (size_t i = 0; i < N; ++i) {
        Box b;
        for (int a = 0; a < 3; ++a) {
            b.lo[a] = boxes[6 * i + a];
            b.hi[a] = boxes[6 * i + 3 + a];
            if (!std::isfinite(b.lo[a]) || !std::isfinite(b.hi[a])) throw std::runtime_error("non-finite instance bounds");
        }
        B.pbox[i] = b;
        for (int a = 0; a < 3; ++a) B.pcent[3 * i + a] = 0.5f * (b.lo[a] + b.hi[a]);
        B.prim[i] = (int)i;
        const int idx = (int)i;
        std::memcpy(&rec[12 * i + 3], &idx, 4);
        owner[i] = idx;
    }
    build_core(B, rec, owner, N, opt, out);
}

static void build_core(Builder& B, const std::vector<float>& tri, const std::vector<int32_t>& tri_mesh_in, size_t N,
                       const BuildOptions& opt, Bvh8& out) {
    // The traversal keeps at most two stack entries per wide level (a node group and a postponed triangle group) on a
    // 40-entry stack (traverse.cuh), so a tree may have at most kMaxWideLevels levels -- the bound the device builder
    // enforces too (lbvh_build.h).  A SAH tree over strongly non-uniform input can be deeper: it is then rebuilt with
    // equal-count splits below a binary depth that is lowered until the wide tree fits (0 = balanced everywhere, which
    // always fits: ceil(log2(2^30 / 2) / 3) + 1 = 11 levels).
    const int kMaxWideLevels = opt.max_levels;
    static const int caps[] = {std::numeric_limits<int>::max(), 40, 24, 12, 0};
    for (int attempt = 0;; ++attempt) {
    B.sah_depth_cap = caps[attempt];
    B.next_node.store(0);
    for (size_t i = 0; i < N; ++i) B.prim[i] = (int)i;
    // ---- 2. binary build
    B.nodes.assign(2 * N, Node2());
    B.max_threads = opt.num_threads > 0 ? opt.num_threads : (int)std::max(1u, std::thread::hardware_concurrency());
    int root = B.alloc_node();
    B.build_range(root, 0, (int)N, 0);
    const std::vector<Node2>& n2 = B.nodes;

    // ---- 3/4. collapse to 8-wide, breadth-first, emitting nodes and triangles
    std::vector<WideNode> wide;
    std::vector<int> wide_src;  // Node2 index each wide node was made from
    wide.reserve(N / 2 + 16);
    wide_src.push_back(root);
    wide.emplace_back();
    std::vector<int> wide_depth(1, 0);
    out.nodes.clear();
    out.triangles.clear();
    out.tri_mesh.clear();
    out.tri_mesh.reserve(N);
    out.triangles.reserve(12 * N);
    std::vector<uint32_t>& W = out.nodes;

    for (size_t wi = 0; wi < wide_src.size(); ++wi) {
        const int src = wide_src[wi];
        int ch[8];
        int nch = 0;
        if (n2[(size_t)src].count > 0) {
            ch[nch++] = src;  // degenerate: the whole scene is one leaf
        } else {
            ch[nch++] = n2[(size_t)src].left;
            ch[nch++] = n2[(size_t)src].right;
            while (nch < 8) {
                int best = -1;
                float best_area = -1.f;
                for (int i = 0; i < nch; ++i) {
                    const Node2& c = n2[(size_t)ch[i]];
                    if (c.count == 0 && c.box.half_area() > best_area) {
                        best_area = c.box.half_area();
                        best = i;
                    }
                }
                if (best < 0) break;
                int c = ch[best];
                ch[best] = n2[(size_t)c].left;
                ch[nch++] = n2[(size_t)c].right;
            }
        }
        const Box& nb = n2[(size_t)src].box;

        // slot assignment: slot s is visited first by rays whose direction signs are s
        // (bit a set = negative along axis a), so it should hold the child that is nearest
        // along that direction.  Greedy maximum of the remaining (child, slot) scores.
        int slot_of[8];
        int child_in_slot[8];
        for (int i = 0; i < 8; ++i) {
            slot_of[i] = -1;
            child_in_slot[i] = -1;
        }
        float pc[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
        float score[8][8];
        for (int i = 0; i < nch; ++i) {
            const Box& cbx = n2[(size_t)ch[i]].box;
            float d[3] = {0.5f * (cbx.lo[0] + cbx.hi[0]) - pc[0], 0.5f * (cbx.lo[1] + cbx.hi[1]) - pc[1],
                          0.5f * (cbx.lo[2] + cbx.hi[2]) - pc[2]};
            for (int s = 0; s < 8; ++s) {
                float proj = ((s & 1) ? -d[0] : d[0]) + ((s & 2) ? -d[1] : d[1]) + ((s & 4) ? -d[2] : d[2]);
                score[i][s] = -proj;  // nearer along the slot's direction = better
            }
        }
        for (int k = 0; k < nch; ++k) {
            int bi = -1, bs = -1;
            float bv = -std::numeric_limits<float>::infinity();
            for (int i = 0; i < nch; ++i) {
                if (slot_of[i] >= 0) continue;
                for (int s = 0; s < 8; ++s) {
                    if (child_in_slot[s] >= 0) continue;
                    if (score[i][s] > bv) {
                        bv = score[i][s];
                        bi = i;
                        bs = s;
                    }
                }
            }
            if (bi < 0) {      // NaN scores (non-finite boxes): first free child into the first free slot
                for (int i = 0; i < nch && bi < 0; ++i)
                    if (slot_of[i] < 0) bi = i;
                for (int s = 0; s < 8 && bs < 0; ++s)
                    if (child_in_slot[s] < 0) bs = s;
            }
            slot_of[bi] = bs;
            child_in_slot[bs] = bi;
        }

        // quantisation frame
        uint32_t words[20];
        std::memset(words, 0, sizeof(words));
        int e[3];
        float scale[3];
        for (int a = 0; a < 3; ++a) {
            double ext = (double)nb.hi[a] - (double)nb.lo[a];
            int ex = -100;
            if (ext > 0.0) {
                ex = (int)std::ceil(std::log2(ext / 255.0));
                // make sure the largest child plane quantises to <= 255 after rounding up
                while (std::ceil(ext / std::ldexp(1.0, ex)) > 255.0) ++ex;
            }
            ex = std::max(-126, std::min(127, ex));
            e[a] = ex;
            scale[a] = std::ldexp(1.0f, ex);
        }
        uint8_t imask = 0;
        uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint8_t q[6][8];
        std::memset(q, 0, sizeof(q));
        const uint32_t child_base = (uint32_t)wide_src.size();
        const uint32_t tri_base = (uint32_t)(out.triangles.size() / 12);
        int tri_off = 0;
        for (int s = 0; s < 8; ++s) {
            int i = child_in_slot[s];
            if (i < 0) continue;
            const Node2& c = n2[(size_t)ch[i]];
            for (int a = 0; a < 3; ++a) {
                double lo = std::floor(((double)c.box.lo[a] - (double)nb.lo[a]) / (double)scale[a]);
                double hi = std::ceil(((double)c.box.hi[a] - (double)nb.lo[a]) / (double)scale[a]);
                // guard fp32 re-expansion: nudge outwards until the decoded planes enclose the box
                int qlo = (int)std::max(0.0, std::min(255.0, lo));
                int qhi = (int)std::max(0.0, std::min(255.0, hi));
                while (qlo > 0 && nb.lo[a] + (float)qlo * scale[a] > c.box.lo[a]) --qlo;
                while (qhi < 255 && nb.lo[a] + (float)qhi * scale[a] < c.box.hi[a]) ++qhi;
                q[a][s] = (uint8_t)qlo;
                q[3 + a][s] = (uint8_t)qhi;
            }
            if (c.count == 0) {
                imask |= (uint8_t)(1u << s);
                meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
                wide_src.push_back(ch[i]);
                wide_depth.push_back(wide_depth[wi] + 1);
            } else {
                static const uint8_t unary[4] = {0, 1, 3, 7};
                meta[s] = (uint8_t)((unary[c.count] << 5) | (uint32_t)tri_off);
                for (int k = 0; k < c.count; ++k) {
                    const float* src_tri = &tri[12 * (size_t)B.prim[(size_t)(c.first + k)]];
                    out.triangles.insert(out.triangles.end(), src_tri, src_tri + 12);
                    out.tri_mesh.push_back(tri_mesh_in[(size_t)B.prim[(size_t)(c.first + k)]]);
                }
                tri_off += c.count;
            }
        }
        std::memcpy(&words[0], &nb.lo[0], 4);
        std::memcpy(&words[1], &nb.lo[1], 4);
        std::memcpy(&words[2], &nb.lo[2], 4);
        words[3] = (uint32_t)(uint8_t)(e[0] + 127) | ((uint32_t)(uint8_t)(e[1] + 127) << 8) |
                   ((uint32_t)(uint8_t)(e[2] + 127) << 16) | ((uint32_t)imask << 24);
        words[4] = child_base;
        words[5] = tri_base;
        std::memcpy(&words[6], meta, 8);
        std::memcpy(&words[8], q[0], 8);    // qlo_x
        std::memcpy(&words[10], q[1], 8);   // qlo_y
        std::memcpy(&words[12], q[2], 8);   // qlo_z
        std::memcpy(&words[14], q[3], 8);   // qhi_x
        std::memcpy(&words[16], q[4], 8);   // qhi_y
        std::memcpy(&words[18], q[5], 8);   // qhi_z
        W.insert(W.end(), words, words + 20);
    }
    out.num_nodes = (int64_t)wide_src.size();
    out.num_tris = (int64_t)(out.triangles.size() / 12);
    out.level_offsets.clear();      // breadth-first emission: depths are non-decreasing
    for (size_t i = 0; i < wide_depth.size(); ++i)
        if (i == 0 || wide_depth[i] != wide_depth[i - 1]) out.level_offsets.push_back((int32_t)i);
    out.level_offsets.push_back((int32_t)wide_depth.size());
    for (int a = 0; a < 3; ++a) {
        out.scene_min[a] = n2[(size_t)root].box.lo[a];
        out.scene_max[a] = n2[(size_t)root].box.hi[a];
    }
    if (out.num_tris != (int64_t)N) throw std::runtime_error("internal error: triangle count mismatch");
    if ((int)out.level_offsets.size() - 1 <= kMaxWideLevels || !opt.depth_guard) break;
    if (caps[attempt] == 0) throw std::runtime_error("internal error: balanced tree deeper than the traversal stack");
    }
}

}  // namespace mcrt
