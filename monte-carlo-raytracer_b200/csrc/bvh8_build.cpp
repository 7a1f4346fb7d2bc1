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
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <limits>
#include <stdexcept>
#include <thread>
#include <chrono>

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
    int pfirst = 0, pcount = 0; // primitive range of the whole subtree (every node)
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
        n.pfirst = first;
        n.pcount = count;
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
            if (count <= 3 && !opt.optimal_collapse) {
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
            // Leaf decision (leaves hold at most 3 triangles, the wide node's per-child maximum).  With the optimal
            // collapse the binary tree goes down to single triangles and the collapse forms the leaves.
            if (count <= 3 && !opt.optimal_collapse) {
                float leaf_cost = (float)count * b.half_area();
                float split_cost = opt.traversal_cost * b.half_area() + best_cost;
                if (best_axis < 0 || leaf_cost <= split_cost) {
                    n.first = first;
                    n.count = count;
                    return;
                }
            }
        } else if (count <= 2 && !opt.optimal_collapse) {
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
            if (count <= 3 && !opt.optimal_collapse) {
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

// development switches (see bvh8.h)
BuildOptions with_env_overrides(BuildOptions o) {
    if (const char* c = std::getenv("MCRT_BVH_COLLAPSE")) o.optimal_collapse = std::strcmp(c, "greedy") != 0;
    if (const char* c = std::getenv("MCRT_BVH_SLOTS")) o.optimal_slots = std::strcmp(c, "greedy") != 0;
    if (const char* c = std::getenv("MCRT_BVH_BINS")) o.num_bins = std::atoi(c);
    if (const char* c = std::getenv("MCRT_BVH_COST")) {
        float a = 0.f, b = 0.f;
        if (std::sscanf(c, "%f:%f", &a, &b) == 2 && a > 0.f && b > 0.f) {
            o.collapse_node_cost = a;
            o.collapse_prim_cost = b;
        }
    }
    return o;
}

}  // namespace

static void build_core(Builder& B, const std::vector<float>& tri, const std::vector<int32_t>& tri_mesh_in, size_t N,
                       const BuildOptions& opt, Bvh8& out);

void build_bvh8(const std::vector<BuildInputMesh>& meshes, const BuildOptions& opt_in, Bvh8& out) {
    const BuildOptions opt = with_env_overrides(opt_in);
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
void build_bvh8_boxes(const float* boxes, int n, const BuildOptions& opt_in, Bvh8& out) {
    const BuildOptions opt = with_env_overrides(opt_in);
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
    auto T0 = std::chrono::steady_clock::now();
    int root = B.alloc_node();
    B.build_range(root, 0, (int)N, 0);
    auto T1 = std::chrono::steady_clock::now();
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

    // Optimal collapse (Ylitie, Karras, Laine 2017, section 3.1): cost[n][i] = the smallest SAH cost of representing
    // binary subtree n by at most i children of one wide node, i = 1..7; a subtree of at most 3 triangles may become
    // one leaf child, any subtree may become an internal child whose own 8 slots are distributed over its two sides.
    // Children have larger indices than their parents (alloc_node), so one descending sweep is a post-order.
    struct Dp {
        float c[8];      // c[i], i = 1..7
        uint8_t k[9];    // k[i], i = 2..7: 0 = same forest as i - 1, else slots given to the left side; k[8]: root split
        uint8_t leaf;    // c[1] is the leaf alternative
    };
    std::vector<Dp> dp;
    const int num_n2 = B.next_node.load();
    if (opt.optimal_collapse) {
        dp.resize((size_t)num_n2);
        const float c_node = opt.collapse_node_cost, c_prim = opt.collapse_prim_cost;
        for (int ni = num_n2 - 1; ni >= 0; --ni) {
            const Node2& n = n2[(size_t)ni];
            Dp& d = dp[(size_t)ni];
            const float area = n.box.half_area();
            std::memset(d.k, 0, sizeof(d.k));
            if (n.count > 0) {
                for (int i = 1; i < 8; ++i) d.c[i] = area * (float)n.pcount * c_prim;
                d.leaf = 1;
                continue;
            }
            const Dp& l = dp[(size_t)n.left];
            const Dp& r = dp[(size_t)n.right];
            float dist[9];
            uint8_t dk[9];
            for (int j = 2; j <= 8; ++j) {
                dist[j] = std::numeric_limits<float>::infinity();
                dk[j] = 1;
                for (int k = std::max(1, j - 7); k <= std::min(7, j - 1); ++k) {
                    const float c = l.c[k] + r.c[j - k];
                    if (c < dist[j]) {
                        dist[j] = c;
                        dk[j] = (uint8_t)k;
                    }
                }
            }
            const float c_int = dist[8] + area * c_node;
            const float c_leaf = n.pcount <= 3 ? area * (float)n.pcount * c_prim : std::numeric_limits<float>::infinity();
            d.leaf = c_leaf <= c_int;
            d.c[1] = d.leaf ? c_leaf : c_int;
            d.k[8] = dk[8];
            for (int i = 2; i < 8; ++i) {
                if (dist[i] < d.c[i - 1]) {
                    d.c[i] = dist[i];
                    d.k[i] = dk[i];
                } else {
                    d.c[i] = d.c[i - 1];
                    d.k[i] = 0;
                }
            }
        }
    }
    auto T2 = std::chrono::steady_clock::now();
    if (std::getenv("MCRT_BUILD_TRACE")) fprintf(stderr, "binary %.3f s, dp %.3f s\n", std::chrono::duration<double>(T1 - T0).count(), std::chrono::duration<double>(T2 - T1).count());
    std::vector<uint8_t> leaf_child((size_t)num_n2, 0);   // binary node emitted as ONE leaf child (its whole range)
    for (int ni = 0; ni < num_n2; ++ni) leaf_child[(size_t)ni] = n2[(size_t)ni].count > 0;
    std::vector<std::pair<int, int>> todo;                // (binary node, slots) during the expansion of one wide node

    for (size_t wi = 0; wi < wide_src.size(); ++wi) {
        const int src = wide_src[wi];
        int ch[8];
        int nch = 0;
        if (n2[(size_t)src].count > 0 || (opt.optimal_collapse && wi == 0 && n2[(size_t)src].pcount <= 3)) {
            ch[nch++] = src;  // degenerate: the whole scene is one leaf
            leaf_child[(size_t)src] = 1;
        } else if (opt.optimal_collapse) {
            todo.clear();
            const int k8 = dp[(size_t)src].k[8];
            todo.emplace_back(n2[(size_t)src].right, 8 - k8);
            todo.emplace_back(n2[(size_t)src].left, k8);
            while (!todo.empty()) {
                int m = todo.back().first, j = todo.back().second;
                todo.pop_back();
                const Dp& d = dp[(size_t)m];
                while (j > 1 && d.k[j] == 0) --j;
                if (j == 1 || n2[(size_t)m].count > 0) {
                    leaf_child[(size_t)m] = d.leaf;
                    ch[nch++] = m;
                } else {
                    todo.emplace_back(n2[(size_t)m].right, j - d.k[j]);
                    todo.emplace_back(n2[(size_t)m].left, (int)d.k[j]);
                }
            }
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
        // the assignment with the largest total score (subset DP over the 8 slots: child i takes a free slot of every
        // set of i occupied slots); the greedy loop below then only fills in what a NaN score left open
        if (opt.optimal_slots) {
            // pull form, branch-free: best[M] for a set M of i + 1 occupied slots = max over its members sl of
            // best[M without sl] + score[i][sl] (child i is the last one placed); sets are visited in increasing order,
            // and M without sl < M
            float best[256];
            int8_t from[256];
            const float ninf = -std::numeric_limits<float>::infinity();
            best[0] = 0.f;
            from[0] = -1;
            for (int m = 1; m < 256; ++m) {
                const int i = __builtin_popcount((unsigned)m) - 1;
                float bv = ninf;
                int bsl = -1;
                if (i < nch) {
                    for (int sl = 0; sl < 8; ++sl) {
                        const bool member = (m >> sl) & 1;
                        const float v = member ? best[m & ~(1 << sl)] + score[i][sl] : ninf;
                        const bool better = v > bv;
                        bv = better ? v : bv;
                        bsl = better ? sl : bsl;
                    }
                }
                best[m] = bv;
                from[m] = (int8_t)bsl;
            }
            int bm = -1;
            for (int m = 0; m < 256; ++m)
                if (__builtin_popcount((unsigned)m) == nch && from[m] >= 0 && (bm < 0 || best[m] > best[bm])) bm = m;
            for (int i = nch - 1; i >= 0 && bm > 0; --i) {
                const int sl = from[bm];
                if (sl < 0) break;
                slot_of[i] = sl;
                child_in_slot[sl] = i;
                bm &= ~(1 << sl);
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
                if (bi < 0) break;   // every child has its slot
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
            if (!leaf_child[(size_t)ch[i]]) {
                imask |= (uint8_t)(1u << s);
                meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
                wide_src.push_back(ch[i]);
                wide_depth.push_back(wide_depth[wi] + 1);
            } else {
                static const uint8_t unary[4] = {0, 1, 3, 7};
                meta[s] = (uint8_t)((unary[c.pcount] << 5) | (uint32_t)tri_off);
                for (int k = 0; k < c.pcount; ++k) {
                    const float* src_tri = &tri[12 * (size_t)B.prim[(size_t)(c.pfirst + k)]];
                    out.triangles.insert(out.triangles.end(), src_tri, src_tri + 12);
                    out.tri_mesh.push_back(tri_mesh_in[(size_t)B.prim[(size_t)(c.pfirst + k)]]);
                }
                tri_off += c.pcount;
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
    if (std::getenv("MCRT_BUILD_TRACE")) fprintf(stderr, "emit %.3f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - T2).count());
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
