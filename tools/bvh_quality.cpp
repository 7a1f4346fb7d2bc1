// bvh_quality.cpp -- host-side measurement of the 8-wide tree's quality: walks a set of rays through a tree built by
// csrc/bvh8_build.cpp with the control flow of csrc/traverse.cuh (octant slot order, postponed triangle groups,
// closest-hit shrinking) and reports node visits and triangle tests per ray.  A development tool: nothing in the
// product or the tests uses it.  Input: three raw files (nodes u32[20*N], triangles f32[12*T], rays f32[8*R]:
// o.xyz tmax d.xyz pad).  g++ -O2 -o /tmp/bvh_quality tools/bvh_quality.cpp
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static std::vector<char> slurp(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) { perror(path); exit(1); }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> b((size_t)n);
    if (fread(b.data(), 1, (size_t)n, f) != (size_t)n) exit(1);
    fclose(f);
    return b;
}

static inline float safe_inv(float d) {
    if (std::fabs(d) <= 1e-20f) d = std::copysign(1e-20f, d);
    return 1.0f / d;
}

static inline bool tri_hit(const float* r, const float* o, const float* d, float tmax, float* tout) {
    float e1[3] = {r[4] - r[0], r[5] - r[1], r[6] - r[2]}, e2[3] = {r[8] - r[0], r[9] - r[1], r[10] - r[2]};
    float s1[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
    float det = s1[0] * e1[0] + s1[1] * e1[1] + s1[2] * e1[2];
    if (det == 0.f) return false;
    float inv = 1.0f / det;
    float dd[3] = {o[0] - r[0], o[1] - r[1], o[2] - r[2]};
    float b1 = (dd[0] * s1[0] + dd[1] * s1[1] + dd[2] * s1[2]) * inv;
    if (b1 < 0.f || b1 > 1.f) return false;
    float s2[3] = {dd[1] * e1[2] - dd[2] * e1[1], dd[2] * e1[0] - dd[0] * e1[2], dd[0] * e1[1] - dd[1] * e1[0]};
    float b2 = (d[0] * s2[0] + d[1] * s2[1] + d[2] * s2[2]) * inv;
    if (b2 < 0.f || b1 + b2 > 1.f) return false;
    float t = (e2[0] * s2[0] + e2[1] * s2[1] + e2[2] * s2[2]) * inv;
    if (t < 0.f || t > tmax) return false;
    *tout = t;
    return true;
}

// The bound on what traversal ORDER can give: children visited nearest first (sorted by entry distance) and
// re-tested against the current closest hit when popped.
static void ideal_walk(const uint32_t* nodes, const float* tris, const float* o, const float* d, const float* inv,
                       float& closest, size_t& visited, size_t& tested, bool cull) {
    struct E { uint32_t node; float t; };
    E stack[256];
    int sp = 0;
    stack[sp++] = {0u, 0.f};
    while (sp) {
        E e = stack[--sp];
        if (cull && e.t > closest) continue;
        const uint32_t* node = nodes + 20 * (size_t)e.node;
        ++visited;
        float p[3];
        std::memcpy(p, node, 12);
        const uint32_t imask = node[3] >> 24;
        const uint8_t* meta = (const uint8_t*)(node + 6);
        const uint8_t* q = meta + 8;
        float sc[3];
        for (int a = 0; a < 3; ++a) {
            uint32_t ex = ((node[3] >> (8 * a)) & 0xFFu) << 23;
            std::memcpy(&sc[a], &ex, 4);
        }
        E kids[8];
        int nk = 0;
        for (int s = 0; s < 8; ++s) {
            const uint32_t m = meta[s];
            if (!m) continue;
            float tmin = 0.f, tmx = closest;
            for (int a = 0; a < 3; ++a) {
                float lo = p[a] + (float)q[8 * a + s] * sc[a], hi = p[a] + (float)q[8 * (3 + a) + s] * sc[a];
                float ta = (lo - o[a]) * inv[a], tb2 = (hi - o[a]) * inv[a];
                if (inv[a] < 0) { float t = ta; ta = tb2; tb2 = t; }
                tmin = std::fmax(tmin, ta);
                tmx = std::fmin(tmx, tb2);
            }
            if (!(tmin <= tmx * 1.0000004f)) continue;
            if ((imask >> s) & 1u) {
                kids[nk++] = {node[4] + (uint32_t)__builtin_popcount(imask & ((1u << s) - 1u)), tmin};
            } else {
                const int cnt = (m >> 5) == 1 ? 1 : (m >> 5) == 3 ? 2 : 3;
                for (int k = 0; k < cnt; ++k) {
                    ++tested;
                    float t;
                    if (tri_hit(tris + 12 * (size_t)(node[5] + (m & 31u) + k), o, d, closest, &t)) closest = t;
                }
            }
        }
        for (int i = 1; i < nk; ++i)
            for (int j = i; j > 0 && kids[j].t > kids[j - 1].t; --j) { E t = kids[j]; kids[j] = kids[j - 1]; kids[j - 1] = t; }
        for (int i = 0; i < nk; ++i) stack[sp++] = kids[i];   // farthest first, nearest on top
    }
}

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: bvh_quality nodes tris rays\n"); return 2; }
    auto nb = slurp(argv[1]), tb = slurp(argv[2]), rb = slurp(argv[3]);
    const uint32_t* nodes = (const uint32_t*)nb.data();
    const float* tris = (const float*)tb.data();
    const float* rays = (const float*)rb.data();
    const size_t R = rb.size() / 32;
    double sum_nodes = 0, sum_tris = 0, sum_groups = 0;
    size_t nhit = 0, max_nodes = 0, max_stack = 0;
    for (size_t ri = 0; ri < R; ++ri) {
        const float* o = rays + 8 * ri;
        const float* d = o + 4;
        float inv[3] = {safe_inv(d[0]), safe_inv(d[1]), safe_inv(d[2])};
        const uint32_t octant = (inv[0] < 0 ? 1u : 0u) | (inv[1] < 0 ? 2u : 0u) | (inv[2] < 0 ? 4u : 0u);
        const uint32_t octinv = 7u - octant;
        float closest = o[3];
        bool hit = false;
        size_t visited = 0, tested = 0;
        if (argc > 4 && (!std::strcmp(argv[4], "ideal") || !std::strcmp(argv[4], "sorted"))) {
            ideal_walk(nodes, tris, o, d, inv, closest, visited, tested, argv[4][0] == 'i');
            sum_nodes += (double)visited;
            sum_tris += (double)tested;
            nhit += closest < o[3];
            continue;
        }
        uint32_t stack[2 * 64];
        float gstack[64];
        float cstack[64][8], ct[8] = {0, 0, 0, 0, 0, 0, 0, 0}, nt[8];
        const bool ccull = argc > 4 && !std::strcmp(argv[4], "childcull");
        const bool gcull = argc > 4 && !std::strncmp(argv[4], "groupcull", 9);
        float gmin = 0.f;
        bool taken = false;
        int sp = 0;
        uint32_t ng0 = 0, ng1 = 0x80000000u;
        for (;;) {
            uint32_t tg0 = 0, tg1 = 0;
            if (gcull && (taken || argv[4][9] == 0) && ng1 > 0x00FFFFFFu && gmin > closest) ng1 &= 0x00FFFFFFu;
            if (ng1 > 0x00FFFFFFu) {
                const uint32_t hits = ng1;
                const int bit = 31 - __builtin_clz(hits);
                taken = true;
                ng1 &= ~(1u << bit);
                if (ccull && ct[bit - 24] > closest) { if (ng1 <= 0x00FFFFFFu) { if (!sp) break; --sp; ng0 = stack[2 * sp]; ng1 = stack[2 * sp + 1]; gmin = gstack[sp]; std::memcpy(ct, cstack[sp], 32); } continue; }
                if (ng1 > 0x00FFFFFFu) { stack[2 * sp] = ng0; stack[2 * sp + 1] = ng1; gstack[sp] = gmin; std::memcpy(cstack[sp], ct, 32); ++sp; if ((size_t)sp > max_stack) max_stack = sp; }
                const uint32_t slot = (uint32_t)(bit - 24) ^ octinv;
                const uint32_t rel = __builtin_popcount(hits & 0xFFu & ((1u << slot) - 1u));
                const uint32_t* node = nodes + 20 * (size_t)(ng0 + rel);
                ++visited;
                float p[3];
                std::memcpy(p, node, 12);
                const uint32_t imask = node[3] >> 24;
                const uint8_t* b = (const uint8_t*)(node + 6);
                const uint8_t* meta = b;
                const uint8_t* q = b + 8;
                float sc[3];
                for (int a = 0; a < 3; ++a) {
                    uint32_t e = ((node[3] >> (8 * a)) & 0xFFu) << 23;
                    std::memcpy(&sc[a], &e, 4);
                }
                uint32_t hitmask = 0;
                float newmin = 3e38f;
                for (int s = 0; s < 8; ++s) {
                    const uint32_t m = meta[s];
                    if (!m) continue;
                    float tmin = 0.f, tmx = closest;
                    for (int a = 0; a < 3; ++a) {
                        float lo = p[a] + (float)q[8 * a + s] * sc[a], hi = p[a] + (float)q[8 * (3 + a) + s] * sc[a];
                        float ta = (lo - o[a]) * inv[a], tb2 = (hi - o[a]) * inv[a];
                        if (inv[a] < 0) { float t = ta; ta = tb2; tb2 = t; }
                        tmin = std::fmax(tmin, ta);
                        tmx = std::fmin(tmx, tb2);
                    }
                    if (tmin <= tmx * 1.0000004f) {
                        const uint32_t inner = (imask >> s) & 1u;
                        if (inner) newmin = std::fmin(newmin, tmin);
                        const uint32_t bi = ((m & 31u) ^ (inner ? octinv : 0u)) & 31u;
                        hitmask |= (m >> 5) << bi;
                        if (inner) nt[bi - 24] = tmin;
                    }
                }
                ng0 = node[4];
                gmin = newmin;
                taken = false;
                if (argc > 4 && !std::strcmp(argv[4], "groupcull3")) {
                    // the cheapest conservative form: only the child in the ray's own octant slot (bit 31, visited first
                    // whenever it is hit) is left out, and the float is truncated to its upper 16 bits (it would travel
                    // in the 16 spare bits of the stack entry's second word)
                    const uint32_t inner_hits = hitmask >> 24;
                    gmin = 3e38f;
                    for (int b2 = 0; b2 < 7; ++b2)
                        if ((inner_hits >> b2) & 1u) gmin = std::fmin(gmin, nt[b2]);
                    uint32_t gb;
                    std::memcpy(&gb, &gmin, 4);
                    gb &= 0xFFFF0000u;
                    std::memcpy(&gmin, &gb, 4);
                }
                if (argc > 4 && !std::strcmp(argv[4], "groupcull2")) {
                    const uint32_t inner_hits = hitmask >> 24;
                    gmin = 3e38f;
                    if (inner_hits) {
                        const int top = 31 - __builtin_clz(inner_hits);
                        for (int b2 = 0; b2 < top; ++b2)
                            if ((inner_hits >> b2) & 1u) gmin = std::fmin(gmin, nt[b2]);
                    }
                }
                std::memcpy(ct, nt, 32);
                ng1 = (hitmask & 0xFF000000u) | imask;
                tg0 = node[5];
                tg1 = hitmask & 0x00FFFFFFu;
                if (tg1) sum_groups += 1;
            }
            while (tg1) {
                const int ti = 31 - __builtin_clz(tg1);
                tg1 &= ~(1u << ti);
                ++tested;
                float t;
                if (tri_hit(tris + 12 * (size_t)(tg0 + ti), o, d, closest, &t)) { closest = t; hit = true; }
            }
            if (ng1 <= 0x00FFFFFFu) {
                if (!sp) break;
                --sp;
                ng0 = stack[2 * sp];
                ng1 = stack[2 * sp + 1];
                gmin = gstack[sp];
                taken = true;
                std::memcpy(ct, cstack[sp], 32);
            }
        }
        sum_nodes += (double)visited;
        sum_tris += (double)tested;
        if (visited > max_nodes) max_nodes = visited;
        nhit += hit;
    }
    const double n = sum_nodes / (double)R, t = sum_tris / (double)R;
    printf("{\"rays\": %zu, \"hit_frac\": %.4f, \"nodes_per_ray\": %.3f, \"tris_per_ray\": %.3f, \"bytes_per_ray\": %.1f, "
           "\"max_nodes\": %zu, \"max_stack\": %zu, \"num_nodes\": %zu, \"num_tris\": %zu}\n",
           R, (double)nhit / (double)R, n, t, 80.0 * n + 48.0 * t, max_nodes, max_stack, nb.size() / 80, tb.size() / 48);
    return 0;
}
