// rrb_api.cu -- implementation of include/rrb.h: the object model behind the C ABI.
//
// Mirrors IntersectionApiImpl / World / Mesh / Instance / Options
// (RadeonRays/src/api/radeon_rays_impl.cpp:43-191, src/world/world.cpp:28-56,
// src/primitive/mesh.cpp:31-113, src/util/options.cpp) with CUDA streams / events / cudaMalloc
// in the role of Calc (RadeonRays/Calc).  No CPU fallback: every query needs the CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cfloat>
#include <cmath>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/rrb.h"
#include "bvh8.h"
#include "lbvh_build.h"
#include "rrb_kernels.h"

namespace {
thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* where) {
    return fail(RRB_ERROR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define RRB_CUDA(call)                                        \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

const float kIdentity[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
}  // namespace

struct rrb_shape {
    bool is_instance = false;
    const rrb_shape* base = nullptr;          // instances only
    std::vector<float> vertices;              // xyz tightly packed (meshes only)
    std::vector<int> indices;                 // 3 per face
    float m[16], minv[16];
    float linear_velocity[3] = {0, 0, 0};
    float angular_velocity[4] = {0, 0, 0, 1};
    int id = 0;
    bool changed = true;                      // ShapeImpl::statechange_
    rrb_shape() {
        std::memcpy(m, kIdentity, sizeof(m));
        std::memcpy(minv, kIdentity, sizeof(minv));
    }
};

struct rrb_buffer {
    void* dptr = nullptr;
    size_t size = 0;
    bool owning = false;
    void* host_map = nullptr;   // pinned staging while mapped
    size_t map_offset = 0, map_size = 0;
    int map_type = 0;
};

struct rrb_event {
    cudaEvent_t ev = nullptr;
    bool recorded = false;
};

struct rrb_api {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::vector<const rrb_shape*> world;      // World::shapes_, attach order
    bool world_changed = true;
    std::map<std::string, std::string> str_options;
    std::map<std::string, float> f32_options;
    int next_id = 1;                          // IntersectionApiImpl::m_nextid, ResetIdCounter -> 1
    // committed acceleration structure
    void* d_nodes = nullptr;
    void* d_tris = nullptr;
    int64_t num_nodes = 0, num_tris = 0;
    double build_seconds = 0.0;
    int64_t launches = 0;
    // device refit state of the last full build (option "bvh.refit")
    std::vector<const rrb_shape*> built_world;     // shapes of the last full build, in build order
    std::vector<int> built_mesh_of_shape;          // world index -> mesh index of the build (-1: contributed nothing)
    std::vector<int32_t> level_offsets;
    std::vector<mcrt::RefitMesh> refit_meshes;
    float* d_ref_vertices = nullptr;
    int* d_ref_indices = nullptr;
    int* d_tri_mesh = nullptr;
    float* d_boxes = nullptr;
    mcrt::RefitMesh* d_refit_meshes = nullptr;
    int64_t refits = 0;
    bool last_commit_was_refit = false;
    bool last_build_on_device = false;                 // acc.type = hlbvh / bvh.builder = lbvh (lbvh_build.cu)
    std::string device_build_fallback_reason;
    std::vector<int> pool_face_prefix;                 // faces before each pooled mesh, total last
    // two-level structure (instances in the world, or option "bvh.force2level"): one bottom-level tree per distinct
    // base mesh in OBJECT space, a top-level tree over the shapes' world bounds, an instance table
    bool two_level = false;
    void* d_instances = nullptr;                       // mcrt::InstanceRecord[num_instances], top-level leaf order
    int num_instances = 0;
    std::map<const rrb_shape*, std::shared_ptr<mcrt::Bvh8>> blas_cache;   // meshes are immutable: built once
    std::vector<const rrb_shape*> blas_order;          // bottom-level trees as laid out behind the top level
    std::vector<int64_t> blas_root;                    // node index of each bottom-level root
    int64_t tlas_capacity = 0;                         // node slots reserved in front for the top level
    int64_t two_level_rebuilds = 0, top_level_updates = 0;
    unsigned int* d_counter = nullptr;
    std::vector<rrb_event*> event_pool;
    // staging for the host-buffer entry points
    void* d_stage_rays = nullptr;
    void* d_stage_out = nullptr;
    size_t stage_rays_bytes = 0, stage_out_bytes = 0;
};

namespace {

int set_device(const rrb_api* api) {
    cudaError_t e = cudaSetDevice(api->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    return RRB_OK;
}

rrb_event* acquire_event(rrb_api* api) {
    rrb_event* ev;
    if (!api->event_pool.empty()) {
        ev = api->event_pool.back();
        api->event_pool.pop_back();
    } else {
        ev = new rrb_event();
        if (cudaEventCreateWithFlags(&ev->ev, cudaEventDisableTiming) != cudaSuccess) {
            delete ev;
            return nullptr;
        }
    }
    ev->recorded = false;
    return ev;
}

int record_done(rrb_api* api, rrb_event** done) {
    if (!done) return RRB_OK;
    rrb_event* ev = acquire_event(api);
    if (!ev) return fail(RRB_ERROR_CUDA, "cudaEventCreate failed");
    RRB_CUDA(cudaEventRecord(ev->ev, api->stream));
    ev->recorded = true;
    *done = ev;
    return RRB_OK;
}

int wait_on(rrb_api* api, const rrb_event* wait) {
    if (wait && wait->recorded) RRB_CUDA(cudaStreamWaitEvent(api->stream, wait->ev, 0));
    return RRB_OK;
}

// the pooled object-space geometry (meshes are immutable after CreateMesh: it stays valid while the world holds
// the same shapes in the same order)
void free_geometry_pool(rrb_api* api) {
    if (api->d_ref_vertices) cudaFree(api->d_ref_vertices);
    if (api->d_ref_indices) cudaFree(api->d_ref_indices);
    if (api->d_refit_meshes) cudaFree(api->d_refit_meshes);
    api->d_ref_vertices = nullptr;
    api->d_ref_indices = nullptr;
    api->d_refit_meshes = nullptr;
    api->built_world.clear();
    api->built_mesh_of_shape.clear();
    api->refit_meshes.clear();
    api->pool_face_prefix.clear();
}
void free_tree_refit_state(rrb_api* api) {
    if (api->d_tri_mesh) cudaFree(api->d_tri_mesh);
    if (api->d_boxes) cudaFree(api->d_boxes);
    api->d_tri_mesh = nullptr;
    api->d_boxes = nullptr;
    api->level_offsets.clear();
}
void free_refit(rrb_api* api) {
    free_geometry_pool(api);
    free_tree_refit_state(api);
}

void free_bvh(rrb_api* api) {
    if (api->d_instances) cudaFree(api->d_instances);
    api->d_instances = nullptr;
    api->num_instances = 0;
    api->two_level = false;
    api->blas_order.clear();
    api->blas_root.clear();
    api->tlas_capacity = 0;
    if (api->d_nodes) cudaFree(api->d_nodes);
    if (api->d_tris) cudaFree(api->d_tris);
    api->d_nodes = api->d_tris = nullptr;
    api->num_nodes = api->num_tris = 0;
    free_refit(api);
}

bool refit_enabled(const rrb_api* api) {
    auto it = api->f32_options.find("bvh.refit");
    return it != api->f32_options.end() && it->second != 0.0f;
}

void fill_refit_transform(mcrt::RefitMesh& rm, const rrb_shape* s) {
    std::memcpy(rm.m, s->m, 12 * sizeof(float));
    rm.shape_id = s->id;
}

// Transform-only update of the committed tree (refit_kernels.cu).  Preconditions checked by the caller.
int refit_committed(rrb_api* api) {
    for (size_t w = 0; w < api->world.size(); ++w) {
        const int mi = api->built_mesh_of_shape[w];
        if (mi >= 0) fill_refit_transform(api->refit_meshes[(size_t)mi], api->world[w]);
    }
    RRB_CUDA(cudaMemcpyAsync(api->d_refit_meshes, api->refit_meshes.data(),
                             api->refit_meshes.size() * sizeof(mcrt::RefitMesh), cudaMemcpyHostToDevice, api->stream));
    cudaError_t e = mcrt::refit_bvh8(api->d_nodes, api->d_tris, api->d_boxes, api->d_tri_mesh, api->d_refit_meshes,
                                     api->d_ref_vertices, api->d_ref_indices, (int)api->num_tris,
                                     api->level_offsets.data(), (int)api->level_offsets.size() - 1, api->stream);
    if (e != cudaSuccess) return cuda_fail(e, "refit_bvh8");
    RRB_CUDA(cudaStreamSynchronize(api->stream));      // Commit is blocking (radeon_rays.h:240-242)
    api->launches += (int64_t)api->level_offsets.size();
    api->refits += 1;
    return RRB_OK;
}

int query_common(rrb_api* api, bool any_hit, const rrb_buffer* rays, const rrb_buffer* numrays_buf, int numrays,
                 rrb_buffer* out, rrb_buffer* counts, const rrb_event* wait, rrb_event** done) {
    if (!api || !rays || !out) return fail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    if (numrays < 0) return fail(RRB_ERROR_INVALID_ARGUMENT, "negative ray count");
    if (!api->d_nodes) return fail(RRB_ERROR_NOT_COMMITTED, "Commit() has not been called on a non-empty world");
    if ((size_t)numrays * 48 > rays->size) return fail(RRB_ERROR_INVALID_ARGUMENT, "ray buffer too small");
    if ((size_t)numrays * (any_hit ? 4 : 32) > out->size) return fail(RRB_ERROR_INVALID_ARGUMENT, "result buffer too small");
    if (counts && (size_t)numrays * 8 > counts->size) return fail(RRB_ERROR_INVALID_ARGUMENT, "count buffer too small");
    int rc = set_device(api);
    if (rc) return rc;
    rc = wait_on(api, wait);
    if (rc) return rc;
    cudaError_t e = mcrt::launch_query(any_hit, counts != nullptr, api->d_nodes, api->d_tris, rays->dptr,
                                       numrays_buf ? (const int*)numrays_buf->dptr : nullptr, numrays, out->dptr,
                                       api->d_counter, counts ? (int*)counts->dptr : nullptr, api->stream,
                                       (int)std::min<int64_t>(api->num_nodes, 0x7fffffff), api->two_level ? api->d_instances : nullptr);
    if (e != cudaSuccess) return cuda_fail(e, "traversal kernel launch");
    if (numrays > 0) api->launches++;
    return record_done(api, done);
}

int ensure_stage(rrb_api* api, size_t rays_bytes, size_t out_bytes) {
    if (rays_bytes > api->stage_rays_bytes) {
        if (api->d_stage_rays) cudaFree(api->d_stage_rays);
        api->d_stage_rays = nullptr;
        api->stage_rays_bytes = 0;
        RRB_CUDA(cudaMalloc(&api->d_stage_rays, rays_bytes));
        api->stage_rays_bytes = rays_bytes;
    }
    if (out_bytes > api->stage_out_bytes) {
        if (api->d_stage_out) cudaFree(api->d_stage_out);
        api->d_stage_out = nullptr;
        api->stage_out_bytes = 0;
        RRB_CUDA(cudaMalloc(&api->d_stage_out, out_bytes));
        api->stage_out_bytes = out_bytes;
    }
    return RRB_OK;
}

}  // namespace

namespace mcrt {
// for the other translation units behind rrb.h (rrb_noalloc.cu): sets the text rrb_last_error() returns
int rrb_fail(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace mcrt

extern "C" {

const char* rrb_last_error(void) { return g_last_error.c_str(); }

int rrb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int rrb_device_info(int devidx, char* name, size_t name_len, char* vendor, size_t vendor_len, int* type) {
    cudaDeviceProp p;
    RRB_CUDA(cudaGetDeviceProperties(&p, devidx));
    if (name && name_len) std::snprintf(name, name_len, "%s", p.name);
    if (vendor && vendor_len) std::snprintf(vendor, vendor_len, "NVIDIA Corporation");
    if (type) *type = 1; /* DeviceInfo::kGpu */
    return RRB_OK;
}

int rrb_create(int cuda_device, void* cuda_stream, rrb_api** out) {
    if (!out) return fail(RRB_ERROR_INVALID_ARGUMENT, "out is null");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(RRB_ERROR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                        " (this library has no CPU fallback)");
    if (cuda_device < 0 || cuda_device >= n) return fail(RRB_ERROR_INVALID_ARGUMENT, "device index out of range");
    std::unique_ptr<rrb_api> api(new rrb_api());
    api->device = cuda_device;
    RRB_CUDA(cudaSetDevice(cuda_device));
    if (cuda_stream) {
        api->stream = (cudaStream_t)cuda_stream;
    } else {
        RRB_CUDA(cudaStreamCreateWithFlags(&api->stream, cudaStreamNonBlocking));
        api->own_stream = true;
    }
    RRB_CUDA(cudaMalloc((void**)&api->d_counter, 256));
    *out = api.release();
    return RRB_OK;
}

int rrb_destroy(rrb_api* api) {
    if (!api) return RRB_OK;
    cudaSetDevice(api->device);
    cudaStreamSynchronize(api->stream);
    free_bvh(api);
    if (api->d_counter) cudaFree(api->d_counter);
    if (api->d_stage_rays) cudaFree(api->d_stage_rays);
    if (api->d_stage_out) cudaFree(api->d_stage_out);
    for (rrb_event* ev : api->event_pool) {
        cudaEventDestroy(ev->ev);
        delete ev;
    }
    if (api->own_stream) cudaStreamDestroy(api->stream);
    delete api;
    return RRB_OK;
}

int rrb_create_mesh(rrb_api* api, const float* vertices, int vnum, int vstride, const int* indices, int istride,
                    const int* numfacevertices, int numfaces, rrb_shape** out) {
    if (!api || !out || !vertices || !indices || vnum < 0 || numfaces < 0)
        return fail(RRB_ERROR_INVALID_ARGUMENT, "CreateMesh: bad argument");
    if (vstride == 0) vstride = 3 * (int)sizeof(float);
    std::unique_ptr<rrb_shape> s(new rrb_shape());
    s->vertices.resize(3 * (size_t)vnum);
    for (int i = 0; i < vnum; ++i) {
        const float* v = (const float*)((const char*)vertices + (size_t)i * vstride);
        s->vertices[3 * (size_t)i + 0] = v[0];
        s->vertices[3 * (size_t)i + 1] = v[1];
        s->vertices[3 * (size_t)i + 2] = v[2];
    }
    s->indices.resize(3 * (size_t)numfaces);
    if (!numfacevertices) {
        if (istride == 0) istride = 3 * (int)sizeof(int);
        for (int f = 0; f < numfaces; ++f) {
            const int* id = (const int*)((const char*)indices + (size_t)f * istride);
            s->indices[3 * (size_t)f + 0] = id[0];
            s->indices[3 * (size_t)f + 1] = id[1];
            s->indices[3 * (size_t)f + 2] = id[2];
        }
    } else {
        const char* p = (const char*)indices;
        for (int f = 0; f < numfaces; ++f) {
            if (numfacevertices[f] == 4)
                return fail(RRB_ERROR_UNSUPPORTED, "quads are not supported (Bvh2 asserts pure triangle meshes)");
            if (numfacevertices[f] != 3) return fail(RRB_ERROR_INVALID_ARGUMENT, "Wrong number of vertices per face");
            const int* id = (const int*)p;
            s->indices[3 * (size_t)f + 0] = id[0];
            s->indices[3 * (size_t)f + 1] = id[1];
            s->indices[3 * (size_t)f + 2] = id[2];
            p += istride == 0 ? 3 * sizeof(int) : (size_t)istride;
        }
    }
    for (int v : s->indices)
        if (v < 0 || v >= vnum) return fail(RRB_ERROR_INVALID_ARGUMENT, "CreateMesh: vertex index out of range");
    s->id = api->next_id++;
    *out = s.release();
    return RRB_OK;
}

int rrb_create_instance(rrb_api* api, const rrb_shape* base, rrb_shape** out) {
    if (!api || !base || !out) return fail(RRB_ERROR_INVALID_ARGUMENT, "CreateInstance: null argument");
    rrb_shape* s = new rrb_shape();
    s->is_instance = true;
    s->base = base->is_instance ? base->base : base;
    s->id = api->next_id++;
    *out = s;
    return RRB_OK;
}

int rrb_delete_shape(rrb_api* api, rrb_shape* shape) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (!shape) return RRB_OK;
    auto it = std::find(api->world.begin(), api->world.end(), (const rrb_shape*)shape);
    if (it != api->world.end()) {
        api->world.erase(it);
        api->world_changed = true;
    }
    delete shape;
    return RRB_OK;
}

int rrb_attach(rrb_api* api, const rrb_shape* shape) {
    if (!api || !shape) return fail(RRB_ERROR_INVALID_ARGUMENT, "AttachShape: null argument");
    if (std::find(api->world.begin(), api->world.end(), shape) == api->world.end()) {
        api->world.push_back(shape);
        api->world_changed = true;
    }
    return RRB_OK;
}

int rrb_detach(rrb_api* api, const rrb_shape* shape) {
    if (!api || !shape) return fail(RRB_ERROR_INVALID_ARGUMENT, "DetachShape: null argument");
    auto it = std::find(api->world.begin(), api->world.end(), shape);
    if (it != api->world.end()) {
        api->world.erase(it);
        api->world_changed = true;
    }
    return RRB_OK;
}

int rrb_detach_all(rrb_api* api) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    api->world.clear();
    api->world_changed = true;
    return RRB_OK;
}

namespace {

// CalcIntersectionDevice::Preprocess (src/device/calc_intersection_device.cpp:68-106): two levels when forced, never
// when forced flat, else whenever the world holds an instance.
bool wants_two_level(const rrb_api* api) {
    auto f2 = api->f32_options.find("bvh.force2level");
    if (f2 != api->f32_options.end() && f2->second > 0.0f) return true;
    auto ff = api->f32_options.find("bvh.forceflat");
    if (ff != api->f32_options.end() && ff->second > 0.0f) return false;
    if (refit_enabled(api)) return false;      // "bvh.refit" (this library's extension) refits the FLAT tree in place
    for (const rrb_shape* s : api->world)
        if (s->is_instance) return true;
    return false;
}

// Two-level commit.  Takes the role of IntersectorTwoLevel::Process (src/intersector/intersector_2level.cpp:257-520):
// a bottom-level tree per distinct mesh in object space (built once per mesh and kept: meshes are immutable), a
// top-level tree over the world bounds of every attached shape, and a shape table holding each shape's id, inverse
// matrix and bottom-level root.  A commit that only changed transforms / ids rebuilds the top level and the table
// (microseconds) and leaves the bottom levels where they are on the device.
int commit_two_level(rrb_api* api, std::chrono::steady_clock::time_point t0) {
    mcrt::BuildOptions opt;
    auto b = api->str_options.find("bvh.builder");
    opt.use_sah = !(b != api->str_options.end() && b->second == "median");
    // stack budget (traverse.cuh): two entries per level of either tree + the two groups parked when a ray enters an
    // instance must fit the 40-entry stack
    mcrt::BuildOptions blas_opt = opt, tlas_opt = opt;
    blas_opt.max_levels = 13;
    tlas_opt.max_levels = 5;
    // distinct base meshes in world order
    std::vector<const rrb_shape*> order;
    std::vector<int> blas_of_shape;
    std::map<const rrb_shape*, int> index_of;
    for (const rrb_shape* s : api->world) {
        const rrb_shape* g = s->is_instance ? s->base : s;
        if (!g || g->indices.empty()) {
            blas_of_shape.push_back(-1);
            continue;
        }
        auto it = index_of.find(g);
        if (it == index_of.end()) {
            it = index_of.emplace(g, (int)order.size()).first;
            order.push_back(g);
        }
        blas_of_shape.push_back(it->second);
    }
    if (order.empty()) return fail(RRB_ERROR_EMPTY_SCENE, "Scene is empty.");
    // bottom levels (cached per mesh)
    for (const rrb_shape* g : order) {
        if (api->blas_cache.count(g)) continue;
        mcrt::BuildInputMesh m;
        m.vertices = g->vertices.data();
        m.vnum = (int)(g->vertices.size() / 3);
        m.vstride = 12;
        m.indices = g->indices.data();
        m.istride = 12;
        m.numfaces = (int)(g->indices.size() / 3);
        m.shape_id = 0;                              // the shape table names the shape, not the triangle record
        std::memcpy(m.m, kIdentity, sizeof(m.m));
        auto bvh = std::make_shared<mcrt::Bvh8>();
        try {
            mcrt::build_bvh8(std::vector<mcrt::BuildInputMesh>(1, m), blas_opt, *bvh);
        } catch (const std::exception& ex) {
            return fail(RRB_ERROR_INVALID_ARGUMENT, ex.what());
        }
        api->blas_cache[g] = bvh;
    }
    // instances: world bounds of the transformed vertices, slightly inflated (the hit is found in object space:
    // the two spaces round differently)
    struct Inst {
        const rrb_shape* s;
        int blas;
    };
    std::vector<Inst> inst;
    std::vector<float> boxes;
    for (size_t w = 0; w < api->world.size(); ++w) {
        if (blas_of_shape[w] < 0) continue;
        const rrb_shape* s = api->world[w];
        const rrb_shape* g = s->is_instance ? s->base : s;
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (size_t v = 0; v + 2 < g->vertices.size(); v += 3) {
            const float x = g->vertices[v], y = g->vertices[v + 1], z = g->vertices[v + 2];
            for (int a = 0; a < 3; ++a) {
                const float* r = s->m + 4 * a;
                const float c = ((r[0] * x + r[1] * y) + r[2] * z) + r[3];
                lo[a] = std::min(lo[a], c);
                hi[a] = std::max(hi[a], c);
            }
        }
        for (int a = 0; a < 3; ++a) {
            const float pad = 1e-5f * (hi[a] - lo[a]) + 4.0f * FLT_EPSILON * std::max(std::fabs(lo[a]), std::fabs(hi[a]));
            lo[a] -= pad;
            hi[a] += pad;
        }
        boxes.insert(boxes.end(), lo, lo + 3);
        boxes.insert(boxes.end(), hi, hi + 3);
        inst.push_back({s, blas_of_shape[w]});
    }
    mcrt::Bvh8 tlas;
    try {
        mcrt::build_bvh8_boxes(boxes.data(), (int)inst.size(), tlas_opt, tlas);
    } catch (const std::exception& ex) {
        return fail(RRB_ERROR_INVALID_ARGUMENT, ex.what());
    }
    const bool same_layout = api->two_level && api->d_nodes && api->blas_order == order &&
                             tlas.num_nodes <= api->tlas_capacity && (int)inst.size() == api->num_instances;
    RRB_CUDA(cudaStreamSynchronize(api->stream));
    if (!same_layout) {
        free_bvh(api);
        api->tlas_capacity = std::max<int64_t>((int64_t)inst.size(), tlas.num_nodes);
        int64_t nodes = api->tlas_capacity, tris = 0;
        api->blas_root.clear();
        for (const rrb_shape* g : order) {
            api->blas_root.push_back(nodes);
            nodes += api->blas_cache[g]->num_nodes;
            tris += api->blas_cache[g]->num_tris;
        }
        std::vector<uint32_t> all_nodes((size_t)nodes * 20, 0u);
        std::vector<float> all_tris((size_t)tris * 12);
        int64_t tri_at = 0;
        for (size_t k = 0; k < order.size(); ++k) {
            const mcrt::Bvh8& bv = *api->blas_cache[order[k]];
            uint32_t* dst = all_nodes.data() + (size_t)api->blas_root[k] * 20;
            std::memcpy(dst, bv.nodes.data(), bv.nodes.size() * sizeof(uint32_t));
            for (int64_t n = 0; n < bv.num_nodes; ++n) {      // child and triangle bases become absolute
                dst[20 * n + 4] += (uint32_t)api->blas_root[k];
                dst[20 * n + 5] += (uint32_t)tri_at;
            }
            std::memcpy(all_tris.data() + (size_t)tri_at * 12, bv.triangles.data(), bv.triangles.size() * sizeof(float));
            tri_at += bv.num_tris;
        }
        RRB_CUDA(cudaMalloc(&api->d_nodes, all_nodes.size() * sizeof(uint32_t)));
        RRB_CUDA(cudaMalloc(&api->d_tris, std::max<size_t>(all_tris.size(), 12) * sizeof(float)));
        RRB_CUDA(cudaMalloc(&api->d_instances, inst.size() * sizeof(mcrt::InstanceRecord)));
        RRB_CUDA(cudaMemcpyAsync(api->d_nodes, all_nodes.data(), all_nodes.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                 api->stream));
        RRB_CUDA(cudaMemcpyAsync(api->d_tris, all_tris.data(), all_tris.size() * sizeof(float), cudaMemcpyHostToDevice,
                                 api->stream));
        RRB_CUDA(cudaStreamSynchronize(api->stream));
        api->num_nodes = nodes;
        api->num_tris = tris;
        api->num_instances = (int)inst.size();
        api->blas_order = order;
        api->two_level = true;
        api->two_level_rebuilds++;
    } else {
        api->top_level_updates++;
    }
    // top level (front of the node array) + the shape table in the top level's leaf order
    std::vector<mcrt::InstanceRecord> table(inst.size());
    for (size_t i = 0; i < inst.size(); ++i) {
        int src;
        std::memcpy(&src, &tlas.triangles[12 * i + 3], 4);
        const Inst& in = inst[(size_t)src];
        mcrt::InstanceRecord& r = table[i];
        std::memcpy(r.minv, in.s->minv, 12 * sizeof(float));
        r.blas_root = (int)api->blas_root[(size_t)in.blas];
        r.shape_id = in.s->id;
        r.pad[0] = r.pad[1] = 0;
    }
    RRB_CUDA(cudaMemcpyAsync(api->d_nodes, tlas.nodes.data(), tlas.nodes.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                             api->stream));
    RRB_CUDA(cudaMemcpyAsync(api->d_instances, table.data(), table.size() * sizeof(mcrt::InstanceRecord),
                             cudaMemcpyHostToDevice, api->stream));
    RRB_CUDA(cudaStreamSynchronize(api->stream));
    api->level_offsets = tlas.level_offsets;
    api->world_changed = false;
    api->last_build_on_device = false;
    for (const rrb_shape* s : api->world) {
        const_cast<rrb_shape*>(s)->changed = false;
        if (s->base) const_cast<rrb_shape*>(s->base)->changed = false;
    }
    api->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return RRB_OK;
}

}  // namespace

int rrb_commit(rrb_api* api) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (api->world.empty()) return fail(RRB_ERROR_EMPTY_SCENE, "Scene is empty.");
    bool dirty = api->world_changed || !api->d_nodes;
    for (const rrb_shape* s : api->world) dirty = dirty || s->changed || (s->base && s->base->changed);
    if (wants_two_level(api) != api->two_level) dirty = true;
    if (!dirty) return RRB_OK;
    int rc = set_device(api);
    if (rc) return rc;
    auto t0 = std::chrono::steady_clock::now();
    api->last_commit_was_refit = false;
    if (wants_two_level(api)) return commit_two_level(api, t0);
    if (api->two_level) free_bvh(api);      // back from two levels to the flat tree
    // Option "bvh.refit" (an extension; the reference always rebuilds, intersector_lds.cpp:166): when the world
    // holds the same shapes as at the last full build and only transforms / ids changed, keep the topology
    // and refit on the device.
    if (refit_enabled(api) && api->d_nodes && !api->world_changed && api->d_boxes && api->built_world == api->world) {
        rc = refit_committed(api);
        if (rc) return rc;
        for (const rrb_shape* s : api->world) {
            const_cast<rrb_shape*>(s)->changed = false;
            if (s->base) const_cast<rrb_shape*>(s->base)->changed = false;
        }
        api->last_commit_was_refit = true;
        api->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        return RRB_OK;
    }
    std::vector<mcrt::BuildInputMesh> meshes;
    std::vector<int> mesh_of_shape;
    meshes.reserve(api->world.size());
    for (const rrb_shape* s : api->world) {
        const rrb_shape* g = s->is_instance ? s->base : s;
        if (!g || g->indices.empty()) {   // Intersection_1Ray_InstanceNoShape: nothing to add
            mesh_of_shape.push_back(-1);
            continue;
        }
        mesh_of_shape.push_back((int)meshes.size());
        mcrt::BuildInputMesh m;
        m.vertices = g->vertices.data();
        m.vnum = (int)(g->vertices.size() / 3);
        m.vstride = 12;
        m.indices = g->indices.data();
        m.istride = 12;
        m.numfaces = (int)(g->indices.size() / 3);
        m.shape_id = s->id;
        std::memcpy(m.m, s->m, sizeof(m.m));
        meshes.push_back(m);
    }
    mcrt::BuildOptions opt;
    auto b = api->str_options.find("bvh.builder");
    opt.use_sah = !(b != api->str_options.end() && b->second == "median");
    // acc.type = "hlbvh" is Radeon Rays' GPU-built hierarchy (intersector_hlbvh.cpp); here it selects the device
    // builder of the 8-wide tree (lbvh_build.cu).  bvh.builder = "lbvh" does the same.
    auto acc = api->str_options.find("acc.type");
    const bool want_device_build = (acc != api->str_options.end() && acc->second == "hlbvh") ||
                                   (b != api->str_options.end() && b->second == "lbvh");
    api->last_build_on_device = false;
    RRB_CUDA(cudaStreamSynchronize(api->stream));
    // a rebuild over the same shapes (moved, re-numbered) keeps the pooled geometry on the device
    const bool keep_pool = api->d_ref_vertices && !api->world_changed && api->built_world == api->world &&
                           api->built_mesh_of_shape == mesh_of_shape;
    if (keep_pool) {
        if (api->d_nodes) cudaFree(api->d_nodes);
        if (api->d_tris) cudaFree(api->d_tris);
        api->d_nodes = api->d_tris = nullptr;
        api->num_nodes = api->num_tris = 0;
        free_tree_refit_state(api);
    } else {
        free_bvh(api);
    }

    const bool trace = getenv("MCRT_BUILD_TRACE") != nullptr;
    auto lap = [&](const char* what) {
        if (trace) fprintf(stderr, "[rrb_commit] %-22s %8.2f ms\n", what,
                           std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e3);
    };
    lap("gathered meshes");
    // pooled object-space geometry on the device: input of the device builder and of later refits
    std::vector<int> face_prefix(1, 0);
    if (keep_pool) {
        size_t w = 0;
        for (const rrb_shape* s : api->world) {
            const int mi = mesh_of_shape[w++];
            if (mi >= 0) fill_refit_transform(api->refit_meshes[(size_t)mi], s);
        }
        RRB_CUDA(cudaMemcpyAsync(api->d_refit_meshes, api->refit_meshes.data(),
                                 api->refit_meshes.size() * sizeof(mcrt::RefitMesh), cudaMemcpyHostToDevice, api->stream));
        face_prefix = api->pool_face_prefix;
        lap("kept geometry pool");
    } else if (want_device_build || refit_enabled(api)) {
        std::vector<float> pv;
        std::vector<int> pi;
        api->refit_meshes.clear();
        size_t w = 0;
        for (const rrb_shape* s : api->world) {
            const rrb_shape* g = s->is_instance ? s->base : s;
            if (mesh_of_shape[w++] < 0) continue;
            mcrt::RefitMesh rm;
            std::memset(&rm, 0, sizeof(rm));
            rm.vertex_base = (int)(pv.size() / 3);
            rm.index_base = (int)pi.size();
            fill_refit_transform(rm, s);
            for (int id : g->indices)
                if (id < 0 || (size_t)id >= g->vertices.size() / 3)
                    return fail(RRB_ERROR_INVALID_ARGUMENT, "vertex index out of range");
            pv.insert(pv.end(), g->vertices.begin(), g->vertices.end());
            pi.insert(pi.end(), g->indices.begin(), g->indices.end());
            face_prefix.push_back(face_prefix.back() + (int)(g->indices.size() / 3));
            api->refit_meshes.push_back(rm);
        }
        if (api->refit_meshes.empty()) return fail(RRB_ERROR_EMPTY_SCENE, "Scene is empty.");
        RRB_CUDA(cudaMalloc((void**)&api->d_ref_vertices, pv.size() * sizeof(float)));
        RRB_CUDA(cudaMalloc((void**)&api->d_ref_indices, pi.size() * sizeof(int)));
        RRB_CUDA(cudaMalloc((void**)&api->d_refit_meshes, api->refit_meshes.size() * sizeof(mcrt::RefitMesh)));
        RRB_CUDA(cudaMemcpyAsync(api->d_ref_vertices, pv.data(), pv.size() * sizeof(float), cudaMemcpyHostToDevice, api->stream));
        RRB_CUDA(cudaMemcpyAsync(api->d_ref_indices, pi.data(), pi.size() * sizeof(int), cudaMemcpyHostToDevice, api->stream));
        RRB_CUDA(cudaMemcpyAsync(api->d_refit_meshes, api->refit_meshes.data(),
                                 api->refit_meshes.size() * sizeof(mcrt::RefitMesh), cudaMemcpyHostToDevice, api->stream));
        RRB_CUDA(cudaStreamSynchronize(api->stream));
        api->pool_face_prefix = face_prefix;
        lap("pooled + uploaded");
    }

    bool built = false;
    if (want_device_build) {
        mcrt::DeviceBvh8 dev;
        std::string why;
        if (mcrt::build_bvh8_device(api->d_ref_vertices, api->d_ref_indices, api->d_refit_meshes, face_prefix, &dev,
                                    api->stream, &why)) {
            api->d_nodes = dev.nodes;
            api->d_tris = dev.tris;
            api->d_tri_mesh = dev.tri_mesh;
            api->d_boxes = dev.boxes;
            api->num_nodes = dev.num_nodes;
            api->num_tris = dev.num_tris;
            api->level_offsets.assign(dev.level_offsets.begin(), dev.level_offsets.end());
            api->launches += dev.launches;
            api->last_build_on_device = true;
            built = true;
            lap("device build");
        } else {
            api->device_build_fallback_reason = why;    // e.g. fewer than 4 triangles: the host builder takes over
        }
    }
    if (!built) {
        mcrt::Bvh8 bvh;
        try {
            mcrt::build_bvh8(meshes, opt, bvh);
        } catch (const std::exception& ex) {
            bool empty = std::string(ex.what()) == "Scene is empty.";
            return fail(empty ? RRB_ERROR_EMPTY_SCENE : RRB_ERROR_INVALID_ARGUMENT, ex.what());
        }
        RRB_CUDA(cudaMalloc(&api->d_nodes, bvh.nodes.size() * sizeof(uint32_t)));
        RRB_CUDA(cudaMalloc(&api->d_tris, bvh.triangles.size() * sizeof(float)));
        RRB_CUDA(cudaMemcpyAsync(api->d_nodes, bvh.nodes.data(), bvh.nodes.size() * sizeof(uint32_t),
                                 cudaMemcpyHostToDevice, api->stream));
        RRB_CUDA(cudaMemcpyAsync(api->d_tris, bvh.triangles.data(), bvh.triangles.size() * sizeof(float),
                                 cudaMemcpyHostToDevice, api->stream));
        api->num_nodes = bvh.num_nodes;
        api->num_tris = bvh.num_tris;
        if (api->d_ref_vertices) {
            // per-triangle mesh index + per-node exact boxes for later refits
            RRB_CUDA(cudaMalloc((void**)&api->d_tri_mesh, bvh.tri_mesh.size() * sizeof(int)));
            RRB_CUDA(cudaMalloc((void**)&api->d_boxes, (size_t)bvh.num_nodes * 6 * sizeof(float)));
            RRB_CUDA(cudaMemcpyAsync(api->d_tri_mesh, bvh.tri_mesh.data(), bvh.tri_mesh.size() * sizeof(int),
                                     cudaMemcpyHostToDevice, api->stream));
            api->level_offsets = bvh.level_offsets;
        }
        RRB_CUDA(cudaStreamSynchronize(api->stream));
    }
    if (api->d_ref_vertices) {
        api->built_world = api->world;
        api->built_mesh_of_shape = mesh_of_shape;
    }
    api->world_changed = false;
    for (const rrb_shape* s : api->world) {
        const_cast<rrb_shape*>(s)->changed = false;
        if (s->base) const_cast<rrb_shape*>(s->base)->changed = false;
    }
    api->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return RRB_OK;
}

int rrb_reset_id_counter(rrb_api* api) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    api->next_id = 1;
    return RRB_OK;
}

int rrb_is_world_empty(rrb_api* api, int* out_empty) {
    if (!api || !out_empty) return fail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    *out_empty = api->world.empty() ? 1 : 0;
    return RRB_OK;
}

int rrb_shape_set_transform(rrb_shape* shape, const float m[16], const float minv[16]) {
    if (!shape || !m) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetTransform: null argument");
    std::memcpy(shape->m, m, sizeof(shape->m));
    if (minv) std::memcpy(shape->minv, minv, sizeof(shape->minv));
    shape->changed = true;
    return RRB_OK;
}

int rrb_shape_get_transform(const rrb_shape* shape, float m[16], float minv[16]) {
    if (!shape) return fail(RRB_ERROR_INVALID_ARGUMENT, "GetTransform: null shape");
    if (m) std::memcpy(m, shape->m, sizeof(shape->m));
    if (minv) std::memcpy(minv, shape->minv, sizeof(shape->minv));
    return RRB_OK;
}

int rrb_shape_set_id(rrb_shape* shape, int id) {
    if (!shape) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetId: null shape");
    shape->id = id;
    shape->changed = true;
    return RRB_OK;
}

int rrb_shape_get_id(const rrb_shape* shape, int* out_id) {
    if (!shape || !out_id) return fail(RRB_ERROR_INVALID_ARGUMENT, "GetId: null argument");
    *out_id = shape->id;
    return RRB_OK;
}

int rrb_shape_set_linear_velocity(rrb_shape* shape, const float v[3]) {
    if (!shape || !v) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetLinearVelocity: null argument");
    std::memcpy(shape->linear_velocity, v, sizeof(shape->linear_velocity));
    return RRB_OK;
}

int rrb_shape_set_angular_velocity(rrb_shape* shape, const float q[4]) {
    if (!shape || !q) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetAngularVelocity: null argument");
    std::memcpy(shape->angular_velocity, q, sizeof(shape->angular_velocity));
    return RRB_OK;
}

int rrb_set_option_str(rrb_api* api, const char* name, const char* value) {
    if (!api || !name || !value) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetOption: null argument");
    auto it = api->str_options.find(name);
    if (it == api->str_options.end() || it->second != value) {
        api->str_options[name] = value;
        api->world_changed = true;   // options feed the builder
    }
    return RRB_OK;
}

int rrb_set_option_f32(rrb_api* api, const char* name, float value) {
    if (!api || !name) return fail(RRB_ERROR_INVALID_ARGUMENT, "SetOption: null argument");
    api->f32_options[name] = value;
    return RRB_OK;
}

int rrb_create_buffer(rrb_api* api, size_t size, const void* initdata, rrb_buffer** out) {
    if (!api || !out) return fail(RRB_ERROR_INVALID_ARGUMENT, "CreateBuffer: null argument");
    int rc = set_device(api);
    if (rc) return rc;
    std::unique_ptr<rrb_buffer> b(new rrb_buffer());
    b->size = size;
    b->owning = true;
    if (size) {
        cudaError_t e = cudaMalloc(&b->dptr, size);
        if (e != cudaSuccess) return fail(RRB_ERROR_OUT_OF_MEMORY, std::string("cudaMalloc: ") + cudaGetErrorString(e));
        if (initdata) {
            RRB_CUDA(cudaMemcpyAsync(b->dptr, initdata, size, cudaMemcpyHostToDevice, api->stream));
            RRB_CUDA(cudaStreamSynchronize(api->stream));
        }
    }
    *out = b.release();
    return RRB_OK;
}

int rrb_wrap_device_ptr(rrb_api* api, void* device_ptr, size_t size, rrb_buffer** out) {
    if (!api || !out || (!device_ptr && size)) return fail(RRB_ERROR_INVALID_ARGUMENT, "wrap: null argument");
    rrb_buffer* b = new rrb_buffer();
    b->dptr = device_ptr;
    b->size = size;
    b->owning = false;
    *out = b;
    return RRB_OK;
}

int rrb_delete_buffer(rrb_api* api, rrb_buffer* buffer) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (!buffer) return RRB_OK;
    if (buffer->host_map) cudaFreeHost(buffer->host_map);
    if (buffer->owning && buffer->dptr) {
        cudaSetDevice(api->device);
        cudaStreamSynchronize(api->stream);
        cudaFree(buffer->dptr);
    }
    delete buffer;
    return RRB_OK;
}

int rrb_buffer_device_ptr(const rrb_buffer* buffer, void** out_ptr, size_t* out_size) {
    if (!buffer) return fail(RRB_ERROR_INVALID_ARGUMENT, "null buffer");
    if (out_ptr) *out_ptr = buffer->dptr;
    if (out_size) *out_size = buffer->size;
    return RRB_OK;
}

int rrb_map(rrb_api* api, rrb_buffer* buffer, int map_type, size_t offset, size_t size, void** data, rrb_event** ev) {
    if (!api || !buffer || !data) return fail(RRB_ERROR_INVALID_ARGUMENT, "MapBuffer: null argument");
    if (offset + size > buffer->size) return fail(RRB_ERROR_INVALID_ARGUMENT, "MapBuffer: range out of bounds");
    if (buffer->host_map) return fail(RRB_ERROR_INVALID_ARGUMENT, "MapBuffer: buffer is already mapped");
    int rc = set_device(api);
    if (rc) return rc;
    RRB_CUDA(cudaMallocHost(&buffer->host_map, size ? size : 1));
    buffer->map_offset = offset;
    buffer->map_size = size;
    buffer->map_type = map_type;
    if ((map_type & RRB_MAP_READ) && size)
        RRB_CUDA(cudaMemcpyAsync(buffer->host_map, (const char*)buffer->dptr + offset, size, cudaMemcpyDeviceToHost,
                                 api->stream));
    *data = buffer->host_map;
    return record_done(api, ev);
}

int rrb_unmap(rrb_api* api, rrb_buffer* buffer, void* ptr, rrb_event** ev) {
    if (!api || !buffer) return fail(RRB_ERROR_INVALID_ARGUMENT, "UnmapBuffer: null argument");
    if (!buffer->host_map || (ptr && ptr != buffer->host_map))
        return fail(RRB_ERROR_INVALID_ARGUMENT, "UnmapBuffer: pointer was not returned by MapBuffer");
    int rc = set_device(api);
    if (rc) return rc;
    if ((buffer->map_type & RRB_MAP_WRITE) && buffer->map_size)
        RRB_CUDA(cudaMemcpyAsync((char*)buffer->dptr + buffer->map_offset, buffer->host_map, buffer->map_size,
                                 cudaMemcpyHostToDevice, api->stream));
    // the staging block must outlive the copy: release it once the stream has passed this point
    RRB_CUDA(cudaStreamSynchronize(api->stream));
    cudaFreeHost(buffer->host_map);
    buffer->host_map = nullptr;
    return record_done(api, ev);
}

int rrb_event_complete(rrb_event* ev, int* out_complete) {
    if (!ev || !out_complete) return fail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    if (!ev->recorded) {
        *out_complete = 1;
        return RRB_OK;
    }
    cudaError_t e = cudaEventQuery(ev->ev);
    if (e == cudaSuccess) *out_complete = 1;
    else if (e == cudaErrorNotReady) *out_complete = 0;
    else return cuda_fail(e, "cudaEventQuery");
    return RRB_OK;
}

int rrb_event_wait(rrb_event* ev) {
    if (!ev) return fail(RRB_ERROR_INVALID_ARGUMENT, "null event");
    if (ev->recorded) RRB_CUDA(cudaEventSynchronize(ev->ev));
    return RRB_OK;
}

int rrb_event_delete(rrb_api* api, rrb_event* ev) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (ev) api->event_pool.push_back(ev);
    return RRB_OK;
}

int rrb_query_intersection(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hits,
                           const rrb_event* wait, rrb_event** done) {
    return query_common(api, false, rays, nullptr, numrays, hits, nullptr, wait, done);
}

int rrb_query_occlusion(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hitresults,
                        const rrb_event* wait, rrb_event** done) {
    return query_common(api, true, rays, nullptr, numrays, hitresults, nullptr, wait, done);
}

int rrb_query_intersection_indirect(rrb_api* api, const rrb_buffer* rays, const rrb_buffer* numrays, int maxrays,
                                    rrb_buffer* hits, const rrb_event* wait, rrb_event** done) {
    if (!numrays) return fail(RRB_ERROR_INVALID_ARGUMENT, "null ray-count buffer");
    return query_common(api, false, rays, numrays, maxrays, hits, nullptr, wait, done);
}

int rrb_query_occlusion_indirect(rrb_api* api, const rrb_buffer* rays, const rrb_buffer* numrays, int maxrays,
                                 rrb_buffer* hitresults, const rrb_event* wait, rrb_event** done) {
    if (!numrays) return fail(RRB_ERROR_INVALID_ARGUMENT, "null ray-count buffer");
    return query_common(api, true, rays, numrays, maxrays, hitresults, nullptr, wait, done);
}

int rrb_query_intersection_counted(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hits,
                                   rrb_buffer* counts) {
    if (!counts) return fail(RRB_ERROR_INVALID_ARGUMENT, "null count buffer");
    return query_common(api, false, rays, nullptr, numrays, hits, counts, nullptr, nullptr);
}

static int query_host(rrb_api* api, bool any_hit, const void* host_rays, int numrays, void* host_out) {
    if (!api || !host_rays || !host_out || numrays < 0) return fail(RRB_ERROR_INVALID_ARGUMENT, "bad argument");
    if (numrays == 0) return RRB_OK;
    int rc = set_device(api);
    if (rc) return rc;
    const size_t rb = (size_t)numrays * 48, ob = (size_t)numrays * (any_hit ? 4 : 32);
    rc = ensure_stage(api, rb, ob);
    if (rc) return rc;
    RRB_CUDA(cudaMemcpyAsync(api->d_stage_rays, host_rays, rb, cudaMemcpyHostToDevice, api->stream));
    // results of inactive rays are "not written": carry the caller's prefilled values through
    RRB_CUDA(cudaMemcpyAsync(api->d_stage_out, host_out, ob, cudaMemcpyHostToDevice, api->stream));
    rrb_buffer r, o;
    r.dptr = api->d_stage_rays;
    r.size = rb;
    o.dptr = api->d_stage_out;
    o.size = ob;
    rc = query_common(api, any_hit, &r, nullptr, numrays, &o, nullptr, nullptr, nullptr);
    if (rc) return rc;
    RRB_CUDA(cudaMemcpyAsync(host_out, api->d_stage_out, ob, cudaMemcpyDeviceToHost, api->stream));
    RRB_CUDA(cudaStreamSynchronize(api->stream));
    return RRB_OK;
}

int rrb_query_intersection_host(rrb_api* api, const void* host_rays, int numrays, void* host_hits) {
    return query_host(api, false, host_rays, numrays, host_hits);
}

int rrb_query_occlusion_host(rrb_api* api, const void* host_rays, int numrays, int* host_results) {
    return query_host(api, true, host_rays, numrays, host_results);
}

int rrb_get_stats(rrb_api* api, rrb_stats* out) {
    if (!api || !out) return fail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    out->num_triangles = api->num_tris;
    out->num_nodes = api->num_nodes;
    out->node_bytes = api->num_nodes * 80;
    out->triangle_bytes = api->num_tris * 48;
    out->build_seconds = api->build_seconds;
    out->launches = api->launches;
    out->refits = api->refits;
    out->last_commit_refit = api->last_commit_was_refit ? 1 : 0;
    out->last_build_on_device = api->last_build_on_device ? 1 : 0;
    out->num_levels = api->level_offsets.empty() ? 0 : (int64_t)api->level_offsets.size() - 1;
    out->two_level = api->two_level ? 1 : 0;
    out->num_instances = api->two_level ? api->num_instances : 0;
    out->two_level_rebuilds = api->two_level_rebuilds;
    out->top_level_updates = api->top_level_updates;
    return RRB_OK;
}

int rrb_get_bvh(rrb_api* api, const void** nodes, const void** triangles, int64_t* num_nodes, int64_t* num_tris) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (!api->d_nodes) return fail(RRB_ERROR_NOT_COMMITTED, "Commit() has not been called on a non-empty world");
    if (nodes) *nodes = api->d_nodes;
    if (triangles) *triangles = api->d_tris;
    if (num_nodes) *num_nodes = api->num_nodes;
    if (num_tris) *num_tris = api->num_tris;
    return RRB_OK;
}

int rrb_get_instances(rrb_api* api, const void** instances, int* num_instances) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (instances) *instances = api->two_level ? api->d_instances : nullptr;
    if (num_instances) *num_instances = api->two_level ? api->num_instances : 0;
    return RRB_OK;
}

int rrb_get_stream(rrb_api* api, int* cuda_device, void** cuda_stream) {
    if (!api) return fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (cuda_device) *cuda_device = api->device;
    if (cuda_stream) *cuda_stream = (void*)api->stream;
    return RRB_OK;
}

struct rrb_host_bvh {
    mcrt::Bvh8 bvh;
};

int rrb_host_bvh_build(const float* vertices, int vnum, const int* indices, int numfaces, const int* face_shape_ids,
                       int use_sah, rrb_host_bvh** out) {
    if (!vertices || !indices || !out) return fail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    // one input mesh per run of equal shape ids, so prim ids restart with every shape
    std::vector<mcrt::BuildInputMesh> meshes;
    int f = 0;
    while (f < numfaces) {
        int g = f;
        const int sid = face_shape_ids ? face_shape_ids[f] : 0;
        while (g < numfaces && (face_shape_ids ? face_shape_ids[g] : 0) == sid) ++g;
        mcrt::BuildInputMesh m;
        m.vertices = vertices;
        m.vnum = vnum;
        m.vstride = 12;
        m.indices = indices + 3 * (size_t)f;
        m.istride = 12;
        m.numfaces = g - f;
        m.shape_id = sid;
        std::memcpy(m.m, kIdentity, sizeof(m.m));
        meshes.push_back(m);
        f = g;
    }
    std::unique_ptr<rrb_host_bvh> h(new rrb_host_bvh());
    mcrt::BuildOptions opt;
    opt.use_sah = (use_sah & 1) != 0;
    opt.depth_guard = (use_sah & 2) == 0;
    try {
        mcrt::build_bvh8(meshes, opt, h->bvh);
    } catch (const std::exception& ex) {
        bool empty = std::string(ex.what()) == "Scene is empty.";
        return fail(empty ? RRB_ERROR_EMPTY_SCENE : RRB_ERROR_INVALID_ARGUMENT, ex.what());
    }
    *out = h.release();
    return RRB_OK;
}

int rrb_host_bvh_get(const rrb_host_bvh* bvh, const uint32_t** nodes, int64_t* num_nodes, const float** triangles,
                     int64_t* num_tris) {
    if (!bvh) return fail(RRB_ERROR_INVALID_ARGUMENT, "null bvh");
    if (nodes) *nodes = bvh->bvh.nodes.data();
    if (num_nodes) *num_nodes = bvh->bvh.num_nodes;
    if (triangles) *triangles = bvh->bvh.triangles.data();
    if (num_tris) *num_tris = bvh->bvh.num_tris;
    return RRB_OK;
}

int rrb_host_bvh_get_refit_info(const rrb_host_bvh* bvh, const int32_t** tri_mesh, const int32_t** level_offsets,
                                int* num_levels) {
    if (!bvh) return fail(RRB_ERROR_INVALID_ARGUMENT, "null bvh");
    if (tri_mesh) *tri_mesh = bvh->bvh.tri_mesh.data();
    if (level_offsets) *level_offsets = bvh->bvh.level_offsets.data();
    if (num_levels) *num_levels = (int)bvh->bvh.level_offsets.size() - 1;
    return RRB_OK;
}

int rrb_host_bvh_free(rrb_host_bvh* bvh) {
    delete bvh;
    return RRB_OK;
}

}  // extern "C"
