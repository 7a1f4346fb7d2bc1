// ptb_api.cu -- implementation of include/ptb.h: device-side scene, per-pixel buffers and the
// stage / frame entry points of the path tracer.
//
// Takes the role of RTScene's upload half (scene/RTScene.cpp:178-197,680-809), of
// RTBufferManager (system/RTBufferManager.h:59-71) and of the kernel-launch halves of
// RTPrimaryRaysPass / RTPathTracingPass / RTReconstructionPass.  No CPU fallback.
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/ptb.h"
#include "bdpt_kernels.h"
#include "pt_kernels.h"

namespace {
thread_local std::string g_ptb_error;

int pfail(int code, const std::string& msg) {
    g_ptb_error = msg;
    return code;
}
#define PTB_CUDA(call)                                                                                 \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
};
}  // namespace

struct ptb_context {
    rrb_api* api = nullptr;
    int device = 0;
    cudaStream_t stream = nullptr;
    int width = 0, height = 0;
    int part = 0, num_parts = 1;
    bool scene_loaded = false;
    mcrt::PtState st;
    mcrt::BdptState bd;
    std::map<std::string, DevBuf> scene_bufs;     // freed on re-upload
    std::map<std::string, DevBuf> pixel_bufs;     // freed on resize
    std::map<std::string, DevBuf> named;          // ptb_get_buffer view (non-owning)
    uint32_t* d_pixels = nullptr;                 // partition pixel list (NULL = identity)
    int n_pixels = 0;
    int bdpt_depth = -1;
    bool last_frame_bdpt = false;                 // the frame in "RadianceBufferCL" came from ptb_bdpt_render_sample
    float* merged = nullptr;                      // ptb_merge_accumulation(root < 0): all-reduce result, 20 B/pixel, lazy
    int path_trace_depth = 0;                     // ptb_set_path_trace: bounces recorded (0 = off)
    int64_t launches = 0;
    float4* denoised = nullptr;                   // RTDenoisePass / RTToneMappingPass outputs
    float4* tonemapped = nullptr;
    int sampler_kind = 1;                         // RT_SAMPLER_RANDOM (samplers.cl:17)
    const uint32_t* d_sobol = nullptr;
    int sobol_dims = 0;
    bool profiling = false;
    mcrt::StageTimer timer;
    mcrt::PtOverlap overlap{nullptr, nullptr, nullptr, nullptr};   // second stream of a frame (ptb_set_overlap)
    // asynchronous read-back (ptb_read_buffer_async): device snapshot + copy stream
    // (two snapshots in rotation: the host waits for the copy issued two calls ago, not for the previous one, so it
    // can keep enqueuing frames while an image is still crossing PCIe)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t snap_ready[2] = {nullptr, nullptr}, copy_done[2] = {nullptr, nullptr};
    void* snapshot[2] = {nullptr, nullptr};
    size_t snapshot_bytes[2] = {0, 0};
    bool copy_pending[2] = {false, false};
    int snap_next = 0;
    int halves = 2;
    int num_materials = 0;
    int shade_split_wanted = 0;        // ptb_set_shade_split / MCRT_SHADE_SPLIT
    bool split_supported = false;      // no material binds a texture the record of the split stage does not carry
    int num_scene_tris = 0;                       // triangles covered by the per-triangle shading records (0 = none)
    const int* d_tri_shape = nullptr;
    // Frames in flight (ptb_set_frames_in_flight): with D > 1 consecutive ptb_pt_render_sample calls run on D side
    // streams with their own wavefront state, so that the latency chain of one frame (the few long rays that keep
    // every traversal launch alive, profiles/r01_c5_tiles.md) is filled by the next frames' work.  ptb_accumulate
    // consumes the frames in the order they were rendered, on the context's stream.
    struct FrameSlot {
        cudaStream_t stream = nullptr;
        cudaEvent_t render_done = nullptr, slot_free = nullptr;
        bool done_recorded = false, free_recorded = false;
        mcrt::PtOverlap overlap{nullptr, nullptr, nullptr, nullptr};
        mcrt::PtState st;     // per-slot pointers filled in by alloc_slots(); scene / options copied at launch
        void* camera = nullptr;
        uint64_t camera_version = 0;
    };
    unsigned char camera_host[176] = {0};   // last ptb_set_camera argument: each slot takes its copy when it renders
    uint64_t camera_version = 0;
    std::vector<FrameSlot> slots;     // empty = one frame at a time on the context's stream
    int next_slot = 0;
    std::deque<int> rendered;         // slots holding a frame that has not been accumulated yet
    ptb_context() {
        std::memset(&st, 0, sizeof(st));
        std::memset(&bd, 0, sizeof(bd));
    }
};

namespace {

int set_dev(ptb_context* c) {
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    return 0;
}

void free_map(std::map<std::string, DevBuf>& m) {
    for (auto& kv : m)
        if (kv.second.p) cudaFree(kv.second.p);
    m.clear();
}

int alloc_into(std::map<std::string, DevBuf>& m, const std::string& name, size_t bytes, bool zero, cudaStream_t s,
               void** out) {
    DevBuf b;
    b.bytes = bytes;
    if (bytes) {
        cudaError_t e = cudaMalloc(&b.p, bytes);
        if (e != cudaSuccess)
            return pfail(RRB_ERROR_OUT_OF_MEMORY, "cudaMalloc(" + name + ", " + std::to_string(bytes) + "): " + cudaGetErrorString(e));
        if (zero) {
            e = cudaMemsetAsync(b.p, 0, bytes, s);
            if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("cudaMemsetAsync: ") + cudaGetErrorString(e));
        }
    }
    m[name] = b;
    *out = b.p;
    return 0;
}

int upload(ptb_context* c, const std::string& name, const void* src, size_t bytes, const void** out) {
    void* p = nullptr;
    // keep a small non-null allocation for empty arrays so kernels never see NULL tables
    int rc = alloc_into(c->scene_bufs, name, bytes ? bytes : 16, bytes == 0, c->stream, &p);
    if (rc) return rc;
    if (bytes) {
        if (!src) return pfail(RRB_ERROR_INVALID_ARGUMENT, "scene array '" + name + "' is null");
        cudaError_t e = cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, c->stream);
        if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e));
    }
    *out = p;
    return 0;
}

// The Sobol table outlives scene uploads (it is not scene data); owned by the context.
int set_sobol_table(ptb_context* c, const uint32_t* matrices, int num_dimensions) {
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    if (c->d_sobol) cudaFree((void*)c->d_sobol);
    c->d_sobol = nullptr;
    void* p = nullptr;
    const size_t bytes = (size_t)num_dimensions * 52 * 4;
    PTB_CUDA(cudaMalloc(&p, bytes));
    PTB_CUDA(cudaMemcpyAsync(p, matrices, bytes, cudaMemcpyHostToDevice, c->stream));
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->d_sobol = (const uint32_t*)p;
    c->sobol_dims = num_dimensions;
    if (c->sampler_kind == 0) c->st.sobol = c->d_sobol;
    return 0;
}

int refresh_bvh(ptb_context* c) {
    const void *nodes = nullptr, *tris = nullptr;
    int64_t num_nodes = 0;
    int rc = rrb_get_bvh(c->api, &nodes, &tris, &num_nodes, nullptr);
    if (rc) return pfail(rc, rrb_last_error());
    c->st.nodes = (const uint4*)nodes;
    c->st.tris = (const float4*)tris;
    c->st.numNodes = (int)std::min<int64_t>(num_nodes, 0x7fffffff);
    const void* inst = nullptr;
    if ((rc = rrb_get_instances(c->api, &inst, nullptr))) return pfail(rc, rrb_last_error());
    c->st.instances = (const float4*)inst;
    static const int stage_top = getenv("MCRT_STAGE_TOP") ? atoi(getenv("MCRT_STAGE_TOP")) : 1;
    c->st.stageTop = stage_top;
    return 0;
}

int ready(ptb_context* c) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    if (!c->scene_loaded) return pfail(RRB_ERROR_INVALID_ARGUMENT, "ptb_upload_scene has not been called");
    if (c->width <= 0) return pfail(RRB_ERROR_INVALID_ARGUMENT, "ptb_resize has not been called");
    int rc = set_dev(c);
    if (rc) return rc;
    return refresh_bvh(c);
}

// ---- frames in flight
// The split shade stage re-reads Kr, Kt, eta and the roughness from the material table in its second kernel
// (pt_shade.cuh: untexturedUberProperties): exact unless a material binds a texture for one of them.
bool materials_allow_split(const void* materials, int n) {
    const unsigned char* m = (const unsigned char*)materials;
    for (int i = 0; i < n; ++i) {
        int ids[8];      // normalMap, diffuse, glossy, specReflection, transmission, opacity, roughness, ior (kernel_data.h:87-113)
        std::memcpy(ids, m + (size_t)i * 128 + 96, sizeof(ids));
        if (ids[3] != -1 || ids[4] != -1 || ids[6] != -1 || ids[7] != -1) return false;
    }
    return true;
}
void apply_shade_split(ptb_context* c) { c->st.shade_split = c->split_supported ? c->shade_split_wanted : 0; }

bool pipelined(const ptb_context* c) { return c->slots.size() > 1 && !c->profiling && !c->st.path_trace; }

// the context's stream waits (on the device) for every frame rendered on a side stream
int join_frames(ptb_context* c) {
    for (ptb_context::FrameSlot& s : c->slots)
        if (s.done_recorded) PTB_CUDA(cudaStreamWaitEvent(c->stream, s.render_done, 0));
    return 0;
}
// ... and forgets them: whatever follows works on the context's own state
int join_and_drop_frames(ptb_context* c) {
    int rc = join_frames(c);
    if (rc) return rc;
    c->rendered.clear();
    auto it = c->named.find("RadianceBufferCL");
    if (it != c->named.end()) it->second.p = c->st.radiance;
    return 0;
}

void destroy_slots(ptb_context* c) {
    for (ptb_context::FrameSlot& s : c->slots) {
        if (s.stream) {
            cudaStreamSynchronize(s.stream);
            cudaStreamDestroy(s.stream);
        }
        if (s.render_done) cudaEventDestroy(s.render_done);
        if (s.slot_free) cudaEventDestroy(s.slot_free);
        if (s.overlap.stream) {
            cudaStreamSynchronize(s.overlap.stream);
            cudaStreamDestroy(s.overlap.stream);
            cudaEventDestroy(s.overlap.fork);
            cudaEventDestroy(s.overlap.join);
            cudaEventDestroy(s.overlap.stagger);
        }
    }
    c->slots.clear();
    c->rendered.clear();
    c->next_slot = 0;
    for (auto it = c->pixel_bufs.begin(); it != c->pixel_bufs.end();) {
        if (it->first.rfind("slot", 0) == 0) {
            if (it->second.p) cudaFree(it->second.p);
            it = c->pixel_bufs.erase(it);
        } else {
            ++it;
        }
    }
}

// per-slot wavefront state (slot 0 reuses the context's own arrays)
int alloc_slots(ptb_context* c, int depth) {
    destroy_slots(c);
    if (depth <= 1) return 0;
    c->slots.resize((size_t)depth);
    const size_t P = (size_t)c->width * c->height;
    for (int k = 0; k < depth; ++k) {
        ptb_context::FrameSlot& s = c->slots[(size_t)k];
        PTB_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        PTB_CUDA(cudaEventCreateWithFlags(&s.render_done, cudaEventDisableTiming));
        PTB_CUDA(cudaEventCreateWithFlags(&s.slot_free, cudaEventDisableTiming));
        PTB_CUDA(cudaStreamCreateWithFlags(&s.overlap.stream, cudaStreamNonBlocking));
        PTB_CUDA(cudaEventCreateWithFlags(&s.overlap.fork, cudaEventDisableTiming));
        PTB_CUDA(cudaEventCreateWithFlags(&s.overlap.join, cudaEventDisableTiming));
        PTB_CUDA(cudaEventCreateWithFlags(&s.overlap.stagger, cudaEventDisableTiming));
        s.st = c->st;
        const std::string pre = "slot" + std::to_string(k) + "_";
        void* p = nullptr;
        int rc;
        if ((rc = alloc_into(c->pixel_bufs, pre + "camera", 176, true, c->stream, &p))) return rc;
        s.camera = p;
        if ((rc = alloc_into(c->pixel_bufs, pre + "cnt", 2 * mcrt::PT_CNT_STRIDE * sizeof(unsigned int), true, c->stream, &p)))
            return rc;
        s.st.cnt = (unsigned int*)p;
        if (k == 0 || P == 0) continue;
        struct {
            const char* name;
            size_t elem;
            void** dst;
        } specs[] = {
            {"ray_o", 16, (void**)&s.st.ray_o},   {"ray_d", 16, (void**)&s.st.ray_d},
            {"hit", 16, (void**)&s.st.hit},       {"throughput", 16, (void**)&s.st.throughput},
            {"radiance", 16, (void**)&s.st.radiance}, {"direct_weight", 16, (void**)&s.st.direct_weight},
            {"queue0", 4, (void**)&s.st.queue[0]}, {"queue1", 4, (void**)&s.st.queue[1]},
            {"bucket", 4 * mcrt::PT_NUM_BUCKETS, (void**)&s.st.bucket},
            {"vtx", 7 * 16, (void**)&s.st.vtx},
            {"strag", 8, (void**)&s.st.strag},
            {"sh_o", 16, (void**)&s.st.sh_o},     {"sh_d", 16, (void**)&s.st.sh_d},
            {"sh_L", 16, (void**)&s.st.sh_L},
        };
        for (auto& sp : specs) {
            if ((rc = alloc_into(c->pixel_bufs, pre + sp.name, P * sp.elem, true, c->stream, &p))) return rc;
            *sp.dst = p;
        }
    }
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

// the PtState a slot renders with: the context's scene / options, the slot's wavefront arrays
mcrt::PtState slot_state(const ptb_context* c, int k) {
    const ptb_context::FrameSlot& s = c->slots[(size_t)k];
    mcrt::PtState st = c->st;
    st.camera = s.camera;
    st.cnt = s.st.cnt;
    if (k > 0) {
        st.ray_o = s.st.ray_o;
        st.ray_d = s.st.ray_d;
        st.hit = s.st.hit;
        st.throughput = s.st.throughput;
        st.radiance = s.st.radiance;
        st.direct_weight = s.st.direct_weight;
        st.queue[0] = s.st.queue[0];
        st.queue[1] = s.st.queue[1];
        st.bucket = s.st.bucket;
        st.vtx = s.st.vtx;
        st.strag = s.st.strag;
        st.sh_o = s.st.sh_o;
        st.sh_d = s.st.sh_d;
        st.sh_L = s.st.sh_L;
    }
    return st;
}

// 64x64 tiles dealt round-robin in row-major tile order (partition A)
void partition_pixels(int width, int height, int part, int num_parts, std::vector<uint32_t>& px) {
    const int tw = (width + 63) / 64, th = (height + 63) / 64;
    for (int ty = 0; ty < th; ++ty)
        for (int tx = 0; tx < tw; ++tx) {
            if ((ty * tw + tx) % num_parts != part) continue;
            for (int y = ty * 64; y < std::min(height, ty * 64 + 64); ++y)
                for (int x = tx * 64; x < std::min(width, tx * 64 + 64); ++x) px.push_back((uint32_t)(y * width + x));
        }
}

int build_partition(ptb_context* c) {
    if (c->d_pixels) {
        cudaFree(c->d_pixels);
        c->d_pixels = nullptr;
    }
    const int P = c->width * c->height;
    if (c->num_parts <= 1) {
        c->n_pixels = P;
        return 0;
    }
    std::vector<uint32_t> px;
    partition_pixels(c->width, c->height, c->part, c->num_parts, px);
    c->n_pixels = (int)px.size();
    if (!px.empty()) {
        PTB_CUDA(cudaMalloc((void**)&c->d_pixels, px.size() * 4));
        PTB_CUDA(cudaMemcpyAsync(c->d_pixels, px.data(), px.size() * 4, cudaMemcpyHostToDevice, c->stream));
        PTB_CUDA(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

}  // namespace

extern "C" {

const char* ptb_last_error(void) { return g_ptb_error.c_str(); }

int ptb_create(rrb_api* api, ptb_context** out) {
    if (!api || !out) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    std::unique_ptr<ptb_context> c(new ptb_context());
    c->api = api;
    void* stream = nullptr;
    int rc = rrb_get_stream(api, &c->device, &stream);
    if (rc) return pfail(rc, rrb_last_error());
    c->stream = (cudaStream_t)stream;
    rc = set_dev(c.get());
    if (rc) return rc;
    void* p = nullptr;
    rc = alloc_into(c->named, "__cnt", 2 * mcrt::PT_CNT_STRIDE * sizeof(unsigned int), true, c->stream, &p);
    if (rc) return rc;
    c->st.cnt = (unsigned int*)p;
    rc = alloc_into(c->named, "__stats", 64, true, c->stream, &p);
    if (rc) return rc;
    c->st.stats = (unsigned long long*)p;
    rc = alloc_into(c->named, "__camera", 176, true, c->stream, &p);
    if (rc) return rc;
    c->st.camera = p;
    if (const char* v = getenv("MCRT_OVERLAP_HALVES")) c->halves = atoi(v) == 1 ? 1 : 2;   // profiling aid
    c->st.coherence = PTB_DEFAULT_COHERENCE;
    if (const char* v = getenv("MCRT_COHERENCE")) c->st.coherence = std::max(0, std::min(4, atoi(v)));   // variant runs
    c->st.coopIters = PTB_DEFAULT_COOPERATIVE_TAIL;
    if (const char* v = getenv("MCRT_COOP_ITERS")) c->st.coopIters = std::max(0, atoi(v));                      // variant runs
    c->shade_split_wanted = PTB_DEFAULT_SHADE_SPLIT;
    if (const char* v = getenv("MCRT_SHADE_SPLIT")) c->shade_split_wanted = std::max(0, std::min(2, atoi(v)));                  // variant runs
    PTB_CUDA(cudaStreamCreateWithFlags(&c->overlap.stream, cudaStreamNonBlocking));
    PTB_CUDA(cudaEventCreateWithFlags(&c->overlap.fork, cudaEventDisableTiming));
    PTB_CUDA(cudaEventCreateWithFlags(&c->overlap.join, cudaEventDisableTiming));
    PTB_CUDA(cudaEventCreateWithFlags(&c->overlap.stagger, cudaEventDisableTiming));
    *out = c.release();
    return 0;
}

int ptb_set_overlap(ptb_context* c, int halves) {
    if (!c || (halves != 1 && halves != 2)) return pfail(RRB_ERROR_INVALID_ARGUMENT, "halves must be 1 or 2");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->halves = halves;
    return 0;
}

int ptb_set_estimator(ptb_context* c, int flags) {
    if (!c || (flags & ~(PTB_EST_RUSSIAN_ROULETTE | PTB_EST_NEE_MIS)))
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "unknown estimator flag");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->st.estimator = flags;
    return 0;
}

int ptb_set_coherence(ptb_context* c, int mode) {
    if (!c || mode < 0 || mode > 4) return pfail(RRB_ERROR_INVALID_ARGUMENT, "coherence mode must be 0 .. 4");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->st.coherence = mode;
    return 0;
}

int ptb_set_cooperative_tail(ptb_context* c, int min_iters) {
    if (!c || min_iters < 0) return pfail(RRB_ERROR_INVALID_ARGUMENT, "min_iters must be >= 0");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->st.coopIters = min_iters;
    return 0;
}

int ptb_set_shade_split(ptb_context* c, int on) {
    if (!c || on < 0 || on > 2) return pfail(RRB_ERROR_INVALID_ARGUMENT, "shade split must be 0, 1 or 2");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->shade_split_wanted = on;
    apply_shade_split(c);
    return 0;
}

int ptb_destroy(ptb_context* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    destroy_slots(c);
    free_map(c->scene_bufs);
    free_map(c->pixel_bufs);
    for (const char* k : {"__cnt", "__stats", "__camera"}) {
        auto it = c->named.find(k);
        if (it != c->named.end() && it->second.p) cudaFree(it->second.p);
    }
    if (c->d_pixels) cudaFree(c->d_pixels);
    if (c->d_sobol) cudaFree((void*)c->d_sobol);
    c->timer.destroy();
    if (c->copy_stream) {
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamDestroy(c->copy_stream);
        for (int k = 0; k < 2; ++k) {
            cudaEventDestroy(c->snap_ready[k]);
            cudaEventDestroy(c->copy_done[k]);
            if (c->snapshot[k]) cudaFree(c->snapshot[k]);
        }
    }
    if (c->overlap.stream) {
        cudaStreamSynchronize(c->overlap.stream);
        cudaStreamDestroy(c->overlap.stream);
        cudaEventDestroy(c->overlap.fork);
        cudaEventDestroy(c->overlap.join);
        cudaEventDestroy(c->overlap.stagger);
    }
    delete c;
    return 0;
}

int ptb_upload_scene(ptb_context* c, const ptb_scene_desc* d) {
    if (!c || !d) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    free_map(c->scene_bufs);
    c->scene_loaded = false;
    const void* p = nullptr;
    if ((rc = upload(c, "shapes", d->shapes, (size_t)d->num_shapes * 160, &p))) return rc;
    c->st.shapes = p;
    if ((rc = upload(c, "indices", d->indices, (size_t)d->num_indices * 4, &p))) return rc;
    c->st.indices = (const uint32_t*)p;
    if ((rc = upload(c, "positions", d->positions, (size_t)d->num_vertices * 16, &p))) return rc;
    c->st.positions = (const float4*)p;
    if ((rc = upload(c, "uvs", d->uvs, (size_t)d->num_vertices * 8, &p))) return rc;
    c->st.uvs = (const float2*)p;
    if ((rc = upload(c, "normals", d->normals, (size_t)d->num_vertices * 16, &p))) return rc;
    c->st.normals = (const float4*)p;
    if ((rc = upload(c, "textures", d->textures, (size_t)d->num_textures * 16, &p))) return rc;
    c->st.textures2D = p;
    if ((rc = upload(c, "texdata", d->texdata, (size_t)d->texdata_bytes, &p))) return rc;
    c->st.texData2D = (const unsigned char*)p;
    if ((rc = upload(c, "lights", d->lights, (size_t)d->num_lights * 80, &p))) return rc;
    c->st.lights = p;
    if ((rc = upload(c, "materials", d->materials, (size_t)d->num_materials * 128, &p))) return rc;
    c->st.materials = p;
    {
        void* a = nullptr;
        // (rounded up to the 16 bytes a bulk copy into shared memory moves: pt_kernels.cu, staged_tables)
        if ((rc = alloc_into(c->scene_bufs, "material_alpha", ((size_t)std::max(d->num_materials, 1) * sizeof(float2) + 15) / 16 * 16, true,
                             c->stream, &a)))
            return rc;
        c->st.material_alpha = (const float2*)a;
        c->num_materials = d->num_materials;
        PTB_CUDA(mcrt::launch_material_alpha(c->st.materials, (float2*)a, d->num_materials, c->stream));
        c->split_supported = materials_allow_split(d->materials, d->num_materials);
        apply_shade_split(c);
    }
    // per-triangle shading records: index-array order, so a hit's (shape, prim) addresses startIdx / 3 + prim
    c->st.tri_shade = nullptr;
    c->num_scene_tris = 0;
    c->d_tri_shape = nullptr;
    // off by default: measured neutral for the default shade kernel (1.991 vs 1.998 ms per step), +6 % for the sorted
    // one (profiles/r02_shade.md); MCRT_TRI_RECORDS=1 or ptb_set_option-style builds switch it on
    static const int tri_records = getenv("MCRT_TRI_RECORDS") ? atoi(getenv("MCRT_TRI_RECORDS")) : 0;
    if (tri_records && d->num_shapes > 0 && d->num_indices >= 3) {
        struct ShapeHdr {      // RTShape, kernel_data.h:36-52
            float m[32];
            uint32_t startIdx, startVertex, numTriangles;
            int materialId, lightID;
            float area;
            int pad[2];
        };
        static_assert(sizeof(ShapeHdr) == 160, "RTShape");
        const ShapeHdr* sh = (const ShapeHdr*)d->shapes;
        const int64_t n_tris = d->num_indices / 3;
        std::vector<int> tri_shape((size_t)n_tris, -1);
        bool ok = true;
        for (int i = 0; i < d->num_shapes && ok; ++i) {
            if (sh[i].startIdx % 3u) ok = false;
            const int64_t first = sh[i].startIdx / 3u;
            if (first + (int64_t)sh[i].numTriangles > n_tris) ok = false;
            for (int64_t t = first; ok && t < first + (int64_t)sh[i].numTriangles; ++t) {
                if (tri_shape[(size_t)t] != -1) ok = false;      // shapes sharing index ranges: records would be ambiguous
                tri_shape[(size_t)t] = i;
            }
        }
        for (int64_t t = 0; ok && t < n_tris; ++t)
            if (tri_shape[(size_t)t] < 0) tri_shape[(size_t)t] = 0, ok = ok;   // unreferenced triangles: any shape, never read
        if (ok) {
            const void* ts = nullptr;
            if ((rc = upload(c, "tri_shape", tri_shape.data(), (size_t)n_tris * 4, &ts))) return rc;
            void* rec = nullptr;
            if ((rc = alloc_into(c->scene_bufs, "tri_shade", (size_t)n_tris * 144, false, c->stream, &rec))) return rc;
            c->d_tri_shape = (const int*)ts;
            c->num_scene_tris = (int)n_tris;
            PTB_CUDA(mcrt::launch_build_tri_shade(c->st, c->d_tri_shape, (float4*)rec, c->num_scene_tris, c->stream));
            c->st.tri_shade = (const float4*)rec;
        }
    }
    c->st.numLights = d->num_lights;
    c->st.numShapes = d->num_shapes;
    c->st.numMaterials = d->num_materials;
    c->st.numTextures = d->num_textures;
    if (d->sobol && (rc = set_sobol_table(c, d->sobol, 1024))) return rc;   // the engine's table (scene/RTScene.cpp:55)
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->scene_loaded = true;
    return 0;
}

int ptb_update_scene_array(ptb_context* c, const char* array, const void* data, size_t bytes) {
    if (!c || !array || !data) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    if (!c->scene_loaded) return pfail(RRB_ERROR_INVALID_ARGUMENT, "ptb_upload_scene has not been called");
    const std::string name(array);
    if (name != "shapes" && name != "lights" && name != "materials")
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "only shapes, lights and materials can be updated in place");
    auto it = c->scene_bufs.find(name);
    if (it == c->scene_bufs.end() || it->second.bytes != bytes || bytes == 0)
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "size of '" + name + "' differs from the uploaded array");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaMemcpyAsync(it->second.p, data, bytes, cudaMemcpyHostToDevice, c->stream));
    if (name == "shapes" && c->st.tri_shade)      // transforms changed: the per-triangle records follow
        PTB_CUDA(mcrt::launch_build_tri_shade(c->st, c->d_tri_shape, (float4*)c->st.tri_shade, c->num_scene_tris, c->stream));
    if (name == "materials") {
        PTB_CUDA(mcrt::launch_material_alpha(c->st.materials, (float2*)c->st.material_alpha, c->num_materials, c->stream));
        c->split_supported = materials_allow_split(data, c->num_materials);
        apply_shade_split(c);
    }
    if (!c->slots.empty()) PTB_CUDA(cudaStreamSynchronize(c->stream));   // frames on side streams must see the update
    return 0;
}

int ptb_set_sampler(ptb_context* c, int kind, const uint32_t* matrices, int num_dimensions) {
    if (!c || (kind != 0 && kind != 1)) return pfail(RRB_ERROR_INVALID_ARGUMENT, "sampler kind must be 0 (Sobol) or 1 (random)");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    if (matrices) {
        if (num_dimensions < 1) return pfail(RRB_ERROR_INVALID_ARGUMENT, "num_dimensions must be >= 1");
        if ((rc = set_sobol_table(c, matrices, num_dimensions))) return rc;
    }
    if (kind == 0 && !c->d_sobol) return pfail(RRB_ERROR_INVALID_ARGUMENT, "the Sobol sampler needs its generator matrices");
    c->sampler_kind = kind;
    c->st.sobol = kind == 0 ? c->d_sobol : nullptr;
    return 0;
}

int ptb_resize(ptb_context* c, int width, int height) {
    if (!c || width <= 0 || height <= 0) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad size");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    const int frames_in_flight = (int)c->slots.size();
    destroy_slots(c);
    free_map(c->pixel_bufs);
    c->st.path_trace = nullptr;
    c->path_trace_depth = 0;
    c->named.erase("PathTraceBuffer");
    c->named.erase("BDPTSampledLightBuffer");
    for (const char* k : {"BDPTConnectionRayO", "BDPTConnectionRayD", "BDPTConnectionContribution", "BDPTConnectionPixel",
                          "BDPTConnectionVisibility", "BDPTConnectionWeight", "BDPTCounters"})
        c->named.erase(k);
    c->bdpt_depth = -1;
    std::memset(&c->bd, 0, sizeof(c->bd));
    c->width = width;
    c->height = height;
    c->st.width = width;
    c->st.height = height;
    c->st.vtx_stride = width * height;
    const size_t P = (size_t)width * height;
    struct Spec {
        const char* name;
        size_t elem;
        void** dst;
        const char* public_name;
    };
    mcrt::PtState& s = c->st;
    const Spec specs[] = {
        {"ray_o", 16, (void**)&s.ray_o, nullptr},
        {"ray_d", 16, (void**)&s.ray_d, nullptr},
        {"hit", 16, (void**)&s.hit, nullptr},
        {"throughput", 16, (void**)&s.throughput, nullptr},
        {"radiance", 16, (void**)&s.radiance, "RadianceBufferCL"},
        {"direct_weight", 16, (void**)&s.direct_weight, nullptr},
        {"queue0", 4, (void**)&s.queue[0], nullptr},
        {"queue1", 4, (void**)&s.queue[1], nullptr},
        {"bucket", 4 * mcrt::PT_NUM_BUCKETS, (void**)&s.bucket, nullptr},
        {"vtx", 7 * 16, (void**)&s.vtx, nullptr},
        {"strag", 8, (void**)&s.strag, nullptr},
        {"sh_o", 16, (void**)&s.sh_o, nullptr},
        {"sh_d", 16, (void**)&s.sh_d, nullptr},
        {"sh_L", 16, (void**)&s.sh_L, nullptr},
        {"aos_rays", 48, &s.aos_rays, "RayBuffer"},
        {"aos_isect", 32, &s.aos_isect, "PrimaryIntersectionBufferCL"},
        {"aos_shadow_rays", 48, &s.aos_shadow_rays, "ShadowRayBuffer"},
        {"aos_occlusion", 4, (void**)&s.aos_occlusion, "ShadowRayOcclusionBuffer"},
        {"aos_temp_radiance", 16, (void**)&s.aos_temp_radiance, "TempRadianceBuffer"},
        {"aos_throughput", 32, &s.aos_throughput, "ThroughputBuffer"},
        {"image", 16, (void**)&s.image, "NextFrameImageCL"},
        {"denoised", 16, (void**)&c->denoised, "DenoisedImageCL"},
        {"tonemapped", 16, (void**)&c->tonemapped, "TonemappedImageCL"},
    };
    for (const Spec& sp : specs) {
        void* p = nullptr;
        // zero-filled at creation: defines what the reference leaves uninitialised
        if ((rc = alloc_into(c->pixel_bufs, sp.name, P * sp.elem, true, c->stream, &p))) return rc;
        *sp.dst = p;
        if (sp.public_name) {
            DevBuf v;
            v.p = p;
            v.bytes = P * sp.elem;
            c->named[sp.public_name] = v;
        }
    }
    {
        // sum(w L) (float4[P]) and sum(w) (float[P]) are ONE 20 B/pixel allocation: the multi-GPU merge reduces it
        // with a single NCCL call and no staging copy (ptb_merge_accumulation)
        void* p = nullptr;
        if ((rc = alloc_into(c->pixel_bufs, "accumulation", P * 20, true, c->stream, &p))) return rc;
        s.weighted_radiance = (float4*)p;
        s.filter_weights = (float*)((char*)p + P * 16);
        DevBuf v;
        v.p = p;
        v.bytes = P * 16;
        c->named["WeightedRadianceBuffer"] = v;
        v.p = s.filter_weights;
        v.bytes = P * 4;
        c->named["FilterWeightsBuffer"] = v;
        v.p = p;
        v.bytes = P * 20;
        c->named["AccumulationBuffer"] = v;
        c->merged = nullptr;
    }
    rc = build_partition(c);
    if (rc) return rc;
    if ((rc = alloc_slots(c, frames_in_flight))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int ptb_set_frames_in_flight(ptb_context* c, int depth) {
    if (!c || depth < 1 || depth > 4) return pfail(RRB_ERROR_INVALID_ARGUMENT, "frames in flight must be 1 .. 4");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return alloc_slots(c, depth);
}

int ptb_set_camera(ptb_context* c, const void* cam) {
    if (!c || !cam) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    int rc = set_dev(c);
    if (rc) return rc;
    // pageable source: the copy is staged before the call returns, so the caller may reuse `cam`
    PTB_CUDA(cudaMemcpyAsync((void*)c->st.camera, cam, 176, cudaMemcpyHostToDevice, c->stream));
    // frames in flight take their own copy on their own stream when they are rendered: it must not queue up behind
    // the accumulation of the previous frame on the context's stream
    std::memcpy(c->camera_host, cam, 176);
    c->camera_version += 1;
    return 0;
}

int ptb_partition_pixels(int width, int height, int part, int num_parts, uint32_t* out, int64_t capacity, int64_t* count) {
    if (width <= 0 || height <= 0 || num_parts < 1 || part < 0 || part >= num_parts || !count)
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad partition");
    std::vector<uint32_t> px;
    partition_pixels(width, height, part, num_parts, px);
    *count = (int64_t)px.size();
    if (out) {
        if (capacity < (int64_t)px.size()) return pfail(RRB_ERROR_INVALID_ARGUMENT, "pixel list capacity too small");
        std::memcpy(out, px.data(), px.size() * sizeof(uint32_t));
    }
    return 0;
}

int ptb_set_partition(ptb_context* c, int part, int num_parts) {
    if (!c || num_parts < 1 || part < 0 || part >= num_parts) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad partition");
    c->part = part;
    c->num_parts = num_parts;
    if (c->width > 0) {
        int rc = set_dev(c);
        if (rc) return rc;
        if ((rc = join_and_drop_frames(c))) return rc;
        PTB_CUDA(cudaStreamSynchronize(c->stream));
        return build_partition(c);
    }
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// stage / frame entry points
// ------------------------------------------------------------------------------------------------
extern "C" {

int ptb_generate_primary_rays(ptb_context* c) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(mcrt::launch_generate_primary_rays(c->st, c->stream));
    c->launches += 1;
    return 0;
}

int ptb_pt_shade(ptb_context* c, int frame, int max_depth, int bounce) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(mcrt::launch_pt_shade_aos(c->st, frame, max_depth, bounce, c->stream));
    c->launches += 1;
    return 0;
}

int ptb_pt_shadow_resolve(ptb_context* c, int bounce) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(mcrt::launch_pt_shadow_resolve_aos(c->st, bounce, c->stream));
    c->launches += 1;
    return 0;
}

int ptb_pt_render_sample(ptb_context* c, int frame, int max_depth) {
    int rc = ready(c);
    if (rc) return rc;
    c->last_frame_bdpt = false;
    if (max_depth < 1) return pfail(RRB_ERROR_INVALID_ARGUMENT, "max_depth must be >= 1");
    cudaError_t e = cudaSuccess;
    if (pipelined(c)) {
        const int k = c->next_slot;
        ptb_context::FrameSlot& s = c->slots[(size_t)k];
        // a slot whose frame was never accumulated is simply rendered over, as the one radiance buffer would be
        for (auto it = c->rendered.begin(); it != c->rendered.end();)
            it = (*it == k) ? c->rendered.erase(it) : it + 1;
        if (s.free_recorded) PTB_CUDA(cudaStreamWaitEvent(s.stream, s.slot_free, 0));
        if (s.camera_version != c->camera_version) {
            PTB_CUDA(cudaMemcpyAsync(s.camera, c->camera_host, 176, cudaMemcpyHostToDevice, s.stream));
            s.camera_version = c->camera_version;
        }
        const mcrt::PtState st = slot_state(c, k);
        c->launches += mcrt::launch_pt_sample(st, c->d_pixels, c->n_pixels, frame, max_depth, s.stream, &e, nullptr,
                                              c->halves == 2 ? &s.overlap : nullptr);
        if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("path tracer launch: ") + cudaGetErrorString(e));
        PTB_CUDA(cudaEventRecord(s.render_done, s.stream));
        s.done_recorded = true;
        c->rendered.push_back(k);
        c->named["RadianceBufferCL"].p = st.radiance;
        c->next_slot = (k + 1) % (int)c->slots.size();
        return 0;
    }
    if ((rc = join_and_drop_frames(c))) return rc;
    if (c->st.path_trace) {
        if (max_depth > c->path_trace_depth) return pfail(RRB_ERROR_INVALID_ARGUMENT, "max_depth exceeds the depth given to ptb_set_path_trace");
        PTB_CUDA(cudaMemsetAsync(c->st.path_trace, 0, (size_t)c->path_trace_depth * c->width * c->height * sizeof(int4), c->stream));
    }
    c->launches += mcrt::launch_pt_sample(c->st, c->d_pixels, c->n_pixels, frame, max_depth, c->stream, &e,
                                          c->profiling ? &c->timer : nullptr, c->halves == 2 ? &c->overlap : nullptr);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("path tracer launch: ") + cudaGetErrorString(e));
    return 0;
}

int ptb_set_path_trace(ptb_context* c, int max_depth) {
    if (!c || max_depth < 0 || max_depth > 64) return pfail(RRB_ERROR_INVALID_ARGUMENT, "path trace depth must be 0 .. 64");
    if (c->width <= 0) return pfail(RRB_ERROR_INVALID_ARGUMENT, "ptb_resize has not been called");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    auto it = c->pixel_bufs.find("path_trace");
    if (it != c->pixel_bufs.end()) {
        if (it->second.p) cudaFree(it->second.p);
        c->pixel_bufs.erase(it);
    }
    c->named.erase("PathTraceBuffer");
    c->st.path_trace = nullptr;
    c->path_trace_depth = 0;
    if (max_depth == 0) return 0;
    void* p = nullptr;
    const size_t bytes = (size_t)max_depth * c->width * c->height * sizeof(int4);
    if ((rc = alloc_into(c->pixel_bufs, "path_trace", bytes, true, c->stream, &p))) return rc;
    c->st.path_trace = (int4*)p;
    c->path_trace_depth = max_depth;
    DevBuf v;
    v.p = p;
    v.bytes = bytes;
    c->named["PathTraceBuffer"] = v;
    return 0;
}

int ptb_profile_enable(ptb_context* c, int on) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    c->profiling = on != 0;
    c->timer.reset();
    return 0;
}

static int profile_collect(ptb_context* c, double* acc, int64_t* n) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    for (int k = 0; k < mcrt::StageTimer::NUM_STAGES; ++k) {
        acc[k] = 0.0;
        n[k] = 0;
    }
    for (const mcrt::StageTimer::Lane& l : c->timer.lanes)
        for (size_t i = 1; i < l.used; ++i) {
            const int tag = l.tags[i];
            if (tag < 0) continue;
            float t = 0.0f;
            PTB_CUDA(cudaEventElapsedTime(&t, l.events[i - 1], l.events[i]));
            acc[tag] += t;
            n[tag] += 1;
        }
    c->timer.reset();
    return 0;
}

int ptb_profile_collect(ptb_context* c, double ms[4], int64_t launches[4]) {
    double acc[mcrt::StageTimer::NUM_STAGES];
    int64_t n[mcrt::StageTimer::NUM_STAGES];
    int rc = profile_collect(c, acc, n);
    if (rc) return rc;
    acc[PTB_STAGE_SHADE] += acc[PTB_STAGE_SHADE_SCATTER];      // the shade stage as a whole; one launch = one bounce
    for (int k = 0; k < 4; ++k) {
        if (ms) ms[k] = acc[k];
        if (launches) launches[k] = n[k];
    }
    return 0;
}

int ptb_profile_collect_kernels(ptb_context* c, double ms[PTB_NUM_STAGES], int64_t launches[PTB_NUM_STAGES]) {
    double acc[mcrt::StageTimer::NUM_STAGES];
    int64_t n[mcrt::StageTimer::NUM_STAGES];
    int rc = profile_collect(c, acc, n);
    if (rc) return rc;
    for (int k = 0; k < PTB_NUM_STAGES; ++k) {
        if (ms) ms[k] = acc[k];
        if (launches) launches[k] = n[k];
    }
    return 0;
}

int ptb_bdpt_reserve(ptb_context* c, int max_depth) {
    int rc = ready(c);
    if (rc) return rc;
    if (c->st.instances)
        return pfail(RRB_ERROR_INVALID_ARGUMENT,
                     "the BDPT kernels trace the flat tree only: commit the world with the option bvh.forceflat = 1");
    if ((rc = join_and_drop_frames(c))) return rc;
    if (max_depth < 1) return pfail(RRB_ERROR_INVALID_ARGUMENT, "max_depth must be >= 1");
    if (c->bdpt_depth == max_depth) return 0;
    // RTBDPTPass::createBuffers (RTBDPTPass.cpp:451-478): sized by maxDepth, zero-filled (E12)
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    for (const char* k : mcrt::bdpt_buffer_names()) {
        auto it = c->pixel_bufs.find(k);
        if (it != c->pixel_bufs.end()) {
            if (it->second.p) cudaFree(it->second.p);
            c->pixel_bufs.erase(it);
        }
    }
    const size_t P = (size_t)c->width * c->height;
    for (const mcrt::BdptBufferSpec& sp : mcrt::bdpt_buffer_specs(c->bd, max_depth)) {
        void* p = nullptr;
        const size_t bytes = sp.bytes_per_pixel ? P * sp.bytes_per_pixel : 64;
        if ((rc = alloc_into(c->pixel_bufs, sp.name, bytes, true, c->stream, &p))) return rc;
        *sp.dst = p;
        if (std::string(sp.name) == "bdpt_sampled_light") {     // the one BDPT buffer whose old contents are read (E12)
            DevBuf v;
            v.p = p;
            v.bytes = bytes;
            c->named["BDPTSampledLightBuffer"] = v;
        }
        // the per-strategy connection records of the frame rendered last (include/ptb.h)
        static const char* const kPublic[][2] = {{"bdpt_conn_o", "BDPTConnectionRayO"}, {"bdpt_conn_d", "BDPTConnectionRayD"},
                                                 {"bdpt_conn_C", "BDPTConnectionContribution"}, {"bdpt_conn_src", "BDPTConnectionPixel"},
                                                 {"bdpt_conn_vis", "BDPTConnectionVisibility"}, {"bdpt_conn_w", "BDPTConnectionWeight"},
                                                 {"bdpt_cnt", "BDPTCounters"}};
        for (const auto& pub : kPublic)
            if (std::string(sp.name) == pub[0]) {
                DevBuf v;
                v.p = p;
                v.bytes = bytes;
                c->named[pub[1]] = v;
            }
    }
    c->bdpt_depth = max_depth;
    return 0;
}

int ptb_bdpt_render_sample(ptb_context* c, int frame, int max_depth) {
    int rc = ptb_bdpt_reserve(c, max_depth);
    if (rc) return rc;
    cudaError_t e = cudaSuccess;
    c->launches += mcrt::launch_bdpt_sample(c->st, c->bd, c->d_pixels, c->n_pixels, frame, max_depth, c->stream, &e);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("bdpt launch: ") + cudaGetErrorString(e));
    c->last_frame_bdpt = true;
    return 0;
}

int ptb_accumulate(ptb_context* c, int frame, const ptb_filter_props* f) {
    int rc = ready(c);
    if (rc) return rc;
    if (!f) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null filter properties");
    mcrt::FilterProps fp;
    fp.filter_type = f->filter_type;
    fp.radius[0] = f->radius[0];
    fp.radius[1] = f->radius[1];
    fp.mitchell_b = f->mitchell_b;
    fp.mitchell_c = f->mitchell_c;
    fp.lanczos_sinc_tau = f->lanczos_sinc_tau;
    fp.gaussian_alpha = f->gaussian_alpha;
    fp.gaussian_exp_x = f->gaussian_exp_x;
    fp.gaussian_exp_y = f->gaussian_exp_y;
    fp.pixel_offset[0] = f->pixel_offset[0];
    fp.pixel_offset[1] = f->pixel_offset[1];
    if (!c->rendered.empty()) {
        // the oldest frame still waiting on a side stream: accumulation keeps the order of the render calls
        const int k = c->rendered.front();
        c->rendered.pop_front();
        ptb_context::FrameSlot& s = c->slots[(size_t)k];
        PTB_CUDA(cudaStreamWaitEvent(c->stream, s.render_done, 0));
        PTB_CUDA(mcrt::launch_accumulate(slot_state(c, k), fp, frame, c->d_pixels, c->n_pixels, c->stream));
        PTB_CUDA(cudaEventRecord(s.slot_free, c->stream));
        s.free_recorded = true;
    } else if (c->last_frame_bdpt && c->num_parts > 1) {
        PTB_CUDA(mcrt::launch_accumulate_splat(c->st, fp, frame, c->part, c->num_parts, c->stream));
    } else {
        PTB_CUDA(mcrt::launch_accumulate(c->st, fp, frame, c->d_pixels, c->n_pixels, c->stream));
    }
    c->launches += 1;
    return 0;
}

namespace {
int named_image(ptb_context* c, const char* name, float4** out) {
    auto it = c->named.find(name ? name : "");
    const size_t P = (size_t)c->width * c->height;
    if (it == c->named.end() || it->second.bytes != P * 16)
        return pfail(RRB_ERROR_INVALID_ARGUMENT, std::string("not a float4[P] image buffer: ") + (name ? name : "(null)"));
    *out = (float4*)it->second.p;
    return 0;
}
}  // namespace

int ptb_denoise(ptb_context* c, const char* src_name, int radius, float sigma_spatial, float sigma_range) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    float4* src = nullptr;
    if ((rc = named_image(c, src_name, &src))) return rc;
    if (src == c->denoised) return pfail(RRB_ERROR_INVALID_ARGUMENT, "denoise cannot run in place");
    if (radius > 32) return pfail(RRB_ERROR_INVALID_ARGUMENT, "denoise radius > 32");
    cudaError_t e = mcrt::launch_bilateral_denoise(c->width, c->height, radius, sigma_spatial, sigma_range, src,
                                                   c->denoised, c->stream);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("denoise launch: ") + cudaGetErrorString(e));
    c->launches += 1;
    return 0;
}

int ptb_tonemap(ptb_context* c, const char* src_name, float lwhite) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    float4* src = nullptr;
    if ((rc = named_image(c, src_name, &src))) return rc;
    cudaError_t e = mcrt::launch_reinhard_tonemap(c->width, c->height, lwhite, src, c->tonemapped, c->stream);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("tonemap launch: ") + cudaGetErrorString(e));
    c->launches += 1;
    return 0;
}

int ptb_reset_accumulation(ptb_context* c) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    const size_t P = (size_t)c->width * c->height;
    PTB_CUDA(cudaMemsetAsync(c->st.weighted_radiance, 0, P * 20, c->stream));    // sum(w L) and sum(w): one allocation
    PTB_CUDA(cudaMemsetAsync(c->st.image, 0, P * 16, c->stream));
    return 0;
}

int ptb_get_buffer(ptb_context* c, const char* name, void** device_ptr, size_t* bytes) {
    if (!c || !name) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    auto it = c->named.find(name);
    if (it == c->named.end() || name[0] == '_') return pfail(RRB_ERROR_INVALID_ARGUMENT, std::string("unknown buffer '") + name + "'");
    if (device_ptr) *device_ptr = it->second.p;
    if (bytes) *bytes = it->second.bytes;
    return 0;
}

int ptb_read_buffer(ptb_context* c, const char* name, void* host_dst, size_t bytes) {
    void* p = nullptr;
    size_t n = 0;
    int rc = ptb_get_buffer(c, name, &p, &n);
    if (rc) return rc;
    if (!host_dst || bytes > n) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad destination or size");
    if ((rc = set_dev(c))) return rc;
    if ((rc = join_frames(c))) return rc;
    PTB_CUDA(cudaMemcpyAsync(host_dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int ptb_write_buffer(ptb_context* c, const char* name, const void* host_src, size_t bytes) {
    void* p = nullptr;
    size_t n = 0;
    int rc = ptb_get_buffer(c, name, &p, &n);
    if (rc) return rc;
    if (!host_src || bytes != n) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad source or size (a named buffer is written whole)");
    if ((rc = set_dev(c))) return rc;
    if ((rc = join_and_drop_frames(c))) return rc;
    PTB_CUDA(cudaMemcpyAsync(p, host_src, bytes, cudaMemcpyHostToDevice, c->stream));
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int ptb_read_wait(ptb_context* c) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    if (!c->copy_pending[0] && !c->copy_pending[1]) return 0;
    int rc = set_dev(c);
    if (rc) return rc;
    for (int k = 0; k < 2; ++k)
        if (c->copy_pending[k]) {
            PTB_CUDA(cudaEventSynchronize(c->copy_done[k]));
            c->copy_pending[k] = false;
        }
    return 0;
}

int ptb_read_buffer_async(ptb_context* c, const char* name, void* host_dst, size_t bytes) {
    void* p = nullptr;
    size_t n = 0;
    int rc = ptb_get_buffer(c, name, &p, &n);
    if (rc) return rc;
    if (!host_dst || bytes > n) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad destination or size");
    if ((rc = set_dev(c))) return rc;
    if ((rc = join_frames(c))) return rc;
    if (!c->copy_stream) {
        PTB_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            PTB_CUDA(cudaEventCreateWithFlags(&c->snap_ready[k], cudaEventDisableTiming));
            PTB_CUDA(cudaEventCreateWithFlags(&c->copy_done[k], cudaEventDisableTiming));
        }
    }
    const int k = c->snap_next;
    c->snap_next ^= 1;
    if (c->copy_pending[k]) {                        // the copy issued two calls ago: this snapshot is free again
        PTB_CUDA(cudaEventSynchronize(c->copy_done[k]));
        c->copy_pending[k] = false;
    }
    if (c->snapshot_bytes[k] < bytes) {
        if (c->snapshot[k]) cudaFree(c->snapshot[k]);
        c->snapshot[k] = nullptr;
        c->snapshot_bytes[k] = 0;
        cudaError_t e = cudaMalloc(&c->snapshot[k], bytes);
        if (e != cudaSuccess) return pfail(RRB_ERROR_OUT_OF_MEMORY, std::string("cudaMalloc(snapshot): ") + cudaGetErrorString(e));
        c->snapshot_bytes[k] = bytes;
    }
    PTB_CUDA(cudaMemcpyAsync(c->snapshot[k], p, bytes, cudaMemcpyDeviceToDevice, c->stream));
    PTB_CUDA(cudaEventRecord(c->snap_ready[k], c->stream));
    PTB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->snap_ready[k], 0));
    PTB_CUDA(cudaMemcpyAsync(host_dst, c->snapshot[k], bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    PTB_CUDA(cudaEventRecord(c->copy_done[k], c->copy_stream));
    c->copy_pending[k] = true;
    return 0;
}

int ptb_synchronize(ptb_context* c) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int ptb_selftest_fdiv(ptb_context* c, const float* a, const float* b, float* out_shortcut, float* out_plain, int n) {
    if (!c || !a || !b || !out_shortcut || !out_plain || n < 0) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad argument");
    int rc = set_dev(c);
    if (rc) return rc;
    if (n == 0) return 0;
    float* d = nullptr;
    const size_t bytes = (size_t)n * sizeof(float);
    PTB_CUDA(cudaMalloc((void**)&d, 4 * bytes));
    cudaError_t e = cudaMemcpyAsync(d, a, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + n, b, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = mcrt::launch_selftest_fdiv(d, d + n, d + 2 * (size_t)n, d + 3 * (size_t)n, n, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_shortcut, d + 2 * (size_t)n, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_plain, d + 3 * (size_t)n, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("fdiv self-test: ") + cudaGetErrorString(e));
    return 0;
}

int ptb_selftest_libm(ptb_context* c, int fn, const float* in, float* out, int n) {
    if (!c || !in || !out || n < 0 || fn < 0 || fn > 2) return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad argument");
    int rc = set_dev(c);
    if (rc) return rc;
    if (n == 0) return 0;
    float* d = nullptr;
    const size_t bytes = (size_t)n * sizeof(float);
    PTB_CUDA(cudaMalloc((void**)&d, 2 * bytes));
    cudaError_t e = cudaMemcpyAsync(d, in, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = mcrt::launch_selftest_libm(fn, d, d + n, n, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d + n, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (e != cudaSuccess) return pfail(RRB_ERROR_CUDA, std::string("libm self-test: ") + cudaGetErrorString(e));
    return 0;
}

int ptb_get_counters(ptb_context* c, ptb_counters* out) {
    if (!c || !out) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    unsigned long long h[8];
    PTB_CUDA(cudaMemcpyAsync(h, c->st.stats, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    PTB_CUDA(cudaStreamSynchronize(c->stream));
    out->closest_rays = (int64_t)h[0];
    out->shadow_rays = (int64_t)h[1];
    out->shaded_vertices = (int64_t)h[2];
    out->samples = (int64_t)h[3];
    out->kernel_launches = c->launches;
    return 0;
}

int ptb_reset_counters(ptb_context* c) {
    if (!c) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null context");
    int rc = set_dev(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    PTB_CUDA(cudaMemsetAsync(c->st.stats, 0, 64, c->stream));
    c->launches = 0;
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// multi-GPU merge: the one collective of the path -- NCCL over NVLink / NVSwitch
// ------------------------------------------------------------------------------------------------
static_assert(sizeof(ncclUniqueId) == PTB_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
#define PTB_NCCL(call)                                                                                  \
    do {                                                                                                \
        ncclResult_t r__ = (call);                                                                      \
        if (r__ != ncclSuccess) return pfail(RRB_ERROR_CUDA, std::string(#call) + ": " + ncclGetErrorString(r__)); \
    } while (0)

extern "C" {

int ptb_nccl_unique_id(void* id_out) {
    if (!id_out) return pfail(RRB_ERROR_INVALID_ARGUMENT, "null argument");
    ncclUniqueId id;
    PTB_NCCL(ncclGetUniqueId(&id));
    std::memcpy(id_out, &id, sizeof(id));
    return 0;
}

int ptb_nccl_comm_create(ptb_context* c, const void* id_in, int rank, int num_ranks, void** comm_out) {
    if (!c || !id_in || !comm_out || num_ranks < 1 || rank < 0 || rank >= num_ranks)
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "bad communicator arguments");
    int rc = set_dev(c);
    if (rc) return rc;
    ncclUniqueId id;
    std::memcpy(&id, id_in, sizeof(id));
    ncclComm_t comm = nullptr;
    PTB_NCCL(ncclCommInitRank(&comm, num_ranks, id, rank));
    *comm_out = comm;
    return 0;
}

int ptb_nccl_comm_destroy(void* comm) {
    if (!comm) return 0;
    PTB_NCCL(ncclCommDestroy((ncclComm_t)comm));
    return 0;
}

int ptb_merge_accumulation(ptb_context* c, void* nccl_comm, int root) {
    int rc = ready(c);
    if (rc) return rc;
    if ((rc = join_frames(c))) return rc;
    const size_t P = (size_t)c->width * c->height;
    float* acc = (float*)c->st.weighted_radiance;      // [4P | P], one allocation (ptb_resize)
    int rank = 0, nranks = 1;
    ncclComm_t comm = (ncclComm_t)nccl_comm;
    if (comm) {
        PTB_NCCL(ncclCommUserRank(comm, &rank));
        PTB_NCCL(ncclCommCount(comm, &nranks));
        if (root >= nranks) return pfail(RRB_ERROR_INVALID_ARGUMENT, "root is not a rank of the communicator");
    } else if (root > 0) {
        return pfail(RRB_ERROR_INVALID_ARGUMENT, "root > 0 needs a communicator");
    }
    if (root < 0) {
        // every rank gets the merged image; the rank-local sums stay as they are and keep accumulating
        const float* src = acc;
        if (comm && nranks > 1) {
            if (!c->merged) {
                void* p = nullptr;
                if ((rc = alloc_into(c->pixel_bufs, "merged_accumulation", P * 20, true, c->stream, &p))) return rc;
                c->merged = (float*)p;
            }
            PTB_NCCL(ncclAllReduce(acc, c->merged, 5 * P, ncclFloat32, ncclSum, comm, c->stream));
            src = c->merged;
        }
        PTB_CUDA(mcrt::launch_divide_accumulation((const float4*)src, src + 4 * P, c->st.image, (int)P, c->stream));
        c->launches += 1;
        return 0;
    }
    // root >= 0: the ranks HAND their sums to the root (in place there), and start again from zero; the root's
    // buffers then hold the sums over every rank and every frame so far
    if (comm && nranks > 1) PTB_NCCL(ncclReduce(acc, acc, 5 * P, ncclFloat32, ncclSum, root, comm, c->stream));
    if (rank == root) {
        PTB_CUDA(mcrt::launch_divide_accumulation(c->st.weighted_radiance, c->st.filter_weights, c->st.image, (int)P, c->stream));
        c->launches += 1;
    } else {
        PTB_CUDA(cudaMemsetAsync(acc, 0, P * 20, c->stream));
    }
    return 0;
}

}  // extern "C"
