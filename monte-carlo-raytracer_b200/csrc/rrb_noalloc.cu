// rrb_noalloc.cu -- rrb_query_intersection_noalloc: closest-hit query that allocates nothing.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The query kernels of rrb_kernels.cu use scratch memory that rrb_create cudaMalloc's (the work counter the
// persistent warps pull rays from) and read the ray buffer sparsely when many rays are inactive.  This variant keeps
// ALL of its scratch in one file-scope device array of fixed size:
//
//   bytes   0..3     work counter of the traversal kernel        (must start at zero: cleared on every call)
//   bytes   4..7     number of active rays found by the compaction pass (likewise)
//   bytes 256..      uint32 index of every active ray, compacted (RRB_NOALLOC_MAX_RAYS entries)
//
// Pass 1 (k_compact_active) writes the indices of the rays whose `extra.y` is non-zero into the list; pass 2
// (k_query_compacted) is the persistent traversal of rrb_kernels.cu:query_kernel pulling from that list, so no lane
// ever fetches an inactive ray.  Per ray the traversal is the same code (traverse.cuh), so hits are bit-identical to
// rrb_query_intersection's; inactive rays and misses leave the same bytes untouched.
//
// The array is shared by every call on a device and keeps the last call's contents: calls whose kernels can overlap
// on the GPU (two streams, two API objects on one device) race on it.  Every GPU that loads the library holds the
// 4 MiB + 256 B until the process exits.
#include <cuda_runtime.h>

#include <cstdlib>
#include <mutex>
#include <string>

#include "../../include/rrb.h"
#include "traverse.cuh"

namespace mcrt {
int rrb_fail(int code, const std::string& msg);   // rrb_api.cu: sets rrb_last_error()
}

namespace {

constexpr size_t kHeaderBytes = 256;
constexpr size_t kScratchBytes = kHeaderBytes + (size_t)RRB_NOALLOC_MAX_RAYS * sizeof(unsigned int);

// In host code this name is a host-side stand-in, not a device address: the host obtains the address with
// cudaGetSymbolAddress (scratch_of_device below) and hands it to the kernels as an argument.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

struct ApiRay {      // RadeonRays::ray, include/math/ray.h:103-107 (as in rrb_kernels.cu)
    float4 o;
    float4 d;
    int4 extra;      // x mask, y active
};

__global__ void __launch_bounds__(256)
k_compact_active(const ApiRay* __restrict__ rays, int numrays, unsigned int* __restrict__ header,
                 unsigned int* __restrict__ list) {
    const unsigned lane = threadIdx.x & 31u;
    const int stride = (int)(gridDim.x * blockDim.x);
    // warp-uniform trip count: every lane of a warp takes part in every ballot
    for (int base = (int)(blockIdx.x * blockDim.x + threadIdx.x) - (int)lane; base < numrays; base += stride) {
        const int i = base + (int)lane;
        const bool active = i < numrays && __ldg(&rays[i].extra).y != 0;
        const unsigned m = __ballot_sync(0xffffffffu, active);
        if (m == 0) continue;
        const int leader = __ffs(m) - 1;
        unsigned at = 0;
        if ((int)lane == leader) at = atomicAdd(header + 1, (unsigned)__popc(m));
        at = __shfl_sync(0xffffffffu, at, leader);
        if (active) list[at + (unsigned)__popc(m & ((1u << lane) - 1u))] = (unsigned)i;
    }
}

template <bool TWO_LEVEL>
__global__ void __launch_bounds__(128, TWO_LEVEL ? 6 : 8)
k_query_compacted(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const ApiRay* __restrict__ rays,
                  void* __restrict__ out, unsigned int* __restrict__ header, const unsigned int* __restrict__ list,
                  int num_nodes, int stage_top, const float4* __restrict__ instances) {
    using namespace mcrt;
    const unsigned num_active = header[1];          // written by the compaction kernel launched before this one
    unsigned int* work_counter = header;
    const unsigned lane = threadIdx.x & 31u;
    __shared__ TopNodes s_top;
    Traverser<false, false, false, TWO_LEVEL> trav{nodes, tris};
    trav.instances = instances;
    trav.top = s_top.nodes;
    trav.top_count = stage_top_nodes(s_top, nodes, num_nodes, stage_top);
    uint2 stack[MCRT_STACK_SIZE];
    TravState st;
    Ray ray;
    int ray_index = -1;
    bool exhausted = false;
    int cnt[2] = {0, 0};

    while (true) {
        const unsigned idle = __ballot_sync(0xffffffffu, ray_index < 0 && !exhausted);
        if (idle) {
            const int n_idle = __popc(idle);
            const int my_rank = __popc(idle & ((1u << lane) - 1u));
            unsigned base = 0;
            const int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(work_counter, (unsigned)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (ray_index < 0 && !exhausted) {
                const unsigned slot = base + (unsigned)my_rank;
                if (slot >= num_active) {
                    exhausted = true;
                } else {
                    const unsigned idx = list[slot];
                    const ApiRay ar = rays[idx];
                    ray.ox = ar.o.x; ray.oy = ar.o.y; ray.oz = ar.o.z; ray.tmax = ar.o.w;
                    ray.dx = ar.d.x; ray.dy = ar.d.y; ray.dz = ar.d.z; ray.mask = ar.extra.x;
                    trav.init(st, ray);
                    ray_index = (int)idx;
                }
            }
        }
        if (__all_sync(0xffffffffu, ray_index < 0 && exhausted)) break;

        if (ray_index >= 0) {
            const bool done = trav.run(st, ray, stack, cnt);
            if (done) {
                // the hit record of rrb_kernels.cu:query_kernel, byte for byte
                int4* h = (int4*)out + 2 * (size_t)ray_index;
                if (st.best_tri >= 0) {
                    const float4 v0 = __ldg(tris + 3 * (size_t)st.best_tri + 0);
                    const float4 v1 = __ldg(tris + 3 * (size_t)st.best_tri + 1);
                    const float4 v2 = __ldg(tris + 3 * (size_t)st.best_tri + 2);
                    const Ray hr = TWO_LEVEL ? hit_space_ray(instances, ray, st.best_inst) : ray;
                    const float px = add_(hr.ox, mul_(st.closest, hr.dx));
                    const float py = add_(hr.oy, mul_(st.closest, hr.dy));
                    const float pz = add_(hr.oz, mul_(st.closest, hr.dz));
                    const float2 uv = barycentrics_ref(px, py, pz, v0, v1, v2);
                    ((int2*)h)[0] = make_int2(TWO_LEVEL ? st.best_sid : __float_as_int(v0.w), __float_as_int(v1.w));
                    ((float4*)h)[1] = make_float4(uv.x, uv.y, 0.0f, st.closest);
                } else {
                    ((int2*)h)[0] = make_int2(-1, -1);   // uvwt untouched on a miss
                }
                ray_index = -1;
            }
        }
    }
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
unsigned char* g_scratch_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch on each device
int g_sm_count[kMaxDevices] = {};

// Set-up, once per device, with that device current: the device address of the static array.
int scratch_of_device(int device, unsigned char** out, int* sms) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return mcrt::rrb_fail(RRB_ERROR_CUDA, std::string("cudaGetSymbolAddress(noalloc scratch): ") + cudaGetErrorString(e));
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || n <= 0)
            return mcrt::rrb_fail(RRB_ERROR_CUDA, std::string("cudaDeviceGetAttribute(SM count): ") + cudaGetErrorString(e));
        g_sm_count[device] = n;
        g_scratch_ptr[device] = (unsigned char*)p;
    }
    *out = g_scratch_ptr[device];
    *sms = g_sm_count[device];
    return RRB_OK;
}

}  // namespace

extern "C" int rrb_query_intersection_noalloc(rrb_api* api, const void* rays_device, int numrays, void* hits_device) {
    if (!api) return mcrt::rrb_fail(RRB_ERROR_INVALID_ARGUMENT, "null api");
    if (numrays < 0) return mcrt::rrb_fail(RRB_ERROR_INVALID_ARGUMENT, "numrays < 0");
    if (numrays > RRB_NOALLOC_MAX_RAYS)
        return mcrt::rrb_fail(RRB_ERROR_INVALID_ARGUMENT,
                              "rrb_query_intersection_noalloc: numrays " + std::to_string(numrays) +
                                  " exceeds RRB_NOALLOC_MAX_RAYS (" + std::to_string(RRB_NOALLOC_MAX_RAYS) +
                                  "): the scratch array is static; split the batch or use rrb_query_intersection");
    const void *nodes = nullptr, *tris = nullptr, *instances = nullptr;
    int64_t num_nodes = 0;
    int rc = rrb_get_bvh(api, &nodes, &tris, &num_nodes, nullptr);      // fails when nothing is committed
    if (rc) return rc;
    if (numrays == 0) return RRB_OK;
    if (!rays_device || !hits_device) return mcrt::rrb_fail(RRB_ERROR_INVALID_ARGUMENT, "null ray or hit pointer");
    int device = 0, n_inst = 0;
    void* stream_v = nullptr;
    if ((rc = rrb_get_stream(api, &device, &stream_v))) return rc;
    if ((rc = rrb_get_instances(api, &instances, &n_inst))) return rc;
    if (device < 0 || device >= kMaxDevices) return mcrt::rrb_fail(RRB_ERROR_INVALID_ARGUMENT, "device index out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return mcrt::rrb_fail(RRB_ERROR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    unsigned char* scratch = nullptr;
    int sms = 0;
    if ((rc = scratch_of_device(device, &scratch, &sms))) return rc;
    cudaStream_t stream = (cudaStream_t)stream_v;
    unsigned int* header = (unsigned int*)scratch;
    unsigned int* list = (unsigned int*)(scratch + kHeaderBytes);

    // the two counters must start at zero; the list needs no clearing (only its first `active` entries are read)
    e = cudaMemsetAsync(header, 0, kHeaderBytes, stream);
    if (e != cudaSuccess) return mcrt::rrb_fail(RRB_ERROR_CUDA, std::string("cudaMemsetAsync(noalloc scratch): ") + cudaGetErrorString(e));
    const ApiRay* rays = (const ApiRay*)rays_device;
    {
        const int want = (numrays + 255) / 256;
        const int blocks = want < sms * 8 ? want : sms * 8;
        k_compact_active<<<blocks, 256, 0, stream>>>(rays, numrays, header, list);
    }
    static const int stage_top = getenv("MCRT_STAGE_TOP") ? atoi(getenv("MCRT_STAGE_TOP")) : 1;
    {
        const int want = ((numrays + 31) / 32 + 3) / 4;
        const int blocks = want < sms * 8 ? want : sms * 8;
        const int nn = (int)(num_nodes < 0x7fffffff ? num_nodes : 0x7fffffff);
        if (instances)
            k_query_compacted<true><<<blocks, 128, 0, stream>>>((const uint4*)nodes, (const float4*)tris, rays, hits_device,
                                                                 header, list, nn, stage_top, (const float4*)instances);
        else
            k_query_compacted<false><<<blocks, 128, 0, stream>>>((const uint4*)nodes, (const float4*)tris, rays, hits_device,
                                                                  header, list, nn, stage_top, nullptr);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return mcrt::rrb_fail(RRB_ERROR_CUDA, std::string("noalloc query launch: ") + cudaGetErrorString(e));
    return RRB_OK;
}
