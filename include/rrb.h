/* rrb.h -- C ABI of the B200 intersection library (librrb_b200 inside libmcrt_b200.so).
 *
 * Drop-in boundary, surface 1: every entry point replaces one method of
 * RadeonRays::IntersectionApi (third_party/RadeonRays/RadeonRays/include/radeon_rays.h:170-305)
 * or of its factory (include/radeon_rays_cl.h:40-41).  The reference-side binding (a header-only
 * C++ adaptor deriving from RadeonRays::IntersectionApi) is in INTEGRATION.md and
 * monte-carlo-raytracer_b200/host/radeon_rays_b200.hpp.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative
 * rrb_status otherwise; the message of the last failure on the calling thread is rrb_last_error().
 * Objects are opaque and created / destroyed only through this API ("to simplify DLL boundary
 * crossing", radeon_rays.h:228-229).  One rrb_api per GPU; an rrb_api is not re-entrant (like the
 * reference: shared counter and lazily grown stack, intersector_lds.cpp:270-277).
 *
 * Query layouts are the reference's: ray 48 B (include/math/ray.h:103-107), Intersection 32 B
 * (radeon_rays.h:140-154), occlusion int32 (1 hit / -1 miss, common.cl:31-32).  Inactive rays
 * (extra.y == 0) and indices >= numrays are NOT written; a miss writes shapeid = primid = -1 and
 * leaves uvwt untouched (intersect_bvh2_lds.cl:84-88,218-223).
 */
#ifndef RRB_H
#define RRB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rrb_api rrb_api;
typedef struct rrb_shape rrb_shape;
typedef struct rrb_buffer rrb_buffer;
typedef struct rrb_event rrb_event;

enum rrb_status {
    RRB_OK = 0,
    RRB_ERROR_INVALID_ARGUMENT = -1,
    RRB_ERROR_EMPTY_SCENE = -2,   /* "Scene is empty." radeon_rays_impl.cpp:113 */
    RRB_ERROR_CUDA = -3,
    RRB_ERROR_OUT_OF_MEMORY = -4,
    RRB_ERROR_NOT_COMMITTED = -5,
    RRB_ERROR_UNSUPPORTED = -6    /* quads: Bvh2 asserts triangles only, bvh2.h:221-222 */
};

enum rrb_map_type { RRB_MAP_READ = 0x1, RRB_MAP_WRITE = 0x2 }; /* radeon_rays.h:156-160 */

/* Thread-local text of the last error; never NULL. */
const char* rrb_last_error(void);

/* ---- device management: IntersectionApi::GetDeviceCount / GetDeviceInfo (radeon_rays.h:189-191) */
int rrb_device_count(void);
int rrb_device_info(int devidx, char* name, size_t name_len, char* vendor, size_t vendor_len, int* type);

/* ---- lifetime: IntersectionApi::Create / CreateFromOpenClContext / Delete (radeon_rays.h:200-203).
 * `cuda_stream` is a cudaStream_t (NULL = a stream owned by the API); all queries are enqueued on it. */
int rrb_create(int cuda_device, void* cuda_stream, rrb_api** out);
int rrb_destroy(rrb_api* api);

/* ---- geometry (radeon_rays.h:214-247).  Strides in bytes, 0 = tightly packed; blocking, copies. */
int rrb_create_mesh(rrb_api* api, const float* vertices, int vnum, int vstride, const int* indices, int istride,
                    const int* numfacevertices, int numfaces, rrb_shape** out);
int rrb_create_instance(rrb_api* api, const rrb_shape* base, rrb_shape** out);
int rrb_delete_shape(rrb_api* api, rrb_shape* shape);
int rrb_attach(rrb_api* api, const rrb_shape* shape);       /* idempotent, world.cpp:28-35 */
int rrb_detach(rrb_api* api, const rrb_shape* shape);
int rrb_detach_all(rrb_api* api);
int rrb_commit(rrb_api* api);                                /* blocking (re)build when anything changed */
int rrb_reset_id_counter(rrb_api* api);
int rrb_is_world_empty(rrb_api* api, int* out_empty);

/* ---- Shape interface (radeon_rays.h:91-110). Matrices are row-major 4x4 (math/matrix.h:71-81). */
int rrb_shape_set_transform(rrb_shape* shape, const float m[16], const float minv[16]);
int rrb_shape_get_transform(const rrb_shape* shape, float m[16], float minv[16]);
int rrb_shape_set_id(rrb_shape* shape, int id);
int rrb_shape_get_id(const rrb_shape* shape, int* out_id);
int rrb_shape_set_linear_velocity(rrb_shape* shape, const float v[3]);   /* stored, no reachable consumer */
int rrb_shape_set_angular_velocity(rrb_shape* shape, const float q[4]); /* stored, no reachable consumer */

/* ---- options: SetOption(char const*, char const*) / (char const*, float) (radeon_rays.h:279-296).
 * Unknown keys are stored and never an error (util/options.cpp).
 * Extension key (f32)  "bvh.refit" != 0 : rrb_commit keeps the tree topology and refits it on the device when the
 *   attached shapes are those of the last full build and only transforms / ids changed (the reference always
 *   rebuilds on the CPU, intersector_lds.cpp:166).  Device pointers returned by rrb_get_bvh stay valid across a
 *   refit.  Set it before the first rrb_commit; rrb_stats.refits counts the commits served this way. */
int rrb_set_option_str(rrb_api* api, const char* name, const char* value);
int rrb_set_option_f32(rrb_api* api, const char* name, float value);

/* ---- buffers (radeon_rays.h:252-259; CreateFromOpenClBuffer radeon_rays_impl.cpp:175-191) */
int rrb_create_buffer(rrb_api* api, size_t size, const void* initdata, rrb_buffer** out);
int rrb_wrap_device_ptr(rrb_api* api, void* device_ptr, size_t size, rrb_buffer** out); /* non-owning, O(1) */
int rrb_delete_buffer(rrb_api* api, rrb_buffer* buffer);
int rrb_buffer_device_ptr(const rrb_buffer* buffer, void** out_ptr, size_t* out_size);
int rrb_map(rrb_api* api, rrb_buffer* buffer, int map_type, size_t offset, size_t size, void** data, rrb_event** ev);
int rrb_unmap(rrb_api* api, rrb_buffer* buffer, void* ptr, rrb_event** ev);

/* ---- events (radeon_rays.h:120-129,264); pooled */
int rrb_event_complete(rrb_event* ev, int* out_complete);
int rrb_event_wait(rrb_event* ev);
int rrb_event_delete(rrb_api* api, rrb_event* ev);

/* ---- ray casting (radeon_rays.h:267-277).  Asynchronous on the API's stream; wait / done may be NULL. */
int rrb_query_intersection(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hits,
                           const rrb_event* wait, rrb_event** done);
int rrb_query_occlusion(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hitresults,
                        const rrb_event* wait, rrb_event** done);
int rrb_query_intersection_indirect(rrb_api* api, const rrb_buffer* rays, const rrb_buffer* numrays, int maxrays,
                                    rrb_buffer* hits, const rrb_event* wait, rrb_event** done);
int rrb_query_occlusion_indirect(rrb_api* api, const rrb_buffer* rays, const rrb_buffer* numrays, int maxrays,
                                 rrb_buffer* hitresults, const rrb_event* wait, rrb_event** done);

/* ---- host-buffer convenience used by the end-to-end timing leg: H2D copy of `numrays` rays from host
 * memory, query, D2H copy of the results, blocking.  (The reference's "fast path" comment,
 * radeon_rays.h:163-166.) */
int rrb_query_intersection_host(rrb_api* api, const void* host_rays, int numrays, void* host_hits);
int rrb_query_occlusion_host(rrb_api* api, const void* host_rays, int numrays, int* host_results);

/* ---- closest-hit query that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same results as rrb_query_intersection, byte for byte (ray 48 B in, Intersection 32 B
 * out, inactive rays and the uvwt of misses untouched), on raw DEVICE pointers the caller owns; asynchronous on the
 * API's stream (rrb_get_stream), no event.  Its scratch -- the work counter of the persistent traversal kernel and
 * a compacted list of the active rays' indices -- lives in one static device array of
 * 256 + 4 * RRB_NOALLOC_MAX_RAYS bytes (4 MiB) that every GPU which loads the library holds until exit; the call
 * makes no cudaMalloc / cudaFree and creates no stream or event.  numrays > RRB_NOALLOC_MAX_RAYS is rejected with
 * RRB_ERROR_INVALID_ARGUMENT (split the batch).  All calls on a device share the array: calls whose kernels can
 * overlap on the GPU (two streams or two API objects on one device) race; calls on one stream do not.
 * (csrc/rrb_noalloc.cu) */
#define RRB_NOALLOC_MAX_RAYS (1 << 20)
int rrb_query_intersection_noalloc(rrb_api* api, const void* rays_device, int numrays, void* hits_device);

/* ---- introspection for measurement (no reference counterpart) */
typedef struct rrb_stats {
    int64_t num_triangles;
    int64_t num_nodes;          /* 80-byte wide nodes */
    int64_t node_bytes;
    int64_t triangle_bytes;
    double build_seconds;       /* wall time of the last rrb_commit */
    int64_t launches;           /* traversal kernels launched since creation */
    int64_t refits;             /* commits served by the device refit (option "bvh.refit") since creation */
    int64_t last_commit_refit;  /* 1 when the last rrb_commit was a refit, 0 when it was a full build */
    int64_t last_build_on_device; /* 1 when the last full build ran on the device (acc.type = "hlbvh" or
                                     bvh.builder = "lbvh": Morton codes, radix sort, radix tree, wide collapse as CUDA
                                     kernels -- Radeon Rays' own GPU builder sits behind the same option,
                                     intersector_hlbvh.cpp); 0 for the multi-threaded host SAH builder */
    int64_t num_levels;         /* breadth-first levels of the tree when known (device build or "bvh.refit"), else 0;
                                   two-level world: levels of the top-level tree */
    int64_t two_level;          /* 1 when the committed world is a top-level tree over shapes with one bottom-level
                                   tree per distinct mesh (instances attached, or option "bvh.force2level"), else 0 */
    int64_t num_instances;      /* shapes in the top level */
    int64_t two_level_rebuilds; /* commits that (re)laid out the bottom-level trees */
    int64_t top_level_updates;  /* commits that only rebuilt the top level and the shape table (transform / id changes) */
} rrb_stats;
int rrb_get_stats(rrb_api* api, rrb_stats* out);
/* The shape table of a two-level world (NULL / 0 for a flat one): 64-byte records in the top-level tree's leaf order,
 * rows 0..2 of the world -> object matrix, bottom-level root node, shape id (csrc/bvh8.h: InstanceRecord).  Like
 * rrb_get_bvh, for the path-tracer half of the library, which traces against the same structure. */
int rrb_get_instances(rrb_api* api, const void** instances, int* num_instances);
/* Instrumented query: per-ray visited-node and tested-triangle counts into int32[2*numrays]. */
int rrb_query_intersection_counted(rrb_api* api, const rrb_buffer* rays, int numrays, rrb_buffer* hits,
                                   rrb_buffer* counts);
/* Device pointers of the committed acceleration structure (used by the path-tracer library). */
int rrb_get_bvh(rrb_api* api, const void** nodes, const void** triangles, int64_t* num_nodes, int64_t* num_tris);
/* The CUDA device index and cudaStream_t every query of this API is enqueued on. */
int rrb_get_stream(rrb_api* api, int* cuda_device, void** cuda_stream);


/* ---- host-only builder access (no CUDA call; used by the CPU test-suite to validate the tree):
 * builds the compressed 8-wide BVH of one tightly packed triangle soup with per-face shape ids.
 * use_sah: bit 0 = binned SAH (else median splits); bit 1 = diagnostic, switches off the depth guard that rebuilds
 * a tree deeper than the traversal stack allows (19 levels) -- the result must then not be traversed. */
typedef struct rrb_host_bvh rrb_host_bvh;
int rrb_host_bvh_build(const float* vertices, int vnum, const int* indices, int numfaces, const int* face_shape_ids,
                       int use_sah, rrb_host_bvh** out);
int rrb_host_bvh_get(const rrb_host_bvh* bvh, const uint32_t** nodes, int64_t* num_nodes, const float** triangles,
                     int64_t* num_tris);
/* Refit metadata of the build: mesh index (run of equal shape ids) of every emitted triangle, and the node index
 * each breadth-first tree level starts at (num_levels + 1 entries). */
int rrb_host_bvh_get_refit_info(const rrb_host_bvh* bvh, const int32_t** tri_mesh, const int32_t** level_offsets,
                                int* num_levels);
int rrb_host_bvh_free(rrb_host_bvh* bvh);

#ifdef __cplusplus
}
#endif
#endif /* RRB_H */
