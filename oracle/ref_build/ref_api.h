/* ref_api.h -- TEST INFRASTRUCTURE (oracle/_ref). C entry points of libmcrt_ref.so, the
 * reference's own sources compiled for the host (see Makefile in this directory). */
#ifndef MCRT_REF_API_H
#define MCRT_REF_API_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Host scene arrays in the reference's device layouts (kernel_data.h:278-294, RTScene.h:44-61).
 * float3 arrays have a 16-byte stride (RadeonRays::float3 is 4 floats). tangents / binormals /
 * colors are bound by the reference but never read on this path: not carried. */
typedef struct ref_scene_desc {
    const void* shapes;          /* RTShape[num_shapes], 160 B */
    const uint32_t* indices;     /* 3 per triangle, relative to RTShape::startVertex */
    const float* positions;      /* float4 stride */
    const float* uvs;            /* float2 */
    const float* normals;        /* float4 stride */
    const void* textures;        /* TextureDesc2D[num_textures], 16 B */
    const unsigned char* texdata;
    const uint32_t* sobol;       /* 1024 x 52, may be NULL for the random sampler */
    const void* lights;          /* RTLight[num_lights], 80 B */
    const void* materials;       /* RTMaterial[num_materials], 128 B */
    const void* camera;          /* RTPinholeCamera, 176 B */
    int32_t num_shapes, num_textures, num_lights, num_materials;
    int64_t num_indices, num_vertices, texdata_bytes;
} ref_scene_desc;

/* OpenMP team size of every kernel loop of this library: set explicitly / counted inside a parallel region */
void ref_set_threads(int n);
int ref_get_threads(void);

/* Radeon Rays host builder + kernels */
void* ref_rr_world_create(void);
void ref_rr_world_destroy(void* w);
int ref_rr_world_add_mesh(void* w, const float* verts, int vnum, int vstride, const int* idx, int istride,
                          int nfaces, int id, const float* m16, const float* minv16);
int ref_rr_world_add_instance(void* w, int base_shape, int id, const float* m16, const float* minv16);
long long ref_rr_world_build(void* w, float traversal_cost, int num_bins, int use_sah);
const void* ref_rr_world_nodes(void* w);
void ref_rr_intersect(const void* nodes, const void* rays, long long numrays, void* hits);
void ref_rr_occluded(const void* nodes, const void* rays, long long numrays, int* hits);

/* assets/kernels stages, one call = one NDRange over W x H */
int ref_pt_struct_sizes(int* out, int n);
void ref_pt_generate_rays(void* rays, void* ray_differentials, const void* camera);
void ref_pt_shade(const ref_scene_desc* scene, int width, int height, void* shadow_rays, void* rays,
                  const void* isects, int frame, int max_depth, int bounce, void* temp_radiance, void* throughput);
void ref_pt_shadow_pass(const void* shadow_rays, const int* occlusion, int width, int height, int bounce,
                        void* throughput, void* temp_radiance, void* radiance);

#ifdef __cplusplus
}
#endif
#endif
