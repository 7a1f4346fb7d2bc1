// ref_rr_bvh.cpp -- TEST INFRASTRUCTURE (oracle/_ref). Not part of the product path.
//
// Drives Radeon Rays' own host-side builder (RadeonRays/src/accelerator/bvh2.{h,cpp},
// primitive/mesh.cpp), compiled from /root/reference, exactly as IntersectorLDS::Process does
// (intersector_lds.cpp:163-264): Bvh2(traversal_cost, num_bins, use_sah).Build(shapes) and a
// copy of the 64-byte nodes.  Bvh2::m_nodes is private with `friend class IntersectorLDS`
// (bvh2.h:172-182); this TU is compiled with -fno-access-control to read it.
#include <cstring>
#include <memory>
#include <vector>

#include "accelerator/bvh2.h"
#include "primitive/mesh.h"
#include "primitive/instance.h"

using namespace RadeonRays;

struct RefWorld {
    std::vector<std::unique_ptr<ShapeImpl>> shapes;
    std::vector<const Shape*> attached;
    std::vector<Bvh2::Node> nodes;
};

static matrix to_matrix(const float* m16) {
    matrix m;
    std::memcpy(&m.m[0][0], m16, 16 * sizeof(float));
    return m;
}

extern "C" {

void* ref_rr_world_create(void) { return new RefWorld(); }
void ref_rr_world_destroy(void* w) { delete (RefWorld*)w; }

// IntersectionApiImpl::CreateMesh (radeon_rays_impl.cpp:60-76) + SetId + SetTransform + AttachShape.
int ref_rr_world_add_mesh(void* w, const float* verts, int vnum, int vstride, const int* idx, int istride,
                          int nfaces, int id, const float* m16, const float* minv16) {
    RefWorld* world = (RefWorld*)w;
    auto mesh = std::make_unique<Mesh>(verts, vnum, vstride, idx, istride, nullptr, nfaces);
    mesh->SetId(id);
    if (m16) mesh->SetTransform(to_matrix(m16), to_matrix(minv16 ? minv16 : m16));
    world->attached.push_back(mesh.get());
    world->shapes.push_back(std::move(mesh));
    return (int)world->shapes.size() - 1;
}

// CreateInstance of a previously added shape (radeon_rays_impl.cpp:79-90), flattened by Bvh2::Build.
int ref_rr_world_add_instance(void* w, int base_shape, int id, const float* m16, const float* minv16) {
    RefWorld* world = (RefWorld*)w;
    auto inst = std::make_unique<Instance>(world->shapes[(size_t)base_shape].get());
    inst->SetId(id);
    if (m16) inst->SetTransform(to_matrix(m16), to_matrix(minv16 ? minv16 : m16));
    world->attached.push_back(inst.get());
    world->shapes.push_back(std::move(inst));
    return (int)world->shapes.size() - 1;
}

// Returns the node count (2T-1); nodes are then read with ref_rr_world_nodes.
long long ref_rr_world_build(void* w, float traversal_cost, int num_bins, int use_sah) {
    RefWorld* world = (RefWorld*)w;
    Bvh2 bvh(traversal_cost, num_bins, use_sah != 0);
    bvh.Build(world->attached.begin(), world->attached.end());
    world->nodes.assign(bvh.m_nodes, bvh.m_nodes + bvh.m_nodecount);
    return (long long)bvh.m_nodecount;
}

const void* ref_rr_world_nodes(void* w) { return ((RefWorld*)w)->nodes.data(); }
}
