// Compile-and-run check of include/radeon_rays_b200.hpp against the reference's own radeon_rays.h:
// the engine-side calling sequence of RTScene::attachMesh / commit and RTPathTracingPass's queries
// (scene/RTScene.cpp:659-670, renderPasses/RTPathTracingPass.cpp:81-84) on one triangle, the
// known-answer case of UnitTest/radeon_rays_apitest_cl.h:82-99.
// Without a CUDA device rrb_create fails; the adaptor must surface that as std::exception.
#include <radeon_rays_b200.hpp>

#include <cstdio>
#include <cstring>

using namespace RadeonRays;

int main() {
    IntersectionApi* api = nullptr;
    try {
        api = RadeonRaysB200::Create(0);
    } catch (std::exception const& e) {
        std::printf("NO_DEVICE %s\n", e.what());
        return 0;
    }
    try {
        const float vertices[] = {-1.f, -1.f, 0.f, 1.f, -1.f, 0.f, 0.f, 1.f, 0.f};
        const int indices[] = {0, 1, 2};
        const int numfaceverts[] = {3};
        Shape* mesh = api->CreateMesh(vertices, 3, 3 * sizeof(float), indices, 0, numfaceverts, 1);
        mesh->SetId(7);
        api->AttachShape(mesh);
        api->Commit();
        ray r(float3(0.f, 0.f, -10.f), float3(0.f, 0.f, 1.f), 10000.f);
        Buffer* rays = api->CreateBuffer(sizeof(ray), &r);
        Intersection init;
        Buffer* hits = api->CreateBuffer(sizeof(Intersection), &init);
        Event* ev = nullptr;
        api->QueryIntersection(rays, 1, hits, nullptr, &ev);
        ev->Wait();
        api->DeleteEvent(ev);
        Intersection* out = nullptr;
        api->MapBuffer(hits, kMapRead, 0, sizeof(Intersection), (void**)&out, &ev);
        ev->Wait();
        api->DeleteEvent(ev);
        const bool ok = out->shapeid == 7 && out->primid == 0 && out->uvwt.w > 9.99f && out->uvwt.w < 10.01f;
        std::printf("HIT shape %d prim %d t %f\n", out->shapeid, out->primid, out->uvwt.w);
        api->UnmapBuffer(hits, out, nullptr);
        // empty world must throw like radeon_rays_impl.cpp:113
        api->DetachAll();
        bool threw = false;
        try {
            api->Commit();
        } catch (RadeonRays::Exception const& e) {
            threw = std::strstr(e.what(), "empty") != nullptr;
        }
        api->DeleteBuffer(rays);
        api->DeleteBuffer(hits);
        api->DeleteShape(mesh);
        RadeonRaysB200::Delete(api);
        if (!ok || !threw) return 1;
        std::printf("OK\n");
        return 0;
    } catch (std::exception const& e) {
        std::printf("FAILED %s\n", e.what());
        return 2;
    }
}
