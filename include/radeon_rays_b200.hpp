// radeon_rays_b200.hpp -- header-only C++ adaptor: RadeonRays::IntersectionApi implemented on the
// C ABI of include/rrb.h (libmcrt_b200.so).  This is the file a maintainer of the reference adds to
// the engine so that `g_isectApi` (source/application/PathTracer/raytracing/system/rt_globals.h:6)
// points at the CUDA implementation; every caller of the interface
// (third_party/RadeonRays/RadeonRays/include/radeon_rays.h:170-305) keeps compiling unchanged.
//
// Needs the reference's own <radeon_rays.h> on the include path (it is NOT shipped here).
//
//   RadeonRays::IntersectionApi* api = RadeonRaysB200::Create(/*cuda device*/0, /*cudaStream_t*/nullptr);
//   RadeonRays::Buffer* rays = RadeonRaysB200::CreateFromDevicePtr(api, d_rays, bytes);   // was CreateFromOpenClBuffer
//   ...
//   RadeonRaysB200::Delete(api);
//
// Error convention: the C ABI returns a status and a thread-local message; the adaptor re-throws it as
// RadeonRaysB200::Error, which derives from BOTH RadeonRays::Exception and std::exception so that the
// engine's existing `catch (std::exception&)` clauses (RTPathTracingPass.cpp:89-104) keep firing.
#pragma once
#include <radeon_rays.h>

#include <exception>
#include <string>
#include <unordered_map>

#include "rrb.h"

namespace RadeonRaysB200 {

class Error : public RadeonRays::Exception, public std::exception {
public:
    Error(int status, std::string msg) : status_(status), msg_(std::move(msg)) {}
    char const* what() const noexcept override { return msg_.c_str(); }
    int status() const noexcept { return status_; }

private:
    int status_;
    std::string msg_;
};

inline void Check(int rc) {
    if (rc != RRB_OK) throw Error(rc, rrb_last_error());
}

class ShapeB200 final : public RadeonRays::Shape {
public:
    explicit ShapeB200(rrb_shape* h) : h_(h) {}
    rrb_shape* handle() const { return h_; }

    // matrix is row-major 4x4 (math/matrix.h:32-60), which is what rrb_shape_set_transform takes
    void SetTransform(RadeonRays::matrix const& m, RadeonRays::matrix const& minv) override {
        Check(rrb_shape_set_transform(h_, &m.m[0][0], &minv.m[0][0]));
    }
    void GetTransform(RadeonRays::matrix& m, RadeonRays::matrix& minv) const override {
        Check(rrb_shape_get_transform(h_, &m.m[0][0], &minv.m[0][0]));
    }
    void SetLinearVelocity(RadeonRays::float3 const& v) override {
        lin_ = v;
        const float a[3] = {v.x, v.y, v.z};
        Check(rrb_shape_set_linear_velocity(h_, a));
    }
    RadeonRays::float3 GetLinearVelocity() const override { return lin_; }
    void SetAngularVelocity(RadeonRays::quaternion const& q) override {
        ang_ = q;
        const float a[4] = {q.x, q.y, q.z, q.w};
        Check(rrb_shape_set_angular_velocity(h_, a));
    }
    RadeonRays::quaternion GetAngularVelocity() const override { return ang_; }
    void SetId(RadeonRays::Id id) override { Check(rrb_shape_set_id(h_, id)); }
    RadeonRays::Id GetId() const override {
        int id = RadeonRays::kNullId;
        Check(rrb_shape_get_id(h_, &id));
        return id;
    }

private:
    rrb_shape* h_;
    RadeonRays::float3 lin_;
    RadeonRays::quaternion ang_;
};

class BufferB200 final : public RadeonRays::Buffer {
public:
    explicit BufferB200(rrb_buffer* h) : h_(h) {}
    rrb_buffer* handle() const { return h_; }

private:
    rrb_buffer* h_;
};

class EventB200 final : public RadeonRays::Event {
public:
    explicit EventB200(rrb_event* h) : h_(h) {}
    rrb_event* handle() const { return h_; }
    bool Complete() const override {
        int done = 0;
        Check(rrb_event_complete(h_, &done));
        return done != 0;
    }
    void Wait() override { Check(rrb_event_wait(h_)); }

private:
    rrb_event* h_;
};

class IntersectionApiB200 final : public RadeonRays::IntersectionApi {
public:
    IntersectionApiB200(int cuda_device, void* cuda_stream) { Check(rrb_create(cuda_device, cuda_stream, &h_)); }
    ~IntersectionApiB200() override {
        for (auto& kv : wrappers_) delete kv.second.buf;
        rrb_destroy(h_);
    }
    rrb_api* handle() const { return h_; }

    // ---- geometry (radeon_rays.h:214-247)
    RadeonRays::Shape* CreateMesh(float const* vertices, int vnum, int vstride, int const* indices, int istride,
                                  int const* numfacevertices, int numfaces) const override {
        rrb_shape* s = nullptr;
        Check(rrb_create_mesh(h_, vertices, vnum, vstride, indices, istride, numfacevertices, numfaces, &s));
        return new ShapeB200(s);
    }
    RadeonRays::Shape* CreateInstance(RadeonRays::Shape const* shape) const override {
        rrb_shape* s = nullptr;
        Check(rrb_create_instance(h_, Cast(shape)->handle(), &s));
        return new ShapeB200(s);
    }
    void DeleteShape(RadeonRays::Shape const* shape) override {
        if (!shape) return;
        Check(rrb_delete_shape(h_, Cast(shape)->handle()));
        delete Cast(shape);
    }
    void AttachShape(RadeonRays::Shape const* shape) override { Check(rrb_attach(h_, Cast(shape)->handle())); }
    void DetachShape(RadeonRays::Shape const* shape) override { Check(rrb_detach(h_, Cast(shape)->handle())); }
    void DetachAll() override { Check(rrb_detach_all(h_)); }
    void Commit() override { Check(rrb_commit(h_)); }          // throws "Scene is empty." like radeon_rays_impl.cpp:113
    void ResetIdCounter() override { Check(rrb_reset_id_counter(h_)); }
    bool IsWorldEmpty() override {
        int e = 1;
        Check(rrb_is_world_empty(h_, &e));
        return e != 0;
    }

    // ---- memory (radeon_rays.h:249-258)
    RadeonRays::Buffer* CreateBuffer(size_t size, void* initdata) const override {
        rrb_buffer* b = nullptr;
        Check(rrb_create_buffer(h_, size, initdata, &b));
        return new BufferB200(b);
    }
    void DeleteBuffer(RadeonRays::Buffer* buffer) const override {
        if (!buffer) return;
        Check(rrb_delete_buffer(h_, Cast(buffer)->handle()));
        delete Cast(buffer);
    }
    void MapBuffer(RadeonRays::Buffer* buffer, RadeonRays::MapType type, size_t offset, size_t size, void** data,
                   RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_map(h_, Cast(buffer)->handle(), (int)type, offset, size, data, event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }
    void UnmapBuffer(RadeonRays::Buffer* buffer, void* ptr, RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_unmap(h_, Cast(buffer)->handle(), ptr, event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }
    void DeleteEvent(RadeonRays::Event* event) const override {
        if (!event) return;
        Check(rrb_event_delete(h_, static_cast<EventB200*>(event)->handle()));
        delete static_cast<EventB200*>(event);
    }

    // ---- queries (radeon_rays.h:260-277)
    void QueryIntersection(RadeonRays::Buffer const* rays, int numrays, RadeonRays::Buffer* hitinfos,
                           RadeonRays::Event const* waitevent, RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_query_intersection(h_, Cast(rays)->handle(), numrays, Cast(hitinfos)->handle(), Ev(waitevent),
                                     event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }
    void QueryOcclusion(RadeonRays::Buffer const* rays, int numrays, RadeonRays::Buffer* hitresults,
                        RadeonRays::Event const* waitevent, RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_query_occlusion(h_, Cast(rays)->handle(), numrays, Cast(hitresults)->handle(), Ev(waitevent),
                                  event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }
    void QueryIntersection(RadeonRays::Buffer const* rays, RadeonRays::Buffer const* numrays, int maxrays,
                           RadeonRays::Buffer* hitinfos, RadeonRays::Event const* waitevent,
                           RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_query_intersection_indirect(h_, Cast(rays)->handle(), Cast(numrays)->handle(), maxrays,
                                              Cast(hitinfos)->handle(), Ev(waitevent), event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }
    void QueryOcclusion(RadeonRays::Buffer const* rays, RadeonRays::Buffer const* numrays, int maxrays,
                        RadeonRays::Buffer* hitresults, RadeonRays::Event const* waitevent,
                        RadeonRays::Event** event) const override {
        rrb_event* e = nullptr;
        Check(rrb_query_occlusion_indirect(h_, Cast(rays)->handle(), Cast(numrays)->handle(), maxrays,
                                           Cast(hitresults)->handle(), Ev(waitevent), event ? &e : nullptr));
        if (event) *event = new EventB200(e);
    }

    // ---- options (radeon_rays.h:279-296): unknown keys are stored, never an error
    void SetOption(char const* name, char const* value) override { Check(rrb_set_option_str(h_, name, value)); }
    void SetOption(char const* name, float value) override { Check(rrb_set_option_f32(h_, name, value)); }

    // Replacement of CreateFromOpenClBuffer (radeon_rays_cl.h:41).  The engine wraps the same device
    // buffer on EVERY query and never frees the wrapper (RTPathTracingPass.cpp:82-83); wrappers are
    // therefore cached by (pointer, size) so the leak costs nothing.
    RadeonRays::Buffer* WrapDevicePtr(void* device_ptr, size_t bytes) {
        auto it = wrappers_.find(device_ptr);
        if (it != wrappers_.end() && it->second.bytes == bytes) return it->second.buf;
        rrb_buffer* b = nullptr;
        Check(rrb_wrap_device_ptr(h_, device_ptr, bytes, &b));
        BufferB200* w = new BufferB200(b);
        if (it != wrappers_.end()) {
            rrb_delete_buffer(h_, it->second.buf->handle());
            delete it->second.buf;
            wrappers_.erase(it);
        }
        wrappers_.emplace(device_ptr, Wrapper{w, bytes});
        return w;
    }

private:
    struct Wrapper {
        BufferB200* buf;
        size_t bytes;
    };
    static ShapeB200 const* Cast(RadeonRays::Shape const* s) { return static_cast<ShapeB200 const*>(s); }
    static BufferB200 const* Cast(RadeonRays::Buffer const* b) { return static_cast<BufferB200 const*>(b); }
    static BufferB200* Cast(RadeonRays::Buffer* b) { return static_cast<BufferB200*>(b); }
    static rrb_event const* Ev(RadeonRays::Event const* e) {
        return e ? static_cast<EventB200 const*>(e)->handle() : nullptr;
    }
    rrb_api* h_ = nullptr;
    std::unordered_map<void*, Wrapper> wrappers_;
};

// Factory functions with the shape of IntersectionApi::Create / Delete (radeon_rays.h:200-203) and
// CreateFromOpenClContext / CreateFromOpenClBuffer (radeon_rays_cl.h:40-41).
inline RadeonRays::IntersectionApi* Create(int cuda_device, void* cuda_stream = nullptr) {
    return new IntersectionApiB200(cuda_device, cuda_stream);
}
inline void Delete(RadeonRays::IntersectionApi* api) { delete static_cast<IntersectionApiB200*>(api); }
inline RadeonRays::Buffer* CreateFromDevicePtr(RadeonRays::IntersectionApi* api, void* device_ptr, size_t bytes) {
    return static_cast<IntersectionApiB200*>(api)->WrapDevicePtr(device_ptr, bytes);
}

}  // namespace RadeonRaysB200
