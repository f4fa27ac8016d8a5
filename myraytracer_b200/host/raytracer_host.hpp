// raytracer_host.hpp — host-side mirror of the reference's Swift interface
// around the render path (the Swift toolchain is absent here, so the host side
// above the C ABI is C++; swift/ holds the Swift glue a maintainer would add).
//
//   rthost::RTContext        <- struct RTContext / init(scene:)   Models/RTContext.swift:74-418
//   rthost::Renderer         <- final class Renderer              Extensions/Object+Extension.swift:13-19, 52-379
//   rthost::RayTracerEngine  <- final class RayTracerEngine       RayTracer.swift:25-205
//   RenderStats / CameraSpec / SceneInfo                          Models/RenderStats.swift, RenderConfig.swift
// RTContext flattens the scene exactly like the reference (triangle set-up,
// unweighted smooth normals, instance matrices and world bounds) but does NOT
// build the reference's CPU BVHs: Renderer hands the flat arrays to
// rt_scene_create, which builds the GPU layout.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "../../include/rtcuda.h"
#include "scene_model.hpp"

namespace rthost {

// std::allocator whose construct() default-initialises: resize() of a vector of doubles or of plain records leaves
// the new elements unwritten instead of zero-filling them on one thread.  The per-triangle arrays of a device-ingested
// mesh are hundreds of MB that rt_ply_ingest / the flatten's own threads overwrite in full; zero-filling them first
// was 100 of the 125 ms a 1M-triangle flatten took (page faults on one thread, profiles/r2al_build_trace.txt).
template <class T>
struct DefaultInitAllocator : std::allocator<T> {
    template <class U> struct rebind { using other = DefaultInitAllocator<U>; };
    DefaultInitAllocator() noexcept {}
    template <class U> DefaultInitAllocator(const DefaultInitAllocator<U>&) noexcept {}
    template <class U> void construct(U* p) { ::new (static_cast<void*>(p)) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};
using RawDoubles = std::vector<double, DefaultInitAllocator<double>>;

struct RTContext {
    double intersectionTestEpsilon = 0, shadowRayEpsilon = 0;
    std::vector<rt_material> materials;
    std::vector<rt_sphere> spheres;
    std::vector<rt_plane> planes;
    RawDoubles triVertices;            // 9 per triangle
    RawDoubles triNormals;             // 9 per triangle
    RawDoubles triMotionBlur;          // 3 per triangle
    std::vector<rt_point_light> pointLights;
    std::vector<rt_area_light> areaLights;
    std::vector<rt_prim, DefaultInitAllocator<rt_prim>> prims;
    std::vector<rt_blas> blas;
    std::vector<rt_instance> instances;
    std::vector<int32_t> instanceObjectIds;
    std::vector<std::string> warnings;
    bool anyMotionBlur = false;
    // local-space bounds of every prim of a BLAS (6 doubles each): Instance.worldBounds is the union of their
    // transformed boxes (makeInstance :437-447, refitTLAS :891-900)
    std::vector<RawDoubles> blasPrimBounds;

    RTContext() {}
    // ingestDevice >= 0: PLY meshes are ingested on that CUDA device (rt_ply_ingest) where the file layout allows;
    // -1: everything on the host.  Same arrays either way.
    explicit RTContext(const pk::Scene& scene, int ingestDevice = -1);
    int64_t deviceIngestedTriangles = 0;
    double deviceIngestMs = 0;
    int64_t triangleCount() const { return (int64_t)triVertices.size() / 9; }
    // Animation: a new localToWorld for instance i — worldToLocal, normalMatrix and worldBounds re-derived exactly as
    // RTContext.refitTLAS step 1 does (Models/RTContext.swift:889-912).  m16: column-major 4x4.
    void setInstanceTransform(size_t i, const double* m16, bool boundsOnDevice = false);
    // borrowed view for rt_scene_create (valid while *this is alive and unmodified)
    rt_scene_desc describe(const pk::Scene& scene) const;
};

struct RenderStats {  // Models/RenderStats.swift:8-23 (+ the ray split the reference never fills)
    int64_t meshes = 0, triangles = 0, spheres = 0, planes = 0, rays = 0, milliseconds = 0;
    int64_t raysPrimary = 0, raysShadow = 0, raysSecondary = 0;
    double deviceMs = 0;
};

struct CameraSpec { int index = 0; int id = 0; std::string imageName; int width = 0, height = 0; };
struct SceneInfo { std::vector<CameraSpec> cameras; int64_t meshes = 0, triangles = 0, spheres = 0, planes = 0; };

struct RenderError {  // NSError(domain:code:) of RayTracer.swift:121-123,142-154
    std::string domain; int code; std::string message;
};

class Renderer {
public:
    // Renderer(ctx:) — uploads the context to `device` (Object+Extension.swift:16-18)
    Renderer(const RTContext& ctx, const pk::Scene& scene, int device);
    // the same scene resident on several GPUs of this process (rt_scene_create_multi): render() shards the frame
    Renderer(const RTContext& ctx, const pk::Scene& scene, const std::vector<int>& devices);
    ~Renderer();
    Renderer(const Renderer&) = delete;
    Renderer& operator=(const Renderer&) = delete;
    bool valid() const { return handle_ != nullptr; }
    const std::string& error() const { return error_; }
    rt_scene_t* handle() const { return handle_; }
    // render(scene:cameraIndex:progressRow:) -> [Vec3] (W*H*3 doubles, 0..255 scale); also fills rgba when given
    int render(const pk::Scene& scene, int cameraIndex, std::vector<double>* linear, std::vector<uint8_t>* rgba,
               std::vector<rt_hit_id>* ids, rt_stats* stats, const rt_render_params* params = nullptr);
    // refitTLAS steps 2-4 (RTContext.swift:914-927): the instances of `ctx` replace the scene's, the TLAS boxes are
    // refitted on the device, the BLASes stay
    int refit(RTContext& ctx);

private:
    rt_scene_t* handle_ = nullptr;
    std::string error_;
};

rt_camera toRtCamera(const pk::Camera& c);

class RayTracerEngine {
public:
    // init(from url:) — RayTracer.swift:30-49; a decode failure leaves the engine empty (:38-40)
    explicit RayTracerEngine(const std::string& url, int device = 0);
    RayTracerEngine(const std::string& url, const std::vector<int>& devices);   // one engine, several GPUs
    bool hasScene() const { return hasScene_; }
    const pk::Scene& scene() const { return scene_; }
    const RTContext& context() const { return ctx_; }
    Renderer* renderer() const { return renderer_.get(); }
    const std::string& loadError() const { return loadError_; }
    SceneInfo inspect() const;                                        // :52-67
    // Animation: n = instance count, 16 doubles (column-major localToWorld) per instance, in
    // the order of context().instances.  The reference's batch driver reloads and rebuilds everything per frame
    // (BatchRender/BatchRender.swift:26-40); here only the instance records and the TLAS boxes change.
    int setInstanceTransforms(const double* l2w, int64_t n, RenderError& err, bool boundsOnDevice = true);
    // sceneMeshAndTriangleCounts(_:) — :208-227
    void sceneMeshAndTriangleCounts(int64_t& meshes, int64_t& tris, int64_t& spheres, int64_t& planes) const;
    // render(format:cameraIndex:progress:) — :115-131 via renderRGBA8Async :137-205.
    // Returns 0 or fills `err` with the reference's domain/code.
    int render(int cameraIndex, std::vector<uint8_t>& rgba, RenderStats& stats, CameraSpec& spec, RenderError& err,
               std::vector<double>* linear = nullptr, std::vector<rt_hit_id>* ids = nullptr,
               const rt_render_params* params = nullptr);

private:
    pk::Scene scene_;
    RTContext ctx_;
    std::unique_ptr<Renderer> renderer_;
    bool hasScene_ = false;
    std::string loadError_;
};

}  // namespace rthost
