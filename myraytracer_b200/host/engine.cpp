// engine.cpp — rthost::Renderer / RayTracerEngine (see raytracer_host.hpp) and the
// small `rth_*` C surface Python tests and bench.py drive them through.
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "raytracer_host.hpp"

namespace rthost {

rt_camera toRtCamera(const pk::Camera& c) {
    rt_camera r{};
    r.width = c.width; r.height = c.height; r.numSamples = c.numSamples;
    r.isLookAt = c.isLookAt ? 1 : 0; r.hasFovy = c.hasFovy ? 1 : 0;
    const pk::V3* src[4] = {&c.position, &c.gaze, &c.gazePoint, &c.up};
    double* dst[4] = {r.position, r.gaze, r.gazePoint, r.up};
    for (int k = 0; k < 4; k++) { dst[k][0] = src[k]->x; dst[k][1] = src[k]->y; dst[k][2] = src[k]->z; }
    for (int k = 0; k < 4; k++) r.nearPlane[k] = c.nearPlane[k];
    r.nearDistance = c.nearDistance; r.fovy = c.fovy;
    r.apertureSize = c.apertureSize; r.focusDistance = c.focusDistance;
    return r;
}

Renderer::Renderer(const RTContext& ctx, const pk::Scene& scene, int device) {
    rt_scene_desc d = ctx.describe(scene);
    handle_ = rt_scene_create(&d, device);
    if (!handle_) error_ = rt_last_error();
}

Renderer::Renderer(const RTContext& ctx, const pk::Scene& scene, const std::vector<int>& devices) {
    rt_scene_desc d = ctx.describe(scene);
    std::vector<int32_t> dev(devices.begin(), devices.end());
    handle_ = rt_scene_create_multi(&d, dev.data(), (int32_t)dev.size());
    if (!handle_) error_ = rt_last_error();
}

Renderer::~Renderer() { rt_scene_destroy(handle_); }

int Renderer::render(const pk::Scene& scene, int cameraIndex, std::vector<double>* linear, std::vector<uint8_t>* rgba,
                     std::vector<rt_hit_id>* ids, rt_stats* stats, const rt_render_params* params) {
    if (cameraIndex < 0 || cameraIndex >= (int)scene.cameras.size()) return RT_ERR_BAD_CAMERA;
    const pk::Camera& cam = scene.cameras[cameraIndex];
    const size_t W = (size_t)std::max(1, cam.width), H = (size_t)std::max(1, cam.height);
    const size_t n = (size_t)std::sqrt((double)std::max(1, cam.numSamples));
    if (linear) linear->assign(W * H * 3, 0.0);
    if (rgba) rgba->assign(W * H * 4, 0);
    if (ids) ids->assign(W * H * n * n, rt_hit_id{});
    rt_camera rc = toRtCamera(cam);
    return rt_render(handle_, &rc, params, rgba ? rgba->data() : nullptr, linear ? linear->data() : nullptr,
                     ids ? ids->data() : nullptr, stats);
}

// Where PLY meshes are ingested: RTCUDA_INGEST=host keeps the host loader; otherwise the engine's CUDA device when
// there is one (the render path needs it anyway).
static int ingestDeviceFor(int device) {
    const char* e = std::getenv("RTCUDA_INGEST");
    if (e && std::strcmp(e, "host") == 0) return -1;
    return rt_device_count() > device && device >= 0 ? device : -1;
}

int Renderer::refit(RTContext& ctx) {
    return rt_scene_update_instances(handle_, ctx.instances.data(), (int64_t)ctx.instances.size(), nullptr);
}

int RayTracerEngine::setInstanceTransforms(const double* l2w, int64_t n, RenderError& err, bool boundsOnDevice) {
    if (!hasScene_) { err = {"Render", -20, "No scene loaded. Can't render."}; return -20; }
    if (!renderer_) { err = {"Render", -21, "Renderer not initialized."}; return -21; }
    if (!l2w || n != (int64_t)ctx_.instances.size()) { err = {"Render", RT_ERR_INVALID, "one transform per instance is required"}; return RT_ERR_INVALID; }
    for (int64_t i = 0; i < n; i++) ctx_.setInstanceTransform((size_t)i, l2w + i * 16, boundsOnDevice);
    const int rc = renderer_->refit(ctx_);
    if (rc != RT_OK) err = {"Render", rc, rt_last_error()};
    return rc;
}

RayTracerEngine::RayTracerEngine(const std::string& url, int device) {
    std::string err;
    if (!pk::loadSceneBlob(url.c_str(), scene_, err)) {   // "Error decoding scene" -> engine stays empty (:38-40)
        loadError_ = err;
        return;
    }
    hasScene_ = true;
    ctx_ = RTContext(scene_, ingestDeviceFor(device));     // RTContext(scene:) :44
    renderer_.reset(new Renderer(ctx_, scene_, device));   // Renderer(ctx:) :45
    if (!renderer_->valid()) { loadError_ = renderer_->error(); renderer_.reset(); }
}

RayTracerEngine::RayTracerEngine(const std::string& url, const std::vector<int>& devices) {
    std::string err;
    if (!pk::loadSceneBlob(url.c_str(), scene_, err)) { loadError_ = err; return; }
    hasScene_ = true;
    ctx_ = RTContext(scene_, ingestDeviceFor(devices.empty() ? 0 : devices[0]));
    renderer_.reset(new Renderer(ctx_, scene_, devices));
    if (!renderer_->valid()) { loadError_ = renderer_->error(); renderer_.reset(); }
}

void RayTracerEngine::sceneMeshAndTriangleCounts(int64_t& meshes, int64_t& tris, int64_t& spheres, int64_t& planes) const {
    meshes = tris = spheres = planes = 0;
    for (const pk::SceneObject& o : scene_.objects) {      // :211-225
        switch (o.kind) {
            case pk::OBJ_SPHERE: spheres++; break;
            case pk::OBJ_MESH: meshes++; tris += (int64_t)o.faces.size() / 3; break;
            case pk::OBJ_TRIANGLE: tris++; break;
            case pk::OBJ_PLANE: planes++; break;
            default: break;
        }
    }
}

SceneInfo RayTracerEngine::inspect() const {
    SceneInfo si;
    for (size_t i = 0; i < scene_.cameras.size(); i++) {
        const pk::Camera& c = scene_.cameras[i];
        si.cameras.push_back(CameraSpec{(int)i, c.id, c.imageName, c.width, c.height});
    }
    sceneMeshAndTriangleCounts(si.meshes, si.triangles, si.spheres, si.planes);
    return si;
}

int RayTracerEngine::render(int cameraIndex, std::vector<uint8_t>& rgba, RenderStats& stats, CameraSpec& spec,
                            RenderError& err, std::vector<double>* linear, std::vector<rt_hit_id>* ids,
                            const rt_render_params* params) {
    if (!hasScene_) { err = {"Render", -20, "No scene loaded. Can't render."}; return -20; }          // :142-145
    if (!renderer_) { err = {"Render", -21, "Renderer not initialized."}; return -21; }              // :146-149
    if (cameraIndex < 0 || cameraIndex >= (int)scene_.cameras.size()) { err = {"Ray", -10, "Invalid camera index"}; return -10; }  // :152-155
    const pk::Camera& cam = scene_.cameras[cameraIndex];
    spec = CameraSpec{cameraIndex, cam.id, cam.imageName, cam.width, cam.height};
    int64_t meshes, tris, spheres, planes;
    sceneMeshAndTriangleCounts(meshes, tris, spheres, planes);
    const auto started = std::chrono::steady_clock::now();                                           // :166
    rt_stats st{};
    int rc = renderer_->render(scene_, cameraIndex, linear, &rgba, ids, &st, params);
    if (rc != RT_OK) { err = {"Render", rc, rt_last_error()}; return rc; }
    stats.meshes = meshes; stats.triangles = tris; stats.spheres = spheres; stats.planes = planes;
    stats.rays = st.rays; stats.raysPrimary = st.raysPrimary; stats.raysShadow = st.raysShadow; stats.raysSecondary = st.raysSecondary;
    stats.deviceMs = st.deviceMs;
    stats.milliseconds = (int64_t)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - started).count();
    return 0;
}

}  // namespace rthost

// ---------------------------------------------------------------------------------------------
// C surface for ctypes (declared in include/rtcuda_host.h)
// ---------------------------------------------------------------------------------------------
#include "../../include/rtcuda_host.h"

namespace {
thread_local std::string g_herr;
}

extern "C" {

const char* rth_last_error(void) { return g_herr.c_str(); }

rth_engine_t* rth_engine_create(const char* sceneUrl, int32_t device) {
    try {
        auto* e = new rthost::RayTracerEngine(sceneUrl ? sceneUrl : "", device);
        if (!e->hasScene() || !e->renderer()) g_herr = e->loadError();
        return reinterpret_cast<rth_engine_t*>(e);
    } catch (const std::exception& ex) {
        g_herr = ex.what();
    } catch (...) {
        g_herr = "unknown exception";
    }
    return nullptr;
}

rth_engine_t* rth_engine_create_multi(const char* sceneUrl, const int32_t* devices, int32_t nDevices) {
    try {
        std::vector<int> dev;
        for (int32_t k = 0; devices && k < nDevices; k++) dev.push_back(devices[k]);
        auto* e = new rthost::RayTracerEngine(sceneUrl ? sceneUrl : "", dev);
        if (!e->hasScene() || !e->renderer()) g_herr = e->loadError();
        return reinterpret_cast<rth_engine_t*>(e);
    } catch (const std::exception& ex) {
        g_herr = ex.what();
    } catch (...) {
        g_herr = "unknown exception";
    }
    return nullptr;
}

void rth_engine_destroy(rth_engine_t* h) { delete reinterpret_cast<rthost::RayTracerEngine*>(h); }

int32_t rth_engine_ready(const rth_engine_t* h) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    return (e && e->hasScene() && e->renderer()) ? 1 : 0;
}

int32_t rth_engine_camera_count(const rth_engine_t* h) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    return (e && e->hasScene()) ? (int32_t)e->scene().cameras.size() : 0;
}

int32_t rth_engine_camera(const rth_engine_t* h, int32_t idx, rt_camera* out) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene() || !out) { g_herr = "No scene loaded. Can't render."; return -20; }
    if (idx < 0 || idx >= (int32_t)e->scene().cameras.size()) { g_herr = "Invalid camera index"; return -10; }
    *out = rthost::toRtCamera(e->scene().cameras[idx]);
    return 0;
}

rt_scene_t* rth_engine_scene(const rth_engine_t* h) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    return (e && e->renderer()) ? e->renderer()->handle() : nullptr;
}

int32_t rth_engine_describe(const rth_engine_t* h, rt_scene_desc* out) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene() || !out) { g_herr = "No scene loaded. Can't render."; return -20; }
    *out = e->context().describe(e->scene());
    return 0;
}

int32_t rth_engine_instance_object_ids(const rth_engine_t* h, int32_t* out, int64_t cap) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene()) return -20;
    const auto& v = e->context().instanceObjectIds;
    for (int64_t i = 0; i < cap && i < (int64_t)v.size(); i++) out[i] = v[i];
    return (int32_t)v.size();
}

int32_t rth_engine_render(rth_engine_t* h, int32_t cameraIndex, const rt_render_params* params, uint8_t* rgbaOut,
                          double* linearOut, rt_hit_id* idsOut, rt_stats* statsOut) {
    auto* e = reinterpret_cast<rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene()) { g_herr = "No scene loaded. Can't render."; return -20; }
    if (!e->renderer()) { g_herr = "Renderer not initialized."; return -21; }
    if (cameraIndex < 0 || cameraIndex >= (int32_t)e->scene().cameras.size()) { g_herr = "Invalid camera index"; return -10; }
    rt_camera cam = rthost::toRtCamera(e->scene().cameras[cameraIndex]);
    int32_t rc = rt_render(e->renderer()->handle(), &cam, params, rgbaOut, linearOut, idsOut, statsOut);
    if (rc != 0) g_herr = rt_last_error();
    else if (statsOut) {
        int64_t m, t, s, p;
        e->sceneMeshAndTriangleCounts(m, t, s, p);
        statsOut->meshes = m; statsOut->spheres = s; statsOut->planes = p;
    }
    return rc;
}

int32_t rth_engine_set_instance_transforms(rth_engine_t* h, const double* l2w, int64_t n, int32_t boundsOnDevice) {
    auto* e = reinterpret_cast<rthost::RayTracerEngine*>(h);
    if (!e) { g_herr = "No scene loaded. Can't render."; return -20; }
    rthost::RenderError err{"", 0, ""};
    const int rc = e->setInstanceTransforms(l2w, n, err, boundsOnDevice != 0);
    if (rc != 0) g_herr = err.message;
    return rc;
}

int32_t rth_engine_inspect(const rth_engine_t* h, int64_t* counts4, int32_t* camerasOut) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene()) { g_herr = "No scene loaded. Can't render."; return -20; }
    const rthost::SceneInfo si = e->inspect();
    if (counts4) { counts4[0] = si.meshes; counts4[1] = si.triangles; counts4[2] = si.spheres; counts4[3] = si.planes; }
    if (camerasOut) *camerasOut = (int32_t)si.cameras.size();
    return 0;
}

int32_t rth_engine_camera_name(const rth_engine_t* h, int32_t idx, char* out, int32_t capacity, int32_t* idOut) {
    auto* e = reinterpret_cast<const rthost::RayTracerEngine*>(h);
    if (!e || !e->hasScene()) { g_herr = "No scene loaded. Can't render."; return -20; }
    if (idx < 0 || idx >= (int32_t)e->scene().cameras.size()) { g_herr = "Invalid camera index"; return -10; }
    const pk::Camera& c = e->scene().cameras[idx];
    if (out && capacity > 0) { std::strncpy(out, c.imageName.c_str(), (size_t)capacity - 1); out[capacity - 1] = 0; }
    if (idOut) *idOut = c.id;
    return 0;
}

// Flatten only (no GPU): lets CPU tests compare the mirror's RTContext with the oracle's.
rth_context_t* rth_context_create(const char* sceneUrl) {
    try {
        auto* s = new std::pair<pk::Scene, rthost::RTContext>();
        std::string err;
        if (!pk::loadSceneBlob(sceneUrl, s->first, err)) { g_herr = err; delete s; return nullptr; }
        const char* ing = std::getenv("RTCUDA_INGEST");   // flatten-only contexts ingest on the host unless asked
        s->second = rthost::RTContext(s->first, (ing && std::strcmp(ing, "device") == 0 && rt_device_count() > 0) ? 0 : -1);
        return reinterpret_cast<rth_context_t*>(s);
    } catch (const std::exception& ex) {
        g_herr = ex.what();
    } catch (...) {
        g_herr = "unknown exception";
    }
    return nullptr;
}
void rth_context_destroy(rth_context_t* c) { delete reinterpret_cast<std::pair<pk::Scene, rthost::RTContext>*>(c); }
int32_t rth_context_describe(const rth_context_t* c, rt_scene_desc* out) {
    auto* s = reinterpret_cast<const std::pair<pk::Scene, rthost::RTContext>*>(c);
    if (!s || !out) return -20;
    *out = s->second.describe(s->first);
    return 0;
}

}  // extern "C"
