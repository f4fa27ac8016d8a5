// oracle/oracle.cpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// C entry points (ctypes) over the FP64 restatement of the reference's render
// path (see vecmath.hpp, bvh.hpp, rtcontext*.hpp, traverse.hpp, render*.hpp for
// the file:line each function follows).  Only tests/, __graft_entry__.smoke()
// and bench.py's CPU-baseline / --impl reference legs may load this library.
//
// PARITY STATUS: the reference has no tests, fixtures or golden vectors
//  and cannot be built here (no Swift toolchain, Apple-only
// imports, ParsingKit absent), so the render restatement is "parity unpinned";
// it is held by per-function file:line traceability, hand-derived KATs and a
// second, independently written restatement in Python without any shared code
// (tests/refmath.py, reftrace.py, refbvh.py: PCG32 and seed, both cameras and
// the thin lens, BVHBuilder, traversal order and tie-breaking, instance
// transforms and motion, smooth normals, area lights, the whole trace
// recursion and the RGBA8 rule agree bit for bit or to rounding; DESIGN.md
// section 7 lists the tests).  The one part pinned by REAL reference
// code is PLY ingest: oracle/_ref/libcply.so is compiled from the reference's
// own Sources/CPly and driven through its ply_reader_* ABI.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>

#include "render_impl.hpp"
#include "rtcontext_init.hpp"

using namespace orc;

namespace {
thread_local std::string g_err;

struct OracleScene {
    pk::Scene scene;
    RTContext ctx;
    double buildMs = 0;
};
}  // namespace

extern "C" {

struct orc_info {
    int64_t triangles, spheres, planes, instances, cameras, materials, pointLights, areaLights;
    int64_t blasNodesTotal, tlasNodes, maxBuildDepth;
    int32_t plyViaReference;
    double buildMs;
};

struct orc_camera_info {
    int32_t width, height, numSamples, sqrtSamples;
};

struct orc_counters {
    int64_t raysPrimary, raysShadow, raysSecondary, nodeVisits, boxTests, triTests, sphereTests, planeTests,
        instEntries, maxStack;
};

struct orc_primary_hit {
    int32_t inst, kind;
    int64_t prim;
    double t;
};

const char* orc_last_error() { return g_err.c_str(); }

// cplyLib: path of oracle/_ref/libcply.so, or NULL/"" to use the built-in PLY reader.
void* orc_scene_load(const char* path, const char* cplyLib) {
    try {
        if (cplyLib && *cplyLib) {
            if (!cplyApi(cplyLib).ok) { g_err = std::string("cannot load CPly from ") + cplyLib; return nullptr; }
        }
        auto* s = new OracleScene();
        std::string err;
        if (!pk::loadSceneBlob(path, s->scene, err)) { g_err = err; delete s; return nullptr; }
        auto t0 = std::chrono::steady_clock::now();
        s->ctx.init(s->scene);
        s->buildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        return s;
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    } catch (...) {
        g_err = "unknown exception";
        return nullptr;
    }
}

void orc_scene_free(void* h) { delete static_cast<OracleScene*>(h); }

int orc_scene_info(void* h, orc_info* out) {
    if (!h || !out) return -1;
    auto* s = static_cast<OracleScene*>(h);
    std::memset(out, 0, sizeof *out);
    out->triangles = (int64_t)s->ctx.triangles.size();
    out->spheres = (int64_t)s->ctx.spheres.size();
    out->planes = (int64_t)s->ctx.planes.size();
    out->instances = (int64_t)s->ctx.instances.size();
    out->cameras = (int64_t)s->scene.cameras.size();
    out->materials = (int64_t)s->ctx.materials.size();
    out->pointLights = (int64_t)s->ctx.pointLights.size();
    out->areaLights = (int64_t)s->ctx.areaLights.size();
    const BLAS* last = nullptr;
    for (const Instance& i : s->ctx.instances) {
        if (i.blas.get() != last) out->blasNodesTotal += (int64_t)i.blas->nodes.size();
        last = i.blas.get();
        if (i.blas->buildDepth > out->maxBuildDepth) out->maxBuildDepth = i.blas->buildDepth;
    }
    out->tlasNodes = s->ctx.hasTlas ? (int64_t)s->ctx.tlas.builder.nodesUsed + 1 : 0;
    out->plyViaReference = s->ctx.plyViaReference ? 1 : 0;
    out->buildMs = s->buildMs;
    return 0;
}

int orc_camera_info(void* h, int cameraIndex, orc_camera_info* out) {
    if (!h || !out) return -1;
    auto* s = static_cast<OracleScene*>(h);
    if (cameraIndex < 0 || cameraIndex >= (int)s->scene.cameras.size()) { g_err = "Invalid camera index"; return -10; }
    const pk::Camera& c = s->scene.cameras[cameraIndex];
    out->width = std::max(1, c.width);
    out->height = std::max(1, c.height);
    out->numSamples = std::max(1, c.numSamples);
    out->sqrtSamples = (int)std::sqrt((double)out->numSamples);
    return 0;
}

// linear: W*H*3 doubles (0..255 scale) or NULL; rgba: W*H*4 or NULL;
// ids: W*H*n*n records or NULL.  renderMs covers render + RGBA8 conversion
// (the reference's own timer scope, RayTracer.swift:166-197).
int orc_render(void* h, int cameraIndex, int threads, int chunkStride, int chunkOffset, double* linear, uint8_t* rgba,
               orc_primary_hit* ids, orc_counters* counters, double* renderMs) {
    if (!h) { g_err = "No scene loaded. Can't render."; return -20; }
    auto* s = static_cast<OracleScene*>(h);
    if (cameraIndex < 0 || cameraIndex >= (int)s->scene.cameras.size()) { g_err = "Invalid camera index"; return -10; }
    try {
        const pk::Camera& cam = s->scene.cameras[cameraIndex];
        const size_t W = (size_t)std::max(1, cam.width), H = (size_t)std::max(1, cam.height);
        std::vector<Vec3> out(W * H);
        Renderer R(s->ctx, s->scene);
        RenderOptions opt;
        opt.threads = threads;
        opt.chunkStride = chunkStride < 1 ? 1 : chunkStride;
        opt.chunkOffset = chunkOffset;
        Counters C;
        static_assert(sizeof(PrimaryHit) == sizeof(orc_primary_hit), "layout");
        std::vector<uint8_t> tmpRgba;
        uint8_t* rg = rgba;
        if (!rg) { tmpRgba.resize(W * H * 4); rg = tmpRgba.data(); }
        auto t0 = std::chrono::steady_clock::now();
        R.render(cameraIndex, out.data(), reinterpret_cast<PrimaryHit*>(ids), C, opt);
        toRGBA8(out.data(), W * H, rg);
        double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (renderMs) *renderMs = ms;
        if (linear) std::memcpy(linear, out.data(), W * H * sizeof(Vec3));
        if (counters) {
            counters->raysPrimary = C.raysPrimary; counters->raysShadow = C.raysShadow;
            counters->raysSecondary = C.raysSecondary; counters->nodeVisits = C.nodeVisits;
            counters->boxTests = C.boxTests; counters->triTests = C.triTests; counters->sphereTests = C.sphereTests;
            counters->planeTests = C.planeTests; counters->instEntries = C.instEntries; counters->maxStack = C.maxStack;
        }
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

// ---- flattened-context readback (host-mirror flatten is compared against this) ----
// v: n*9 doubles (v0,v1,v2); nrm: n*9 doubles (n0,n1,n2); either may be NULL.
int orc_get_triangles(void* h, double* v, double* nrm) {
    if (!h) return -1;
    auto* s = static_cast<OracleScene*>(h);
    size_t k = 0;
    for (const Triangle& t : s->ctx.triangles) {
        const Vec3 vs[3] = {t.v0, t.v1, t.v2}, ns[3] = {t.n0, t.n1, t.n2};
        for (int c = 0; c < 3; c++) {
            if (v) { v[k * 9 + c * 3] = vs[c].x; v[k * 9 + c * 3 + 1] = vs[c].y; v[k * 9 + c * 3 + 2] = vs[c].z; }
            if (nrm) { nrm[k * 9 + c * 3] = ns[c].x; nrm[k * 9 + c * 3 + 1] = ns[c].y; nrm[k * 9 + c * 3 + 2] = ns[c].z; }
        }
        k++;
    }
    return 0;
}

// per instance: l2w[16], w2l[16], nmat[9] (column-major), bounds[6], meta[4] = {objectId, materialOverride, primCount, firstPrimIndex}
int orc_get_instances(void* h, double* l2w, double* w2l, double* nmat, double* bounds, int64_t* meta) {
    if (!h) return -1;
    auto* s = static_cast<OracleScene*>(h);
    size_t k = 0;
    for (const Instance& i : s->ctx.instances) {
        if (l2w) std::memcpy(l2w + k * 16, i.localToWorld.m, 16 * 8);
        if (w2l) std::memcpy(w2l + k * 16, i.worldToLocal.m, 16 * 8);
        if (nmat)
            for (int c = 0; c < 3; c++) {
                nmat[k * 9 + c * 3] = i.normalMatrix.c[c].x; nmat[k * 9 + c * 3 + 1] = i.normalMatrix.c[c].y;
                nmat[k * 9 + c * 3 + 2] = i.normalMatrix.c[c].z;
            }
        if (bounds) {
            bounds[k * 6] = i.worldBounds.minP.x; bounds[k * 6 + 1] = i.worldBounds.minP.y; bounds[k * 6 + 2] = i.worldBounds.minP.z;
            bounds[k * 6 + 3] = i.worldBounds.maxP.x; bounds[k * 6 + 4] = i.worldBounds.maxP.y; bounds[k * 6 + 5] = i.worldBounds.maxP.z;
        }
        if (meta) {
            meta[k * 4] = i.objectId;
            meta[k * 4 + 1] = i.hasMaterialOverride ? i.materialOverride : INT64_MIN;
            meta[k * 4 + 2] = (int64_t)i.blas->prims.size();
            meta[k * 4 + 3] = i.blas->prims.empty() ? -1 : i.blas->prims[0].primitiveIndex;
        }
        k++;
    }
    return 0;
}

}  // extern "C"
