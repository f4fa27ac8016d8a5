// rt_context.cpp — rthost::RTContext: scene flattening of RTContext.init(scene:)
// (Models/RTContext.swift:94-418) into the flat arrays of include/rtcuda.h.
// Compile with -ffp-contract=off (no FMA contraction, like the reference's x86-64 Swift).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <system_error>
#include <thread>
#ifdef __linux__
#include <sys/mman.h>
#endif

#include "ply_loader.hpp"
#include "raytracer_host.hpp"

namespace rthost {
namespace {

struct D3 { double x, y, z; };
inline D3 d3(const pk::V3& v) { return {v.x, v.y, v.z}; }
inline D3 operator+(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline D3 operator-(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline D3 operator*(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline double dot(D3 a, D3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 normalize(D3 a) { return a * (1.0 / std::sqrt(dot(a, a))); }
inline double mn(double a, double b) { return a < b ? a : b; }
inline double mx(double a, double b) { return a > b ? a : b; }
inline void put(double* d, D3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

// simd_double4x4.inverse stand-in: Gauss-Jordan with partial pivoting (RTContext.swift:451)
void inverse4(const double* M, double* R) {
    double a[4][8];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) { a[r][c] = M[c * 4 + r]; a[r][c + 4] = (r == c) ? 1.0 : 0.0; }
    for (int col = 0; col < 4; col++) {
        int piv = col;
        for (int r = col + 1; r < 4; r++)
            if (std::fabs(a[r][col]) > std::fabs(a[piv][col])) piv = r;
        if (piv != col)
            for (int c = 0; c < 8; c++) { double t = a[col][c]; a[col][c] = a[piv][c]; a[piv][c] = t; }
        const double d = 1.0 / a[col][col];
        for (int c = 0; c < 8; c++) a[col][c] *= d;
        for (int r = 0; r < 4; r++) {
            if (r == col) continue;
            const double f = a[r][col];
            if (f == 0.0) continue;
            for (int c = 0; c < 8; c++) a[r][c] -= f * a[col][c];
        }
    }
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) R[c * 4 + r] = a[r][c + 4];
}

// normalTransformMatrix(from:) = (upper 3x3).inverse.transpose (RTContext.swift:23-31), column-major out[9]
void normalMatrix(const double* M, double* out) {
    const D3 c0{M[0], M[1], M[2]}, c1{M[4], M[5], M[6]}, c2{M[8], M[9], M[10]};
    D3 r0 = cross(c1, c2), r1 = cross(c2, c0), r2 = cross(c0, c1);
    const double inv = 1.0 / dot(c0, r0);
    r0 = r0 * inv; r1 = r1 * inv; r2 = r2 * inv;
    // inverse rows are r0,r1,r2; the transpose of the inverse has them as columns
    put(out, r0); put(out + 3, r1); put(out + 6, r2);
}

// AABB.transformed(by:) (Accelearion/AABB.swift:71-92)
void transformBox(const double* lo, const double* hi, const double* M, double* olo, double* ohi) {
    const D3 cOld = D3{lo[0] + hi[0], lo[1] + hi[1], lo[2] + hi[2]} * 0.5;
    const D3 eOld = D3{hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]} * 0.5;
    double cN[3], eN[3];
    for (int r = 0; r < 3; r++) {
        cN[r] = ((M[r] * cOld.x + M[4 + r] * cOld.y) + M[8 + r] * cOld.z) + M[12 + r];
        eN[r] = (std::fabs(M[r]) * eOld.x + std::fabs(M[4 + r]) * eOld.y) + std::fabs(M[8 + r]) * eOld.z;
    }
    for (int r = 0; r < 3; r++) { olo[r] = cN[r] - eN[r]; ohi[r] = cN[r] + eN[r]; }
}

// fn(begin, end, part) over [0, n) in contiguous parts, one thread each: the per-triangle loops of big meshes (every
// part writes — and so first touches — its own pages).  Small n: one call on this thread.
template <class F>
void forParts(int64_t n, F fn, int64_t threadedFrom = 32768) {
    unsigned hw = std::thread::hardware_concurrency();
    const int64_t parts = n < threadedFrom ? 1 : (int64_t)(hw < 1 ? 1 : (hw > 16 ? 16 : hw));
    if (parts == 1) { fn((int64_t)0, n, 0); return; }
    std::vector<std::thread> th;
    int64_t started = 1;
    try {
        for (; started < parts; started++) th.emplace_back(fn, n * started / parts, n * (started + 1) / parts, (int)started);
    } catch (const std::system_error&) {}          // no more threads to be had: the parts not started run here
    fn((int64_t)0, n / parts, 0);
    for (int64_t p = started; p < parts; p++) fn(n * p / parts, n * (p + 1) / parts, (int)p);
    for (std::thread& t : th) t.join();
}
const int kMaxParts = 16;

// first touch of freshly allocated (unwritten) elements [from, to) of a default-initialised vector, on all threads
void touchPages(double* base, size_t from, size_t to) {
#ifdef __linux__
    {   // 2 MB pages where the kernel hands them out on request (transparent_hugepage = madvise): 512 times fewer faults
        const uintptr_t huge = (uintptr_t)2 << 20;
        const uintptr_t lo = ((uintptr_t)(base + from) + huge - 1) & ~(huge - 1), hi = (uintptr_t)(base + to) & ~(huge - 1);
        if (hi > lo) madvise((void*)lo, hi - lo, MADV_HUGEPAGE);
    }
#endif
    const size_t step = 4096 / sizeof(double);
    forParts((int64_t)((to - from + step - 1) / step), [&](int64_t b, int64_t e, int) {
        for (int64_t k = b; k < e; k++) base[from + (size_t)k * step] = 0.0;
    }, 1024);
}

// localToWorld -> worldToLocal, normalMatrix, worldBounds of an instance (makeInstance :437-456, refitTLAS :889-912)
void placeInstance(rt_instance& in, const double* m16, const RawDoubles& pb) {
    std::memcpy(in.localToWorld, m16, sizeof in.localToWorld);
    inverse4(m16, in.worldToLocal);
    normalMatrix(m16, in.normalMatrix);
    const double inf = std::numeric_limits<double>::infinity();
    // the union of the transformed prim boxes; min / max are exact, so the parts combine to the one-thread result
    double plo[kMaxParts][3], phi[kMaxParts][3];
    for (int p = 0; p < kMaxParts; p++) for (int a = 0; a < 3; a++) { plo[p][a] = inf; phi[p][a] = -inf; }
    forParts((int64_t)(pb.size() / 6), [&](int64_t b, int64_t e, int part) {
        double lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
        for (int64_t t = b; t < e; t++) {
            double tl[3], th[3];
            transformBox(&pb[(size_t)t * 6], &pb[(size_t)t * 6 + 3], m16, tl, th);
            for (int a = 0; a < 3; a++) { lo[a] = mn(lo[a], tl[a]); hi[a] = mx(hi[a], th[a]); }
        }
        for (int a = 0; a < 3; a++) { plo[part][a] = lo[a]; phi[part][a] = hi[a]; }
    });
    double lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
    for (int p = 0; p < kMaxParts; p++) for (int a = 0; a < 3; a++) { lo[a] = mn(lo[a], plo[p][a]); hi[a] = mx(hi[a], phi[p][a]); }
    for (int a = 0; a < 3; a++) { in.worldBounds[a] = lo[a]; in.worldBounds[3 + a] = hi[a]; }
}

}  // namespace

void RTContext::setInstanceTransform(size_t i, const double* m16, bool boundsOnDevice) {
    rt_instance& in = instances.at(i);
    const RawDoubles& pb = blasPrimBounds.at((size_t)in.blas);
    if (boundsOnDevice && pb.size() > 6) {
        // a mesh: the union over its triangles is left to rt_scene_update_instances (inverted box = "derive"),
        // which computes the same values on the device and writes them back here
        static const RawDoubles none;
        placeInstance(in, m16, none);
    } else {
        placeInstance(in, m16, pb);
    }
}

// RTCUDA_TRACE=1: phase times of engine creation on stderr (diagnostics)
static double traceNow() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static bool traceOn() { static const bool on = std::getenv("RTCUDA_TRACE") != nullptr; return on; }
static void trace(const char* what, double t0) { if (traceOn()) std::fprintf(stderr, "[rtcuda trace] %-32s %8.2f ms\n", what, traceNow() - t0); }

RTContext::RTContext(const pk::Scene& scene, int ingestDevice) {
    const double tCtx = traceNow();
    intersectionTestEpsilon = scene.intersectionTestEpsilon;
    shadowRayEpsilon = scene.shadowRayEpsilon;
    for (const pk::Material& m : scene.materials) {
        rt_material q{};
        q.type = m.type;
        put(q.ambient, d3(m.ambient)); put(q.diffuse, d3(m.diffuse)); put(q.specular, d3(m.specular));
        put(q.mirror, d3(m.mirror)); put(q.absorption, d3(m.absorption));
        q.phong = m.phong; q.ior = m.ior; q.absorptionIndex = m.absorptionIndex; q.roughness = m.roughness;
        materials.push_back(q);
    }
    for (const pk::PointLight& l : scene.pointLights) { rt_point_light q{}; put(q.position, d3(l.position)); put(q.intensity, d3(l.intensity)); pointLights.push_back(q); }
    for (const pk::AreaLight& l : scene.areaLights) { rt_area_light q{}; put(q.position, d3(l.position)); put(q.normal, d3(l.normal)); q.size = l.size; put(q.radiance, d3(l.radiance)); areaLights.push_back(q); }

    struct Entry { int32_t blas; bool hasMat; int64_t material; pk::M4 transform; };
    std::map<int32_t, Entry> byId;
    std::vector<int32_t> meshOrder;
    std::vector<const pk::SceneObject*> meshInstances;
    std::vector<pk::SceneObject> inherited;
    inherited.reserve(scene.objects.size());

    auto vtx = [&](int32_t idx1) -> D3 {
        const size_t k = (size_t)(idx1 - 1);
        if (k >= scene.vertexData.size()) { warnings.push_back("vertex index out of range"); return D3{0, 0, 0}; }
        return d3(scene.vertexData[k]);
    };

    auto addInstance = [&](int32_t blasIdx, const pk::M4& M, bool hasOverride, int64_t matOverride, D3 motion, int32_t objectId) {
        rt_instance in{};
        in.blas = blasIdx;
        in.hasMaterialOverride = hasOverride ? 1 : 0;
        in.materialOverride = matOverride;
        placeInstance(in, M.m, blasPrimBounds[blasIdx]);
        put(in.instanceMotion, motion);
        if (motion.x != 0 || motion.y != 0 || motion.z != 0) anyMotionBlur = true;
        instances.push_back(in);
        instanceObjectIds.push_back(objectId);
    };
    auto newBlas = [&](int64_t firstPrim, int64_t count) -> int32_t {
        blas.push_back(rt_blas{firstPrim, count});
        blasPrimBounds.emplace_back();
        return (int32_t)blas.size() - 1;
    };
    auto pushTriangle = [&](D3 v0, D3 v1, D3 v2, D3 n0, D3 n1, D3 n2, D3 mb) {
        const size_t k = triVertices.size();
        triVertices.resize(k + 9); triNormals.resize(k + 9);
        put(&triVertices[k], v0); put(&triVertices[k + 3], v1); put(&triVertices[k + 6], v2);
        put(&triNormals[k], n0); put(&triNormals[k + 3], n1); put(&triNormals[k + 6], n2);
        triMotionBlur.push_back(mb.x); triMotionBlur.push_back(mb.y); triMotionBlur.push_back(mb.z);
        if (mb.x != 0 || mb.y != 0 || mb.z != 0) anyMotionBlur = true;
    };

    for (const pk::SceneObject& object : scene.objects) {
        const int64_t matID = object.hasMaterial ? object.material : -1;    // materialIndex(for:in:) :423-426
        if (object.kind == pk::OBJ_SPHERE) {                                 // :122-160
            const D3 c = vtx(object.centerIdx);
            rt_sphere s{}; put(s.center, c); s.radius = object.radius;
            spheres.push_back(s);
            prims.push_back(rt_prim{RT_GEO_SPHERE, 1, (int64_t)spheres.size() - 1, matID});
            const int32_t b = newBlas((int64_t)prims.size() - 1, 1);
            blasPrimBounds[b] = {c.x - s.radius, c.y - s.radius, c.z - s.radius, c.x + s.radius, c.y + s.radius, c.z + s.radius};
            addInstance(b, object.transform, true, matID, D3{0, 0, 0}, object.id);
        } else if (object.kind == pk::OBJ_PLANE) {                           // :161-192
            rt_plane p{}; put(p.center, vtx(object.centerIdx)); put(p.normal, d3(object.normal));
            planes.push_back(p);
            prims.push_back(rt_prim{RT_GEO_PLANE, 1, (int64_t)planes.size() - 1, matID});
            const int32_t b = newBlas((int64_t)prims.size() - 1, 1);
            blasPrimBounds[b] = {-1e5, -1e5, -1e5, 1e5, 1e5, 1e5};
            addInstance(b, object.transform, true, matID, D3{0, 0, 0}, object.id);
        } else if (object.kind == pk::OBJ_TRIANGLE) {                        // :193-235
            const D3 v0 = vtx(object.indices[0]), v1 = vtx(object.indices[1]), v2 = vtx(object.indices[2]);
            const D3 n = normalize(cross(v1 - v0, v2 - v0));
            pushTriangle(v0, v1, v2, n, n, n, D3{0, 0, 0});
            prims.push_back(rt_prim{RT_GEO_TRIANGLE, 0, triangleCount() - 1, matID});
            const int32_t b = newBlas((int64_t)prims.size() - 1, 1);
            blasPrimBounds[b] = {mn(v0.x, mn(v1.x, v2.x)), mn(v0.y, mn(v1.y, v2.y)), mn(v0.z, mn(v1.z, v2.z)),
                                 mx(v0.x, mx(v1.x, v2.x)), mx(v0.y, mx(v1.y, v2.y)), mx(v0.z, mx(v1.z, v2.z))};
            addInstance(b, object.transform, true, matID, D3{0, 0, 0}, object.id);
        } else if (object.kind == pk::OBJ_MESH_INSTANCE) {                   // :236-241
            if (!object.hasMaterial) {
                pk::SceneObject mI = object;
                for (const pk::SceneObject& o : scene.objects)
                    if (o.id == mI.baseMeshID) { mI.hasMaterial = o.hasMaterial; mI.material = o.material; break; }
                inherited.push_back(mI);
                meshInstances.push_back(&inherited.back());
            } else {
                meshInstances.push_back(&object);
            }
        } else if (object.kind == pk::OBJ_MESH) {                            // :242-377
            const int64_t start = triangleCount();
            const bool isSmooth = object.smooth;
            const D3 mb = d3(object.motionBlur);
            const bool blur = mb.x != 0 || mb.y != 0 || mb.z != 0;
            std::vector<int64_t> faceIdx;
            std::vector<double> posStore;
            const double* pos = nullptr;
            size_t nPos = 0;
            std::vector<double> normals;
            const bool fromPLY = !object.plyPath.empty();
            // PLY -> device ingest (rt_ply_ingest): the file's raw vertex / face blocks go to the GPU, which returns
            // the per-triangle arrays the loops below would produce (same bits); files the device path does not take
            // (ascii, big endian, mixed polygons ...) are read and flattened here
            int64_t devTris = 0;
            bool fromDevice = false;
            if (fromPLY && ingestDevice >= 0) {
                PlyBlocks blocks; std::string why;
                const double tRead = traceNow();
                const bool haveBlocks = loadPlyBlocks(pk::dirOf(scene.path) + object.plyPath, blocks, why);
                trace("flatten: PLY file -> raw blocks", tRead);
                if (haveBlocks) {
                    // the kernels' output lands directly behind the triangles flattened so far
                    const int64_t nT = blocks.layout.faceCount * (blocks.layout.verticesPerFace == 4 ? 2 : 1);
                    const size_t old = triVertices.size();
                    triVertices.resize(old + (size_t)nT * 9); triNormals.resize(old + (size_t)nT * 9);      // unwritten (RawDoubles)
                    touchPages(triVertices.data(), old, triVertices.size()); touchPages(triNormals.data(), old, triNormals.size());
                    rt_ply_stats ps{};
                    trace("flatten: output arrays resized", tRead);
                    const double tIng = traceNow();
                    const int rc = rt_ply_ingest(&blocks.layout, blocks.vertexBlock.data(), blocks.faceBlock.data(), isSmooth ? 1 : 0,
                                                 ingestDevice, triVertices.data() + old, triNormals.data() + old, &ps);
                    trace("flatten: rt_ply_ingest", tIng);
                    if (rc == RT_OK) { fromDevice = true; devTris = nT; deviceIngestedTriangles += nT; deviceIngestMs += ps.deviceMs; }
                    else {
                        triVertices.resize(old); triNormals.resize(old);
                        if (std::string(rt_last_error()).find("index out of range") != std::string::npos) { warnings.push_back("mesh face index out of range"); continue; }
                        warnings.push_back(std::string("device ingest refused ") + object.plyPath + ": " + rt_last_error());
                    }
                }
            }
            if (fromDevice) {
                // nothing to load on the host
            } else if (fromPLY) {
                PlyMesh pm; std::string err;
                if (!loadPly(pk::dirOf(scene.path) + object.plyPath, pm, err)) { warnings.push_back("PLY load failed: " + object.plyPath + " (" + err + ")"); continue; }
                faceIdx = std::move(pm.indices); posStore = std::move(pm.positions); normals = std::move(pm.normals);
                pos = posStore.data(); nPos = posStore.size() / 3;
            } else {
                faceIdx.assign(object.faces.begin(), object.faces.end());
                pos = scene.vertexData.empty() ? nullptr : &scene.vertexData[0].x;   // pk::V3 is 3 packed doubles
                nPos = scene.vertexData.size();
            }
            auto P = [&](int64_t i) -> D3 { return D3{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]}; };
            const int64_t triCount = fromDevice ? devTris : (int64_t)faceIdx.size() / 3;
            const int64_t offset = (!normals.empty() || fromPLY) ? 0 : 1;    // :300-301 (with normals: used as-is, :270-272)
            bool bad = false;
            for (int64_t k = 0; k < triCount * 3 && !fromDevice; k++)
                if (faceIdx[k] - offset < 0 || (size_t)(faceIdx[k] - offset) >= nPos) bad = true;
            if (bad) { warnings.push_back("mesh face index out of range"); continue; }
            std::vector<D3> vtxNormals;
            std::vector<D3> faceNormals;
            if (normals.empty() && !fromDevice) {
                faceNormals.resize((size_t)triCount);
                for (int64_t t = 0; t < triCount; t++) {                     // :307-314
                    const D3 v0 = P(faceIdx[3 * t] - offset), v1 = P(faceIdx[3 * t + 1] - offset), v2 = P(faceIdx[3 * t + 2] - offset);
                    faceNormals[t] = normalize(cross(v1 - v0, v2 - v0));
                }
                if (isSmooth) {                                              // :316-326 unweighted unit-normal sum
                    vtxNormals.assign(nPos, D3{0, 0, 0});
                    for (int64_t t = 0; t < triCount; t++)
                        for (int c = 0; c < 3; c++) { D3& a = vtxNormals[faceIdx[3 * t + c] - offset]; a = a + faceNormals[t]; }
                    for (D3& a : vtxNormals) a = normalize(a);
                }
            }
            const int64_t firstPrim = (int64_t)prims.size();
            RawDoubles pb;
            if (!fromDevice) pb.reserve((size_t)triCount * 6);
            if (fromDevice) {
                // vertices and normals are in place; motion vectors, prim records and prim boxes in bulk
                const int64_t first = triangleCount() - triCount;
                triMotionBlur.resize((size_t)(first + triCount) * 3);
                prims.resize((size_t)firstPrim + (size_t)triCount);
                pb.resize((size_t)triCount * 6);
                if (blur) anyMotionBlur = true;
                forParts(triCount, [&](int64_t tb, int64_t te, int) {
                    for (int64_t t = tb; t < te; t++) {
                        double* m = &triMotionBlur[(size_t)(first + t) * 3];
                        m[0] = mb.x; m[1] = mb.y; m[2] = mb.z;
                        prims[(size_t)(firstPrim + t)] = rt_prim{RT_GEO_TRIANGLE, isSmooth ? 1 : 0, first + t, matID};
                        const double* v = &triVertices[(size_t)(first + t) * 9];
                        double* b = &pb[(size_t)t * 6];
                        for (int a = 0; a < 3; a++) {
                            double lo = mn(v[a], mn(v[3 + a], v[6 + a])), hi = mx(v[a], mx(v[3 + a], v[6 + a]));
                            if (blur) { const double m3 = a == 0 ? mb.x : (a == 1 ? mb.y : mb.z); lo = mn(lo, lo + m3); hi = mx(hi, hi + m3); }
                            b[a] = lo; b[3 + a] = hi;
                        }
                    }
                });
            }
            for (int64_t t = 0; t < triCount && !fromDevice; t++) {
                const int64_t i0 = faceIdx[3 * t] - offset, i1 = faceIdx[3 * t + 1] - offset, i2 = faceIdx[3 * t + 2] - offset;
                const D3 v0 = P(i0), v1 = P(i1), v2 = P(i2);
                D3 n0, n1, n2;
                if (!normals.empty()) {
                    n0 = D3{normals[3 * i0], normals[3 * i0 + 1], normals[3 * i0 + 2]};
                    n1 = D3{normals[3 * i1], normals[3 * i1 + 1], normals[3 * i1 + 2]};
                    n2 = D3{normals[3 * i2], normals[3 * i2 + 1], normals[3 * i2 + 2]};
                } else if (isSmooth) {
                    n0 = vtxNormals[i0]; n1 = vtxNormals[i1]; n2 = vtxNormals[i2];
                } else {
                    n0 = n1 = n2 = normalize(faceNormals[t]);                // :343 (normalised twice)
                }
                pushTriangle(v0, v1, v2, n0, n1, n2, mb);
                prims.push_back(rt_prim{RT_GEO_TRIANGLE, isSmooth ? 1 : 0, triangleCount() - 1, matID});
                double lo[3] = {mn(v0.x, mn(v1.x, v2.x)), mn(v0.y, mn(v1.y, v2.y)), mn(v0.z, mn(v1.z, v2.z))};
                double hi[3] = {mx(v0.x, mx(v1.x, v2.x)), mx(v0.y, mx(v1.y, v2.y)), mx(v0.z, mx(v1.z, v2.z))};
                if (blur) {                                                  // :283-289 / :348-354
                    const double m3[3] = {mb.x, mb.y, mb.z};
                    for (int a = 0; a < 3; a++) { lo[a] = mn(lo[a], lo[a] + m3[a]); hi[a] = mx(hi[a], hi[a] + m3[a]); }
                }
                pb.insert(pb.end(), {lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]});
            }
            meshOrder.push_back(object.id);
            if (triangleCount() > start) {                                   // :372-376
                const int32_t b = newBlas(firstPrim, triCount);
                blasPrimBounds[b] = std::move(pb);
                byId[object.id] = Entry{b, true, matID, object.transform};
            }
        }
    }

    for (const pk::SceneObject* io : meshInstances) {                        // :384-401
        auto it = byId.find(io->baseMeshID);
        if (it == byId.end()) continue;
        const Entry base = it->second;
        const int64_t matOverride = io->hasMaterial ? io->material : -1;
        addInstance(base.blas, io->transform, true, matOverride, d3(io->motionBlur), io->id);
        byId[io->id] = Entry{base.blas, true, matOverride, io->transform};
    }
    for (int32_t meshID : meshOrder) {                                       // :403-410 (object order)
        auto it = byId.find(meshID);
        if (it == byId.end()) continue;
        addInstance(it->second.blas, it->second.transform, true, it->second.material, D3{0, 0, 0}, meshID);
    }
    if (instances.empty()) warnings.push_back("No instances to build TLAS");  // :416
    trace("flatten: RTContext(scene:) total", tCtx);
}

rt_scene_desc RTContext::describe(const pk::Scene& scene) const {
    rt_scene_desc d{};
    d.abiVersion = RT_ABI_VERSION;
    d.nTriangles = triangleCount();
    d.triVertices = triVertices.data(); d.triNormals = triNormals.data();
    d.triMotionBlur = anyMotionBlur ? triMotionBlur.data() : nullptr;
    d.nSpheres = (int64_t)spheres.size(); d.spheres = spheres.data();
    d.nPlanes = (int64_t)planes.size(); d.planes = planes.data();
    d.nPrims = (int64_t)prims.size(); d.prims = prims.data();
    d.nBlas = (int64_t)blas.size(); d.blas = blas.data();
    d.nInstances = (int64_t)instances.size(); d.instances = instances.data();
    d.nMaterials = (int64_t)materials.size(); d.materials = materials.data();
    d.nPointLights = (int64_t)pointLights.size(); d.pointLights = pointLights.data();
    d.nAreaLights = (int64_t)areaLights.size(); d.areaLights = areaLights.data();
    d.ambient[0] = scene.ambient.x; d.ambient[1] = scene.ambient.y; d.ambient[2] = scene.ambient.z;
    d.background[0] = scene.backgroundColor.x; d.background[1] = scene.backgroundColor.y; d.background[2] = scene.backgroundColor.z;
    d.intersectionTestEpsilon = intersectionTestEpsilon; d.shadowRayEpsilon = shadowRayEpsilon;
    d.maxRecursionDepth = scene.maxRecursionDepth;
    return d;
}

}  // namespace rthost
