// scene_model.hpp — stand-in for the reference's absent ParsingKit dependency.
//
// The reference takes its scene model (Scene, Camera, Material, PointLight,
// AreaLight, Sphere, Plane, Triangle, Mesh, MeshInstance, VertexData) from the
// SwiftPM package ParsingKit @ c2027f72 (Package.resolved:4-12), which is not
// vendored in /root/reference.  Only the fields the render path reads are
// known, from their call sites:
//   Scene      RTContext.swift:95-104,120-127,199-201,252-265,374,388-390
//              Object+Extension.swift:58,88,97,102,116
//   Camera     Object+Extension.swift:58-63,385-425
//   Material   Object+Extension.swift:105-106,112-116,129-131,189-274
//   lights     Object+Extension.swift:118-186
//   objects    RTContext.swift:120-378
// This header defines exactly that field set as plain C++ structs plus a
// binary ".rtscene" container so that seeded synthetic scenes can be written
// by Python and read by both the host-side mirror and the CPU oracle.
// `Scene.composeTransform` (ParsingKit, absent) is bypassed: every object
// carries its already-composed local-to-world matrix.
//
// Header-only, no dependencies beyond the C++17 standard library.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace pk {

struct V3 {
    double x = 0, y = 0, z = 0;
};

// Column-major 4x4, like simd_double4x4: m[c*4 + r].
struct M4 {
    double m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
};

struct Camera {
    int32_t id = 0;
    std::string imageName;
    int32_t width = 0, height = 0;  // imageResolution.(0,1)
    int32_t numSamples = 1;
    double apertureSize = 0, focusDistance = 0;
    V3 position, gaze, gazePoint, up;
    double nearPlane[4] = {0, 0, 0, 0};  // l, r, b, t
    double nearDistance = 1;
    bool hasFovy = false;
    double fovy = 0;
    bool isLookAt = false;  // cam.type?.lowercased() == "lookat"
};

// Material.type is a String in ParsingKit; the render loop compares it with
// "mirror", "dielectric", "conductor" and treats anything else as plain
// Blinn-Phong (Object+Extension.swift:189,207,252).
enum MaterialType : int32_t { MAT_PLAIN = 0, MAT_MIRROR = 1, MAT_DIELECTRIC = 2, MAT_CONDUCTOR = 3 };

struct Material {
    int32_t id = 0;
    int32_t type = MAT_PLAIN;
    V3 ambient, diffuse, specular, mirror, absorption;
    double phong = 1, ior = 0, absorptionIndex = 0, roughness = 0;
};

struct PointLight {
    V3 position, intensity;
};

struct AreaLight {
    V3 position, normal;
    double size = 0;
    V3 radiance;
};

enum ObjectKind : int32_t { OBJ_SPHERE = 0, OBJ_PLANE = 1, OBJ_TRIANGLE = 2, OBJ_MESH = 3, OBJ_MESH_INSTANCE = 4 };

struct SceneObject {
    int32_t kind = OBJ_SPHERE;
    int32_t id = 0;  // numeric object id ("" in the reference when missing)
    // `material: String?` parsed with Int(id) (RTContext.swift:423-426):
    // hasMaterial == false stands for nil / "" / unparseable -> index -1
    // (for a MeshInstance: inherit the base object's, RTContext.swift:237-239).
    bool hasMaterial = false;
    int32_t material = -1;
    M4 transform;  // composeTransform(tokens:reset:[base:]) result
    // sphere / plane
    int32_t centerIdx = 0;  // 1-based into vertexData
    double radius = 0;
    V3 normal;
    // lone triangle
    int32_t indices[3] = {0, 0, 0};  // 1-based into vertexData
    // mesh
    bool smooth = false;  // shadingMode == "smooth"
    V3 motionBlur;        // Mesh.motionBlur / MeshInstance.motionBlur
    std::string plyPath;  // faces.plyPath (relative to the scene file), "" = none
    std::vector<int32_t> faces;  // faces.data, 1-based into vertexData
    // mesh instance
    int32_t baseMeshID = 0;
};

struct Scene {
    double intersectionTestEpsilon = 1e-6;
    double shadowRayEpsilon = 1e-3;
    int32_t maxRecursionDepth = 0;
    V3 backgroundColor, ambient;
    std::vector<Camera> cameras;
    std::vector<Material> materials;
    std::vector<PointLight> pointLights;
    std::vector<AreaLight> areaLights;
    std::vector<V3> vertexData;
    std::vector<SceneObject> objects;
    std::string path;  // scene file path (PLY paths resolve against its directory)
};

// ---------------------------------------------------------------------------
// .rtscene reader.  Layout (little endian, packed, written by
// myraytracer_b200/scenegen.py::write_rtscene):
//   char[8] "RTSCENE1"
//   f64 intersectionTestEpsilon, shadowRayEpsilon; i32 maxRecursionDepth
//   f64[3] background; f64[3] ambient
//   u32 nCameras  { i32 id; str imageName; i32 w,h,numSamples; f64 aperture,focus;
//                   f64[3] position,gaze,gazePoint,up; f64[4] nearPlane; f64 nearDistance;
//                   i32 hasFovy; f64 fovy; i32 isLookAt }
//   u32 nMaterials{ i32 id,type; f64[3] ambient,diffuse,specular,mirror,absorption;
//                   f64 phong,ior,absorptionIndex,roughness }
//   u32 nPoint    { f64[3] position,intensity }
//   u32 nArea     { f64[3] position,normal; f64 size; f64[3] radiance }
//   u32 nVertices ; f64[3*n]
//   u32 nObjects  { i32 kind,id,hasMaterial,material; f64[16] transform;
//                   i32 centerIdx; f64 radius; f64[3] normal; i32[3] indices;
//                   i32 smooth; f64[3] motionBlur; str plyPath; u32 nFaces; i32[nFaces];
//                   i32 baseMeshID }
//   str = u32 length + bytes
// ---------------------------------------------------------------------------
namespace detail {
struct Reader {
    FILE* f = nullptr;
    bool ok = true;
    explicit Reader(const char* p) { f = std::fopen(p, "rb"); ok = (f != nullptr); }
    ~Reader() { if (f) std::fclose(f); }
    void raw(void* d, size_t n) {
        if (!ok) return;
        if (n && std::fread(d, 1, n, f) != n) ok = false;
    }
    int32_t i32() { int32_t v = 0; raw(&v, 4); return v; }
    uint32_t u32() { uint32_t v = 0; raw(&v, 4); return v; }
    double f64() { double v = 0; raw(&v, 8); return v; }
    V3 v3() { V3 v; v.x = f64(); v.y = f64(); v.z = f64(); return v; }
    std::string str() {
        uint32_t n = u32();
        if (!ok || n > (1u << 20)) { ok = false; return {}; }
        std::string s(n, '\0');
        raw(s.data(), n);
        return s;
    }
};
}  // namespace detail

inline bool loadSceneBlob(const char* path, Scene& sc, std::string& err) {
    detail::Reader r(path);
    if (!r.ok) { err = std::string("cannot open ") + path; return false; }
    char magic[8];
    r.raw(magic, 8);
    if (!r.ok || std::memcmp(magic, "RTSCENE1", 8) != 0) { err = "bad magic"; return false; }
    sc = Scene();
    sc.path = path;
    sc.intersectionTestEpsilon = r.f64();
    sc.shadowRayEpsilon = r.f64();
    sc.maxRecursionDepth = r.i32();
    sc.backgroundColor = r.v3();
    sc.ambient = r.v3();
    const uint32_t kMax = 1u << 30;
    uint32_t nc = r.u32();
    if (!r.ok || nc > kMax) { err = "truncated (cameras)"; return false; }
    for (uint32_t i = 0; i < nc && r.ok; i++) {
        Camera c;
        c.id = r.i32();
        c.imageName = r.str();
        c.width = r.i32(); c.height = r.i32(); c.numSamples = r.i32();
        c.apertureSize = r.f64(); c.focusDistance = r.f64();
        c.position = r.v3(); c.gaze = r.v3(); c.gazePoint = r.v3(); c.up = r.v3();
        for (double& v : c.nearPlane) v = r.f64();
        c.nearDistance = r.f64();
        c.hasFovy = r.i32() != 0;
        c.fovy = r.f64();
        c.isLookAt = r.i32() != 0;
        sc.cameras.push_back(c);
    }
    uint32_t nm = r.u32();
    if (!r.ok || nm > kMax) { err = "truncated (materials)"; return false; }
    for (uint32_t i = 0; i < nm && r.ok; i++) {
        Material m;
        m.id = r.i32(); m.type = r.i32();
        m.ambient = r.v3(); m.diffuse = r.v3(); m.specular = r.v3(); m.mirror = r.v3(); m.absorption = r.v3();
        m.phong = r.f64(); m.ior = r.f64(); m.absorptionIndex = r.f64(); m.roughness = r.f64();
        sc.materials.push_back(m);
    }
    uint32_t np = r.u32();
    if (!r.ok || np > kMax) { err = "truncated (point lights)"; return false; }
    for (uint32_t i = 0; i < np && r.ok; i++) {
        PointLight l; l.position = r.v3(); l.intensity = r.v3();
        sc.pointLights.push_back(l);
    }
    uint32_t na = r.u32();
    if (!r.ok || na > kMax) { err = "truncated (area lights)"; return false; }
    for (uint32_t i = 0; i < na && r.ok; i++) {
        AreaLight l; l.position = r.v3(); l.normal = r.v3(); l.size = r.f64(); l.radiance = r.v3();
        sc.areaLights.push_back(l);
    }
    uint32_t nv = r.u32();
    if (!r.ok || nv > kMax) { err = "truncated (vertices)"; return false; }
    sc.vertexData.resize(nv);
    r.raw(sc.vertexData.data(), size_t(nv) * sizeof(V3));
    uint32_t no = r.u32();
    if (!r.ok || no > kMax) { err = "truncated (objects)"; return false; }
    for (uint32_t i = 0; i < no && r.ok; i++) {
        SceneObject o;
        o.kind = r.i32(); o.id = r.i32();
        o.hasMaterial = r.i32() != 0; o.material = r.i32();
        r.raw(o.transform.m, sizeof(o.transform.m));
        o.centerIdx = r.i32(); o.radius = r.f64(); o.normal = r.v3();
        o.indices[0] = r.i32(); o.indices[1] = r.i32(); o.indices[2] = r.i32();
        o.smooth = r.i32() != 0; o.motionBlur = r.v3();
        o.plyPath = r.str();
        uint32_t nf = r.u32();
        if (!r.ok || nf > kMax) { err = "truncated (faces)"; return false; }
        o.faces.resize(nf);
        r.raw(o.faces.data(), size_t(nf) * 4);
        o.baseMeshID = r.i32();
        sc.objects.push_back(std::move(o));
    }
    if (!r.ok) { err = "truncated scene file"; return false; }
    return true;
}

// Directory part of a path, with trailing '/' ("" when there is none):
// stands in for scene.path?.deletingLastPathComponent() (RTContext.swift:252).
inline std::string dirOf(const std::string& p) {
    size_t k = p.find_last_of('/');
    return k == std::string::npos ? std::string() : p.substr(0, k + 1);
}

}  // namespace pk
