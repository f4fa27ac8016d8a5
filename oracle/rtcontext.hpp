// oracle/rtcontext.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Restates the reference's scene flattening + acceleration-structure set-up:
//   BLAS / Instance / TLAS / RTContext   Models/RTContext.swift:34-92
//   RTContext.init(scene:)               Models/RTContext.swift:94-418
//   materialIndex                        Models/RTContext.swift:423-426
//   buildBLASForMesh / makeInstance / buildTLAS   :430-474
//   normalTransformMatrix                :23-31
//   decomposeUniformScale                :875-880
// Deviations (documented, none changes arithmetic on the render path):
//   * the unused global BVH (RTContext.swift:380-382) is not built;
//   * base-mesh instances are appended in scene-object order instead of Swift
//     Dictionary iteration order (RTContext.swift:403);
//   * object ids are integers (ParsingKit uses optional strings).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../myraytracer_b200/host/scene_model.hpp"
#include "bvh.hpp"
#include "ply.hpp"

namespace orc {

inline Vec3 toV(const pk::V3& v) { return Vec3(v.x, v.y, v.z); }
inline Mat4 toM(const pk::M4& m) { Mat4 r; for (int i = 0; i < 16; i++) r.m[i] = m.m[i]; return r; }

struct Triangle {  // ParsingKit Triangle fields used at RTContext.swift:204-209,275-281,479-510
    Vec3 v0, v1, v2, e1, e2, centroid, n0, n1, n2, motionBlur;
};
struct Sphere { Vec3 center; double radius; };
struct Plane { Vec3 center; Vec3 normal; };

struct Material {
    int type;
    Vec3 ambient, diffuse, specular, mirror, absorption;
    double phong, ior, absorptionIndex, roughness;
};
struct PointLight { Vec3 position, intensity; };
struct AreaLight { Vec3 position, normal; double size; Vec3 radiance; };

struct BLAS {  // RTContext.swift:34-42
    std::vector<BVHNode> nodes;
    std::vector<uint64_t> primIdx;
    std::vector<PrimitiveInfo> prims;
    uint64_t root = 0;
    int buildDepth = 0;
};

struct Instance {  // RTContext.swift:43-61
    std::shared_ptr<const BLAS> blas;
    Mat4 localToWorld, worldToLocal;
    AABB worldBounds;
    bool hasMaterialOverride = false;
    int64_t materialOverride = 0;
    Vec3 instanceMotion;
    Mat3 normalMatrix;
    int32_t objectId = 0;  // oracle extra: id of the scene object this instance came from
};

struct TLAS {  // RTContext.swift:63-71
    BVHBuilder builder;
    std::vector<Instance> instances;
};

// RTContext.swift:23-31
inline Mat3 normalTransformMatrix(const Mat4& M) { return transpose(inverse(upper3(M))); }

// RTContext.swift:875-880
inline double decomposeUniformScale(const Mat4& M) {
    double sx = length(Vec3(M.m[0], M.m[1], M.m[2]));
    double sy = length(Vec3(M.m[4], M.m[5], M.m[6]));
    double sz = length(Vec3(M.m[8], M.m[9], M.m[10]));
    return (sx + sy + sz) / 3.0;
}

// RTContext.swift:437-457
inline Instance makeInstance(std::shared_ptr<const BLAS> base, const Mat4& M, bool hasOverride, int64_t override_,
                             Vec3 instanceMotion = Vec3()) {
    Vec3 minW(kInf), maxW(-kInf);
    for (const PrimitiveInfo& p : base->prims) {
        AABB worldB = p.bounds.transformed(M);
        minW = vmin(minW, worldB.minP);
        maxW = vmax(maxW, worldB.maxP);
    }
    Instance inst;
    inst.blas = std::move(base);
    inst.localToWorld = M;
    inst.worldToLocal = inverse(M);
    inst.worldBounds = AABB(minW, maxW);
    inst.hasMaterialOverride = hasOverride;
    inst.materialOverride = override_;
    inst.instanceMotion = instanceMotion;
    inst.normalMatrix = normalTransformMatrix(M);
    return inst;
}

// RTContext.swift:459-474
inline TLAS buildTLAS(const std::vector<Instance>& instances) {
    std::vector<PrimitiveInfo> tlasPrims;
    tlasPrims.reserve(instances.size());
    for (size_t i = 0; i < instances.size(); i++) {
        const AABB& b = instances[i].worldBounds;
        PrimitiveInfo p;
        p.type = GEO_MESH;
        p.primitiveIndex = (int64_t)i;
        p.bounds = b;
        p.centroid = (b.minP + b.maxP) * 0.5;
        p.materialID = instances[i].hasMaterialOverride ? instances[i].materialOverride : 0;
        p.shadingMode = SHADE_SMOOTH;
        tlasPrims.push_back(p);
    }
    TLAS t;
    t.builder = BVHBuilder(std::move(tlasPrims), 2, 12);
    t.instances = instances;
    return t;
}

inline std::shared_ptr<const BLAS> blasFromBuilder(BVHBuilder&& b, std::vector<PrimitiveInfo> prims) {
    auto r = std::make_shared<BLAS>();
    r->nodes = std::move(b.bvhNode);
    r->primIdx = std::move(b.primitiveIdx);
    r->prims = std::move(prims);
    r->root = b.rootNodeIdx;
    r->buildDepth = b.maxDepthSeen;
    return r;
}

struct RTContext {  // RTContext.swift:74-92
    double intersectionTestEpsilon = 0, shadowRayEpsilon = 0;
    std::vector<Material> materials;
    std::vector<Sphere> spheres;
    std::vector<Plane> planes;
    std::vector<Triangle> triangles;
    std::vector<PointLight> pointLights;
    std::vector<AreaLight> areaLights;
    std::vector<PrimitiveInfo> bvhPrims;
    std::vector<Instance> instances;
    bool hasTlas = false;
    TLAS tlas;
    // oracle extras
    std::vector<std::string> warnings;
    bool plyViaReference = false;

    RTContext() {}
    explicit RTContext(const pk::Scene& scene) { init(scene); }
    void init(const pk::Scene& scene);
};

}  // namespace orc
