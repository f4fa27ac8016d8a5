// oracle/rtcontext_init.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// RTContext.init(scene:) restated: Models/RTContext.swift:94-418.
#pragma once
#include "rtcontext.hpp"

namespace orc {

inline void RTContext::init(const pk::Scene& scene) {
    intersectionTestEpsilon = scene.intersectionTestEpsilon;  // :95
    shadowRayEpsilon = scene.shadowRayEpsilon;                // :96
    for (const pk::Material& m : scene.materials) {           // :97
        Material q;
        q.type = m.type;
        q.ambient = toV(m.ambient); q.diffuse = toV(m.diffuse); q.specular = toV(m.specular);
        q.mirror = toV(m.mirror); q.absorption = toV(m.absorption);
        q.phong = m.phong; q.ior = m.ior; q.absorptionIndex = m.absorptionIndex; q.roughness = m.roughness;
        materials.push_back(q);
    }
    for (const pk::PointLight& l : scene.pointLights) pointLights.push_back({toV(l.position), toV(l.intensity)});  // :103
    for (const pk::AreaLight& l : scene.areaLights)                                                                  // :104
        areaLights.push_back({toV(l.position), toV(l.normal), l.size, toV(l.radiance)});

    int64_t currentSphereIndex = 0, currentPlaneIndex = 0, currentTriangleIndex = 0;  // :106-108

    struct Entry { std::shared_ptr<const BLAS> blas; bool hasMat; int64_t material; Mat4 transform; };
    std::map<int32_t, Entry> instanceByID;          // :116
    std::vector<int32_t> meshOrder;                 // stands in for meshTriRangeByID keys (:111), object order
    std::vector<Instance> instArray;                // :117
    std::vector<pk::SceneObject> meshInstances;     // :118

    auto vtx = [&](int32_t idx1) -> Vec3 {
        size_t k = (size_t)(idx1 - 1);
        if (k >= scene.vertexData.size()) { warnings.push_back("vertex index out of range"); return Vec3(); }
        return toV(scene.vertexData[k]);
    };

    for (const pk::SceneObject& object : scene.objects) {                 // :120
        // materialIndex(for:in:) :423-426 — Int(id) or -1
        const int64_t matID = object.hasMaterial ? object.material : -1;  // :121
        const Mat4 M = toM(object.transform);
        if (object.kind == pk::OBJ_SPHERE) {                              // :122-160
            Vec3 localCenter = vtx(object.centerIdx);                     // :124
            Sphere sph{localCenter, object.radius};                       // :137
            spheres.push_back(sph);
            PrimitiveInfo prim;                                           // :141-146
            prim.type = GEO_SPHERE;
            prim.primitiveIndex = currentSphereIndex;
            prim.bounds = AABB(sph.center - Vec3(sph.radius), sph.center + Vec3(sph.radius));
            prim.centroid = sph.center;
            prim.materialID = matID;
            BVHBuilder b(std::vector<PrimitiveInfo>{prim});               // :149
            Instance inst = makeInstance(blasFromBuilder(std::move(b), {prim}), M, true, matID);  // :152-157
            inst.objectId = object.id;
            instArray.push_back(inst);
            currentSphereIndex += 1;
        } else if (object.kind == pk::OBJ_PLANE) {                        // :161-192
            Plane pln{vtx(object.centerIdx), toV(object.normal)};         // :167-169
            planes.push_back(pln);
            const double maxCoord = 1e5;                                  // :172
            PrimitiveInfo prim;
            prim.type = GEO_PLANE;
            prim.primitiveIndex = currentPlaneIndex;
            prim.bounds = AABB(Vec3(-maxCoord), Vec3(maxCoord));
            prim.centroid = Vec3();
            prim.materialID = matID;
            BVHBuilder b(std::vector<PrimitiveInfo>{prim});
            Instance inst = makeInstance(blasFromBuilder(std::move(b), {prim}), M, true, matID);
            inst.objectId = object.id;
            instArray.push_back(inst);
            currentPlaneIndex += 1;
        } else if (object.kind == pk::OBJ_TRIANGLE) {                     // :193-235
            Vec3 v0 = vtx(object.indices[0]), v1 = vtx(object.indices[1]), v2 = vtx(object.indices[2]);  // :199-201
            Triangle tri;
            tri.v0 = v0; tri.v1 = v1; tri.v2 = v2;
            tri.e1 = v1 - v0; tri.e2 = v2 - v0;                           // :205
            tri.centroid = (v0 + v1 + v2) / 3.0;                          // :206
            Vec3 n = normalize(cross(tri.e1, tri.e2));                    // :207
            tri.n0 = n; tri.n1 = n; tri.n2 = n;
            tri.motionBlur = Vec3();
            triangles.push_back(tri);
            PrimitiveInfo prim;                                           // :216-221
            prim.type = GEO_TRIANGLE;
            prim.primitiveIndex = currentTriangleIndex;
            prim.bounds = AABB(vmin(v0, vmin(v1, v2)), vmax(v0, vmax(v1, v2)));  // :212-213
            prim.centroid = tri.centroid;
            prim.materialID = matID;
            prim.shadingMode = SHADE_FLAT;
            BVHBuilder b(std::vector<PrimitiveInfo>{prim});
            Instance inst = makeInstance(blasFromBuilder(std::move(b), {prim}), M, true, matID);
            inst.objectId = object.id;
            instArray.push_back(inst);
            currentTriangleIndex += 1;
        } else if (object.kind == pk::OBJ_MESH_INSTANCE) {                // :236-241
            pk::SceneObject mI = object;
            if (!mI.hasMaterial) {                                        // :237-239 inherit the base object's material
                for (const pk::SceneObject& o : scene.objects)
                    if (o.id == mI.baseMeshID) { mI.hasMaterial = o.hasMaterial; mI.material = o.material; break; }
            }
            meshInstances.push_back(mI);
        } else if (object.kind == pk::OBJ_MESH) {                         // :242-377
            const size_t start = triangles.size();
            const bool isSmooth = object.smooth;                          // :245
            const Vec3 mb = toV(object.motionBlur);
            const bool hasBlur = mb != Vec3();
            std::vector<int64_t> faceIndices;
            std::vector<Vec3> meshPositions, meshNormals;
            const bool fromPLY = !object.plyPath.empty();
            if (fromPLY) {                                                // :251-261
                PlyMesh pm;
                bool viaRef = false;
                if (!loadPly(pk::dirOf(scene.path) + object.plyPath, pm, &viaRef)) {
                    warnings.push_back("PLY load failed: " + object.plyPath);
                    continue;                                             // :260
                }
                plyViaReference = viaRef;
                faceIndices = std::move(pm.indices);
                meshPositions = std::move(pm.positions);
                meshNormals = std::move(pm.normals);
            } else {                                                      // :262-265
                faceIndices.assign(object.faces.begin(), object.faces.end());
                meshPositions.reserve(scene.vertexData.size());
                for (const pk::V3& v : scene.vertexData) meshPositions.push_back(toV(v));
            }
            auto pushPrim = [&](const Triangle& t, Vec3 v0, Vec3 v1, Vec3 v2) {  // :283-296 / :348-361
                Vec3 b0 = hasBlur ? v0 + mb : v0, b1 = hasBlur ? v1 + mb : v1, b2 = hasBlur ? v2 + mb : v2;
                PrimitiveInfo p;
                p.type = GEO_TRIANGLE;
                p.primitiveIndex = currentTriangleIndex;
                p.bounds = AABB(vmin(vmin(v0, vmin(v1, v2)), vmin(b0, vmin(b1, b2))),
                                vmax(vmax(v0, vmax(v1, v2)), vmax(b0, vmax(b1, b2))));
                p.centroid = t.centroid;
                p.materialID = matID;
                p.shadingMode = isSmooth ? SHADE_SMOOTH : SHADE_FLAT;
                bvhPrims.push_back(p);
                currentTriangleIndex += 1;
            };
            const size_t triCount = faceIndices.size() / 3;
            if (!meshNormals.empty()) {                                   // :267-297 (indices used as-is)
                for (size_t i = 0; i < triCount; i++) {
                    int64_t i0 = faceIndices[3 * i], i1 = faceIndices[3 * i + 1], i2 = faceIndices[3 * i + 2];
                    Vec3 v0 = meshPositions[i0], v1 = meshPositions[i1], v2 = meshPositions[i2];
                    Triangle t;
                    t.v0 = v0; t.v1 = v1; t.v2 = v2;
                    t.e1 = v1 - v0; t.e2 = v2 - v0;
                    t.centroid = (v0 + v1 + v2) / 3.0;
                    t.n0 = meshNormals[i0]; t.n1 = meshNormals[i1]; t.n2 = meshNormals[i2];
                    t.motionBlur = mb;
                    triangles.push_back(t);
                    pushPrim(t, v0, v1, v2);
                }
            } else {                                                      // :298-363
                const int64_t offset = fromPLY ? 0 : 1;                   // :300-301
                std::vector<Vec3> faceNormals(triCount);
                for (size_t t = 0; t < triCount; t++) {                   // :307-314
                    int64_t i0 = faceIndices[3 * t] - offset, i1 = faceIndices[3 * t + 1] - offset,
                            i2 = faceIndices[3 * t + 2] - offset;
                    Vec3 v0 = meshPositions[i0], v1 = meshPositions[i1], v2 = meshPositions[i2];
                    faceNormals[t] = normalize(cross(v1 - v0, v2 - v0));
                }
                std::vector<Vec3> vtxNormals(meshPositions.size());
                if (isSmooth) {                                           // :317-326 unweighted sum of unit normals
                    for (size_t t = 0; t < triCount; t++) {
                        Vec3 fn = faceNormals[t];
                        vtxNormals[faceIndices[3 * t] - offset] += fn;
                        vtxNormals[faceIndices[3 * t + 1] - offset] += fn;
                        vtxNormals[faceIndices[3 * t + 2] - offset] += fn;
                    }
                    for (Vec3& n : vtxNormals) n = normalize(n);
                }
                for (size_t t = 0; t < triCount; t++) {                   // :328-362
                    int64_t i0 = faceIndices[3 * t] - offset, i1 = faceIndices[3 * t + 1] - offset,
                            i2 = faceIndices[3 * t + 2] - offset;
                    Vec3 v0 = meshPositions[i0], v1 = meshPositions[i1], v2 = meshPositions[i2];
                    Triangle tri;
                    tri.v0 = v0; tri.v1 = v1; tri.v2 = v2;
                    tri.e1 = v1 - v0; tri.e2 = v2 - v0;
                    tri.centroid = (v0 + v1 + v2) / 3.0;
                    tri.motionBlur = mb;
                    if (isSmooth) {
                        tri.n0 = vtxNormals[i0]; tri.n1 = vtxNormals[i1]; tri.n2 = vtxNormals[i2];
                    } else {
                        Vec3 n = normalize(faceNormals[t]);               // :343 (normalised twice)
                        tri.n0 = n; tri.n1 = n; tri.n2 = n;
                    }
                    triangles.push_back(tri);
                    pushPrim(tri, v0, v1, v2);
                }
            }
            const size_t end = triangles.size();                          // :365
            meshOrder.push_back(object.id);                               // :367-368
            if (end > start) {                                            // :372-376
                // buildBLASForMesh :430-435 — the subset is this mesh's contiguous slice of bvhPrims
                std::vector<PrimitiveInfo> subset(bvhPrims.begin() + (bvhPrims.size() - (end - start)), bvhPrims.end());
                BVHBuilder b(subset, 2, 12);
                instanceByID[object.id] = Entry{blasFromBuilder(std::move(b), std::move(subset)), true, matID, M};
            }
        }
    }

    for (const pk::SceneObject& instObj : meshInstances) {                // :384-401
        auto it = instanceByID.find(instObj.baseMeshID);
        if (it == instanceByID.end()) continue;
        Entry baseData = it->second;
        const Mat4 instanceTransform = toM(instObj.transform);            // composeTransform(..., base:) pre-composed
        const int64_t matOverride = instObj.hasMaterial ? instObj.material : -1;  // :392
        Instance inst = makeInstance(baseData.blas, instanceTransform, true, matOverride, toV(instObj.motionBlur));
        inst.objectId = instObj.id;
        instArray.push_back(inst);
        instanceByID[instObj.id] = Entry{baseData.blas, true, matOverride, instanceTransform};  // :400
    }

    for (int32_t meshID : meshOrder) {                                    // :403-410
        auto it = instanceByID.find(meshID);
        if (it == instanceByID.end()) continue;
        Instance inst = makeInstance(it->second.blas, it->second.transform, true, it->second.material);
        inst.objectId = meshID;
        instArray.push_back(inst);
    }

    if (!instArray.empty()) {                                             // :412-417
        tlas = buildTLAS(instArray);
        instances = instArray;
        hasTlas = true;
    } else {
        warnings.push_back("No instances to build TLAS");
    }
}

}  // namespace orc
