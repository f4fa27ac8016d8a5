// oracle/traverse.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Restates the reference's intersection and traversal routines:
//   intersectTriangle / Sphere / Plane   Models/RTContext.swift:479-538
//   intersectBLAS                        :544-610
//   intersectTLAS                        :619-720
//   occludedTLAS / occludedBLAS          :724-829
//   triShadowHit / sphereShadowHit / planeShadowHit   :832-870
// Ray / Hit are ParsingKit types (absent); their defaults are the inferred
// ones of SURVEY.md §2.3: invDir = 1/dir, tMin = 0, tMax = +inf, hit.t = +inf,
// hit.kind = none.  Parity unpinned at that boundary.
// Oracle extras (outputs only): hit.prim / hit.primType / hit.inst, counters.
#pragma once
#include "rtcontext.hpp"

namespace orc {

enum HitKind : int32_t { HIT_NONE = 0, HIT_TRIANGLE = 1, HIT_SPHERE = 2, HIT_PLANE = 3 };

struct Hit {
    double t = kInf;
    Vec3 p, n;
    int32_t kind = HIT_NONE;
    int64_t mat = 0;
    int64_t prim = -1;  // extra: primitiveIndex of the committed primitive
    int64_t inst = -1;  // extra: instance index set by intersectTLAS
};

struct Ray {
    Vec3 origin, dir, invDir;
    double tMin = 0, tMax = kInf, time = 0;
    Hit hit;
    Ray() {}
    Ray(Vec3 o, Vec3 d, double tm = 0) : origin(o), dir(d), invDir(1.0 / d), time(tm) {}
};

struct Counters {
    int64_t raysPrimary = 0, raysShadow = 0, raysSecondary = 0;
    int64_t nodeVisits = 0;   // the reference's `visits` (closest-hit only)
    int64_t boxTests = 0;     // every hitAABB evaluation (closest-hit and any-hit)
    int64_t triTests = 0, sphereTests = 0, planeTests = 0;
    int64_t instEntries = 0;  // world->local ray transforms
    int64_t maxStack = 0;
    void add(const Counters& o) {
        raysPrimary += o.raysPrimary; raysShadow += o.raysShadow; raysSecondary += o.raysSecondary;
        nodeVisits += o.nodeVisits; boxTests += o.boxTests; triTests += o.triTests;
        sphereTests += o.sphereTests; planeTests += o.planeTests; instEntries += o.instEntries;
        if (o.maxStack > maxStack) maxStack = o.maxStack;
    }
};

// RTContext.swift:557-565 (copies at :632-640, :731-739, :791-799)
static inline double hitAABB(const BVHNode& n, const Ray& r, double eps) {
    Vec3 t1 = (n.aabbMin - r.origin) * r.invDir;
    Vec3 t2 = (n.aabbMax - r.origin) * r.invDir;
    Vec3 tminv = vmin(t1, t2);
    Vec3 tmaxv = vmax(t1, t2);
    double tmin = dmax(dmax(tminv.x, tminv.y), tminv.z);
    double tmax = dmin(tmaxv.x, dmin(tmaxv.y, tmaxv.z));
    return (tmax >= dmax(tmin, eps)) ? tmin : kInf;
}

// RTContext.swift:479-510
static inline void intersectTriangle(Ray& ray, const Triangle& tr, bool isSmooth, int64_t matIdx, double eps,
                                     int64_t primIndex) {
    Vec3 offset = tr.motionBlur * ray.time;
    Vec3 origin = ray.origin - offset;
    Vec3 pvec = cross(ray.dir, tr.e2);
    double det = dot(tr.e1, pvec);
    if (std::fabs(det) < eps) return;
    double invDet = 1.0 / det;
    Vec3 tvec = origin - tr.v0;
    double u = dot(tvec, pvec) * invDet;
    if (u < 0.0 || u > 1.0) return;
    Vec3 q = cross(tvec, tr.e1);
    double v = dot(ray.dir, q) * invDet;
    if (v < 0.0 || u + v > 1.0) return;
    double t = dot(tr.e2, q) * invDet;
    if (t <= dmax(eps, ray.tMin) || t >= ray.hit.t) return;
    double w = 1.0 - u - v;
    Vec3 n;
    if (isSmooth) n = normalize(w * tr.n0 + u * tr.n1 + v * tr.n2);
    else n = normalize(cross(tr.e1, tr.e2));
    Vec3 p = tr.v0 + u * tr.e1 + v * tr.e2;
    ray.hit.t = t; ray.hit.p = p; ray.hit.n = n; ray.hit.kind = HIT_TRIANGLE; ray.hit.mat = matIdx;
    ray.hit.prim = primIndex;
}

// RTContext.swift:512-527
static inline void intersectSphere(Ray& ray, const Sphere& s, int64_t matIdx, double eps, int64_t primIndex) {
    Vec3 oc = ray.origin - s.center;
    double a = dot(ray.dir, ray.dir);
    double b = 2.0 * dot(oc, ray.dir);
    double c = dot(oc, oc) - s.radius * s.radius;
    double disc = b * b - 4 * a * c;
    if (disc < 0) return;
    double sd = std::sqrt(disc);
    double t = (-b - sd) / (2 * a);
    if (t < eps) t = (-b + sd) / (2 * a);
    if (t <= dmax(eps, ray.tMin) || t >= ray.hit.t) return;
    Vec3 p = ray.origin + ray.dir * t;
    Vec3 n = normalize(p - s.center);
    ray.hit.t = t; ray.hit.p = p; ray.hit.n = n; ray.hit.kind = HIT_SPHERE; ray.hit.mat = matIdx;
    ray.hit.prim = primIndex;
}

// RTContext.swift:529-538
static inline void intersectPlane(Ray& ray, const Plane& pl, int64_t matIdx, double eps, int64_t primIndex) {
    double denom = dot(pl.normal, ray.dir);
    if (std::fabs(denom) < eps) return;
    double t = dot(pl.center - ray.origin, pl.normal) / denom;
    if (t <= dmax(eps, ray.tMin) || t >= ray.hit.t) return;
    Vec3 p = ray.origin + ray.dir * t;
    Vec3 n = normalize(pl.normal);
    ray.hit.t = t; ray.hit.p = p; ray.hit.n = n; ray.hit.kind = HIT_PLANE; ray.hit.mat = matIdx;
    ray.hit.prim = primIndex;
}

// RTContext.swift:832-848
static inline bool triShadowHit(const Ray& ray, const Triangle& tr, double eps) {
    Vec3 offset = tr.motionBlur * ray.time;
    Vec3 origin = ray.origin - offset;
    Vec3 pvec = cross(ray.dir, tr.e2);
    double det = dot(tr.e1, pvec);
    if (std::fabs(det) < eps) return false;
    double invDet = 1.0 / det;
    Vec3 tvec = origin - tr.v0;
    double u = dot(tvec, pvec) * invDet;
    if (u < 0.0 || u > 1.0) return false;
    Vec3 q = cross(tvec, tr.e1);
    double v = dot(ray.dir, q) * invDet;
    if (v < 0.0 || u + v > 1.0) return false;
    double t = dot(tr.e2, q) * invDet;
    return (t > dmax(eps, ray.tMin) && t < ray.tMax);
}

// RTContext.swift:850-862
static inline bool sphereShadowHit(const Ray& ray, const Sphere& s, double eps) {
    Vec3 oc = ray.origin - s.center;
    double a = dot(ray.dir, ray.dir);
    double b = 2.0 * dot(oc, ray.dir);
    double c = dot(oc, oc) - s.radius * s.radius;
    double disc = b * b - 4 * a * c;
    if (disc < 0) return false;
    double sd = std::sqrt(disc);
    double t = (-b - sd) / (2 * a);
    if (t < dmax(eps, ray.tMin)) t = (-b + sd) / (2 * a);
    return (t > dmax(eps, ray.tMin) && t < ray.tMax);
}

// RTContext.swift:864-870
static inline bool planeShadowHit(const Ray& ray, const Plane& pl, double eps) {
    double denom = dot(pl.normal, ray.dir);
    if (std::fabs(denom) < eps) return false;
    double t = dot(pl.center - ray.origin, pl.normal) / denom;
    return (t > dmax(eps, ray.tMin) && t < ray.tMax);
}

// The reference's closest-hit stacks hold 64 entries with no overflow check
// (RTContext.swift:550,626); the oracle uses 256 and records the high-water mark.
static constexpr int kStackCap = 256;

// RTContext.swift:544-610
static inline int64_t intersectBLAS(const RTContext& ctx, Ray& ray, const Instance& inst, double eps, Counters& C) {
    const BLAS& bl = *inst.blas;
    const BVHNode* nodes = bl.nodes.data();
    uint64_t stack[kStackCap];
    int64_t visits = 0;
    int sp = 0;
    stack[sp] = bl.root;
    while (sp >= 0) {
        visits += 1;
        uint64_t idx = stack[sp]; sp -= 1;
        const BVHNode& node = nodes[idx];
        C.boxTests++;
        if (hitAABB(node, ray, eps) == kInf) continue;
        if (node.isLeaf()) {
            size_t first = node.leftFirst, count = node.primitiveCount;
            for (size_t k = 0; k < count; k++) {
                const PrimitiveInfo& p = bl.prims[bl.primIdx[first + k]];
                switch (p.type) {
                    case GEO_TRIANGLE:
                        C.triTests++;
                        intersectTriangle(ray, ctx.triangles[p.primitiveIndex], p.shadingMode == SHADE_SMOOTH,
                                          p.materialID, eps, p.primitiveIndex);
                        break;
                    case GEO_SPHERE:
                        C.sphereTests++;
                        intersectSphere(ray, ctx.spheres[p.primitiveIndex], p.materialID, eps, p.primitiveIndex);
                        break;
                    case GEO_PLANE:
                        C.planeTests++;
                        intersectPlane(ray, ctx.planes[p.primitiveIndex], p.materialID, eps, p.primitiveIndex);
                        break;
                    case GEO_MESH: break;
                }
            }
            continue;
        }
        uint64_t L = node.leftFirst, R = L + 1;
        C.boxTests += 2;
        double d1 = hitAABB(nodes[L], ray, eps), d2 = hitAABB(nodes[R], ray, eps);
        if (d1 > d2) { std::swap(d1, d2); std::swap(L, R); }
        if (d2 != kInf) { sp += 1; stack[sp] = R; }
        if (d1 != kInf) { sp += 1; stack[sp] = L; }
        if (sp + 1 > C.maxStack) C.maxStack = sp + 1;
    }
    return visits;
}

// world -> local ray, RTContext.swift:657-673 (and :755-769)
static inline Ray toLocal(const Ray& worldRay, const Instance& inst) {
    Vec3 instOffset = inst.instanceMotion * worldRay.time;
    Ray rL = worldRay;
    rL.origin -= instOffset;
    Vec4 o4 = mul(inst.worldToLocal, Vec4{rL.origin.x, rL.origin.y, rL.origin.z, 1.0});
    Vec4 d4 = mul(inst.worldToLocal, Vec4{rL.dir.x, rL.dir.y, rL.dir.z, 0.0});
    rL.origin = Vec3(o4.x, o4.y, o4.z);
    rL.dir = Vec3(d4.x, d4.y, d4.z);
    rL.invDir = 1.0 / rL.dir;
    return rL;  // tMax, hit copied from worldRay (:671-673)
}

// RTContext.swift:619-720
static inline int64_t intersectTLAS(const RTContext& ctx, Ray& worldRay, double eps, Counters& C) {
    const TLAS& tlas = ctx.tlas;
    const BVHNode* nodes = tlas.builder.bvhNode.data();
    const std::vector<PrimitiveInfo>& prims = tlas.builder.primitives;
    const std::vector<uint64_t>& pidx = tlas.builder.primitiveIdx;
    uint64_t stack[kStackCap];
    int64_t visits = 0;
    int sp = 0;
    stack[sp] = tlas.builder.rootNodeIdx;
    while (sp >= 0) {
        visits += 1;
        uint64_t idx = stack[sp]; sp -= 1;
        const BVHNode& node = nodes[idx];
        C.boxTests++;
        if (hitAABB(node, worldRay, eps) == kInf) continue;
        if (node.isLeaf()) {
            size_t first = node.leftFirst, count = node.primitiveCount;
            for (size_t k = 0; k < count; k++) {
                const PrimitiveInfo& p = prims[pidx[first + k]];
                if (p.type != GEO_MESH) continue;
                const Instance& inst = tlas.instances[p.primitiveIndex];
                C.instEntries++;
                Ray tmp = toLocal(worldRay, inst);
                visits += intersectBLAS(ctx, tmp, inst, eps, C);
                if (tmp.hit.t < worldRay.hit.t) {                         // :679
                    Vec3 instOffset = inst.instanceMotion * worldRay.time;
                    Vec4 wp = mul(inst.localToWorld, Vec4{tmp.hit.p.x, tmp.hit.p.y, tmp.hit.p.z, 1.0});
                    Vec3 wn = normalize(mul(inst.normalMatrix, tmp.hit.n));  // :686
                    if (determinant(upper3(inst.localToWorld)) < 0.0) wn = -wn;  // :691-698
                    worldRay.hit.t = tmp.hit.t;
                    worldRay.hit.p = Vec3(wp.x, wp.y, wp.z) + instOffset;
                    worldRay.hit.n = wn;
                    worldRay.hit.kind = tmp.hit.kind;
                    worldRay.hit.mat = inst.hasMaterialOverride ? inst.materialOverride : tmp.hit.mat;
                    worldRay.hit.prim = tmp.hit.prim;
                    worldRay.hit.inst = p.primitiveIndex;
                }
            }
            continue;
        }
        uint64_t L = node.leftFirst, R = L + 1;
        C.boxTests += 2;
        double d1 = hitAABB(nodes[L], worldRay, eps), d2 = hitAABB(nodes[R], worldRay, eps);
        if (d1 > d2) { std::swap(d1, d2); std::swap(L, R); }
        if (d2 != kInf) { sp += 1; stack[sp] = R; }
        if (d1 != kInf) { sp += 1; stack[sp] = L; }
        if (sp + 1 > C.maxStack) C.maxStack = sp + 1;
    }
    C.nodeVisits += visits;
    return visits;
}

// RTContext.swift:784-829.  The reference's heap [UInt] stack is a fixed array
// here (same pop order; faster than the reference, which favours the baseline).
static inline bool occludedBLAS(const RTContext& ctx, const Ray& r, const Instance& inst, double eps, Counters& C) {
    const BLAS& bl = *inst.blas;
    const BVHNode* nodes = bl.nodes.data();
    uint64_t stack[kStackCap];
    int sp = 0;
    stack[sp++] = bl.root;
    while (sp > 0) {
        uint64_t idx = stack[--sp];
        const BVHNode& node = nodes[idx];
        C.boxTests++;
        if (hitAABB(node, r, eps) == kInf) continue;
        if (node.isLeaf()) {
            size_t first = node.leftFirst, count = node.primitiveCount;
            for (size_t k = 0; k < count; k++) {
                const PrimitiveInfo& p = bl.prims[bl.primIdx[first + k]];
                switch (p.type) {
                    case GEO_TRIANGLE:
                        C.triTests++;
                        if (triShadowHit(r, ctx.triangles[p.primitiveIndex], eps)) return true;
                        break;
                    case GEO_SPHERE:
                        C.sphereTests++;
                        if (sphereShadowHit(r, ctx.spheres[p.primitiveIndex], eps)) return true;
                        break;
                    case GEO_PLANE:
                        C.planeTests++;
                        if (planeShadowHit(r, ctx.planes[p.primitiveIndex], eps)) return true;
                        break;
                    case GEO_MESH: break;
                }
            }
        } else {
            uint64_t L = node.leftFirst, R = L + 1;
            stack[sp++] = R; stack[sp++] = L;                            // :823-825
            if (sp > C.maxStack) C.maxStack = sp;
        }
    }
    return false;
}

// RTContext.swift:724-781
static inline bool occludedTLAS(const RTContext& ctx, const Ray& worldRay, double eps, Counters& C) {
    const TLAS& tlas = ctx.tlas;
    const BVHNode* nodes = tlas.builder.bvhNode.data();
    const std::vector<PrimitiveInfo>& prims = tlas.builder.primitives;
    const std::vector<uint64_t>& pidx = tlas.builder.primitiveIdx;
    uint64_t stack[kStackCap];
    int sp = 0;
    stack[sp++] = tlas.builder.rootNodeIdx;
    while (sp > 0) {
        uint64_t idx = stack[--sp];
        const BVHNode& node = nodes[idx];
        C.boxTests++;
        if (hitAABB(node, worldRay, eps) == kInf) continue;
        if (node.isLeaf()) {
            size_t first = node.leftFirst, count = node.primitiveCount;
            for (size_t k = 0; k < count; k++) {
                const PrimitiveInfo& p = prims[pidx[first + k]];
                if (p.type != GEO_MESH) continue;
                const Instance& inst = tlas.instances[p.primitiveIndex];
                C.instEntries++;
                Ray rL = toLocal(worldRay, inst);
                if (occludedBLAS(ctx, rL, inst, eps, C)) return true;
            }
        } else {
            uint64_t L = node.leftFirst, R = L + 1;
            stack[sp++] = R; stack[sp++] = L;
            if (sp > C.maxStack) C.maxStack = sp;
        }
    }
    return false;
}

}  // namespace orc
