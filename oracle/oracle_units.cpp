// oracle/oracle_units.cpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// Function-level C entry points used by the known-answer tests in tests/
// (the reference has none of its own).
#include <cstring>

#include "render_impl.hpp"
#include "rtcontext_init.hpp"

using namespace orc;

extern "C" {

// Object+Extension.swift:556-589 — n raw draws of PCG32(seed)
void orc_pcg32(uint64_t seed, int n, uint32_t* out) {
    PCG32 r(seed);
    for (int i = 0; i < n; i++) out[i] = r.next();
}
// per-pixel seed expression, Object+Extension.swift:294
uint64_t orc_pixel_seed(uint64_t i, uint64_t j) { return ((j << 32) ^ i) + 0x9E3779B97F4A7C15ULL; }

void orc_jitter_table(double* xs, double* ys) {
    PCG32 rng(0x123456789ABCDEFULL);
    buildStratifiedJitter(rng, xs, ys);
}

// out = {R, hasCosT, cosT, sin2T}
void orc_fresnel_dielectric(double n1, double n2, double cosI, double* out) {
    double R, cosT, s2; bool has;
    fresnelDielectric(n1, n2, cosI, R, has, cosT, s2);
    out[0] = R; out[1] = has ? 1.0 : 0.0; out[2] = cosT; out[3] = s2;
}
double orc_fresnel_conductor(double eta, double k, double cosI) { return fresnelConductorRGB(eta, k, cosI).x; }

void orc_orthonormal_basis(const double* n, double* t, double* b) {
    Vec3 T, B;
    orthonormalBasis(Vec3(n[0], n[1], n[2]), T, B);
    t[0] = T.x; t[1] = T.y; t[2] = T.z; b[0] = B.x; b[1] = B.y; b[2] = B.z;
}

// slab test, RTContext.swift:557-565: returns tmin or +inf
double orc_hit_aabb(const double* bmin, const double* bmax, const double* o, const double* d, double eps) {
    BVHNode n;
    n.aabbMin = Vec3(bmin[0], bmin[1], bmin[2]);
    n.aabbMax = Vec3(bmax[0], bmax[1], bmax[2]);
    Ray r(Vec3(o[0], o[1], o[2]), Vec3(d[0], d[1], d[2]));
    return hitAABB(n, r, eps);
}

// closest-hit Moeller-Trumbore, RTContext.swift:479-510.  v: 9 doubles.  Returns t or +inf; uvn = {nx,ny,nz,px,py,pz}
double orc_intersect_triangle(const double* v, const double* o, const double* d, double tMin, double hitT, double eps,
                              double* outNP) {
    Triangle tr;
    tr.v0 = Vec3(v[0], v[1], v[2]); tr.v1 = Vec3(v[3], v[4], v[5]); tr.v2 = Vec3(v[6], v[7], v[8]);
    tr.e1 = tr.v1 - tr.v0; tr.e2 = tr.v2 - tr.v0;
    Ray r(Vec3(o[0], o[1], o[2]), Vec3(d[0], d[1], d[2]));
    r.tMin = tMin; r.hit.t = hitT;
    intersectTriangle(r, tr, false, 1, eps, 0);
    if (r.hit.kind == HIT_NONE) return kInf;
    if (outNP) { outNP[0] = r.hit.n.x; outNP[1] = r.hit.n.y; outNP[2] = r.hit.n.z; outNP[3] = r.hit.p.x; outNP[4] = r.hit.p.y; outNP[5] = r.hit.p.z; }
    return r.hit.t;
}

// mode 0: intersectSphere (:512-527), mode 1: sphereShadowHit (:850-862, returns 1/0)
double orc_sphere(int mode, const double* c, double radius, const double* o, const double* d, double tMin, double tMaxOrHitT,
                  double eps) {
    Sphere s{Vec3(c[0], c[1], c[2]), radius};
    Ray r(Vec3(o[0], o[1], o[2]), Vec3(d[0], d[1], d[2]));
    r.tMin = tMin;
    if (mode == 0) { r.hit.t = tMaxOrHitT; intersectSphere(r, s, 1, eps, 0); return r.hit.kind == HIT_NONE ? kInf : r.hit.t; }
    r.tMax = tMaxOrHitT;
    return sphereShadowHit(r, s, eps) ? 1.0 : 0.0;
}

double orc_plane(int mode, const double* c, const double* n, const double* o, const double* d, double tMin,
                 double tMaxOrHitT, double eps) {
    Plane p{Vec3(c[0], c[1], c[2]), Vec3(n[0], n[1], n[2])};
    Ray r(Vec3(o[0], o[1], o[2]), Vec3(d[0], d[1], d[2]));
    r.tMin = tMin;
    if (mode == 0) { r.hit.t = tMaxOrHitT; intersectPlane(r, p, 1, eps, 0); return r.hit.kind == HIT_NONE ? kInf : r.hit.t; }
    r.tMax = tMaxOrHitT;
    return planeShadowHit(r, p, eps) ? 1.0 : 0.0;
}

// BVHBuilder on n boxes (bounds: n*6 doubles, centroid = given n*3).  Writes up to 2n-1 nodes:
// nodeBounds[k*6], nodeMeta[k*2] = {leftFirst, primitiveCount}; primIdx[n].  Returns nodesUsed+1.
int64_t orc_bvh_build(int64_t n, const double* bounds, const double* centroids, double* nodeBounds, int64_t* nodeMeta,
                      int64_t* primIdx) {
    std::vector<PrimitiveInfo> prims((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        prims[i].type = GEO_TRIANGLE;
        prims[i].primitiveIndex = i;
        prims[i].bounds = AABB(Vec3(bounds[i * 6], bounds[i * 6 + 1], bounds[i * 6 + 2]),
                               Vec3(bounds[i * 6 + 3], bounds[i * 6 + 4], bounds[i * 6 + 5]));
        prims[i].centroid = Vec3(centroids[i * 3], centroids[i * 3 + 1], centroids[i * 3 + 2]);
    }
    BVHBuilder b(std::move(prims), 2, 12);
    int64_t used = (int64_t)b.nodesUsed + 1;
    for (int64_t k = 0; k < used; k++) {
        const BVHNode& nd = b.bvhNode[k];
        nodeBounds[k * 6] = nd.aabbMin.x; nodeBounds[k * 6 + 1] = nd.aabbMin.y; nodeBounds[k * 6 + 2] = nd.aabbMin.z;
        nodeBounds[k * 6 + 3] = nd.aabbMax.x; nodeBounds[k * 6 + 4] = nd.aabbMax.y; nodeBounds[k * 6 + 5] = nd.aabbMax.z;
        nodeMeta[k * 2] = (int64_t)nd.leftFirst; nodeMeta[k * 2 + 1] = (int64_t)nd.primitiveCount;
    }
    for (int64_t i = 0; i < n; i++) primIdx[i] = (int64_t)b.primitiveIdx[i];
    return used;
}

// AABB.transformed(by:), AABB.swift:71-92.  box: 6 doubles, M: 16 doubles column-major, out: 6 doubles
void orc_aabb_transformed(const double* box, const double* M, double* out) {
    AABB a(Vec3(box[0], box[1], box[2]), Vec3(box[3], box[4], box[5]));
    Mat4 m; std::memcpy(m.m, M, sizeof m.m);
    AABB r = a.transformed(m);
    out[0] = r.minP.x; out[1] = r.minP.y; out[2] = r.minP.z; out[3] = r.maxP.x; out[4] = r.maxP.y; out[5] = r.maxP.z;
}

// PLY load through the reference's CPly (useRef=1, cplyLib required) or the built-in reader (useRef=0).
// Two-call protocol: pass NULL buffers to get counts in nv/ni/nn.
int orc_ply_load(const char* path, int useRef, const char* cplyLib, int64_t* nv, int64_t* ni, int64_t* nn, double* pos,
                 int64_t* idx, double* nrm) {
    PlyMesh m;
    bool ok;
    if (useRef) {
        CPlyApi& A = cplyApi(cplyLib);
        if (!A.ok) return -2;
        ok = loadPlyViaCPly(A, path, m);
    } else {
        ok = loadPlyBuiltin(path, m);
    }
    if (!ok) return -1;
    *nv = (int64_t)m.positions.size(); *ni = (int64_t)m.indices.size(); *nn = (int64_t)m.normals.size();
    if (pos) for (size_t i = 0; i < m.positions.size(); i++) { pos[3 * i] = m.positions[i].x; pos[3 * i + 1] = m.positions[i].y; pos[3 * i + 2] = m.positions[i].z; }
    if (idx) for (size_t i = 0; i < m.indices.size(); i++) idx[i] = m.indices[i];
    if (nrm) for (size_t i = 0; i < m.normals.size(); i++) { nrm[3 * i] = m.normals[i].x; nrm[3 * i + 1] = m.normals[i].y; nrm[3 * i + 2] = m.normals[i].z; }
    return 0;
}

}  // extern "C"
