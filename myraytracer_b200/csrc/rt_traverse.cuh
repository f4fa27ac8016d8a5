// rt_traverse.cuh — two-level traversal and primitive tests on the device.
//
// Replaces, for the GPU layout of gpu_layout.h:
//   intersectTLAS / intersectBLAS     Models/RTContext.swift:619-720, 544-610
//   occludedTLAS / occludedBLAS       Models/RTContext.swift:724-829
//   intersectTriangle/Sphere/Plane    Models/RTContext.swift:479-538
//   tri/sphere/planeShadowHit         Models/RTContext.swift:832-870
//   hitAABB                           Models/RTContext.swift:557-565
// Precision policy: primitive tests and the world->local ray are
// FP64 in the reference's exact operation order (bit-identical t, u, v);
// BLAS box tests are FP32 against boxes padded/rounded outward, evaluated from
// a ray origin advanced to the BLAS in FP64 — conservative, so no hit the
// reference finds is culled; TLAS boxes are FP64 with a widened interval.
#pragma once
#include "gpu_layout.h"
#include "rt_kernels.h"
#include "rt_math.cuh"

namespace rtg {


struct Counters {
    unsigned long long box, tri, sphere, plane, inst, smooth;
    int maxStack;
};

struct HitRec {
    double t, u, v;   // u, v: barycentrics of a triangle hit
    int32_t inst;     // -1 = miss
    int32_t slot;     // triangle slot (leaf order) within the global GTri/GTriD array
};

// FP64 slab test of the reference, bit-exact (used for analytic instances' 1-primitive BLAS root):
// returns true iff tmax >= max(tmin, eps).
__device__ __forceinline__ bool hitAABBExact(const double* lo, const double* hi, V3 o, V3 inv, double eps) {
    double t1x = (lo[0] - o.x) * inv.x, t1y = (lo[1] - o.y) * inv.y, t1z = (lo[2] - o.z) * inv.z;
    double t2x = (hi[0] - o.x) * inv.x, t2y = (hi[1] - o.y) * inv.y, t2z = (hi[2] - o.z) * inv.z;
    double tmin = dmax(dmax(dmin(t1x, t2x), dmin(t1y, t2y)), dmin(t1z, t2z));
    double mx = dmax(t1x, t2x), my = dmax(t1y, t2y), mz = dmax(t1z, t2z);
    double tmax = dmin(mx, dmin(my, mz));
    return tmax >= dmax(tmin, eps);
}

// Widened FP64 slab interval using an approximate reciprocal (relative error ~1e-13, see refinedInv):
// [tn, tf] contains the exact interval.  NaN lanes (0*inf) drop out through fmin/fmax.
__device__ __forceinline__ void slabApprox(const double* lo, const double* hi, V3 o, V3 ia, double& tn, double& tf) {
    double t1x = (lo[0] - o.x) * ia.x, t1y = (lo[1] - o.y) * ia.y, t1z = (lo[2] - o.z) * ia.z;
    double t2x = (hi[0] - o.x) * ia.x, t2y = (hi[1] - o.y) * ia.y, t2z = (hi[2] - o.z) * ia.z;
    tn = fmax(fmax(fmin(t1x, t2x), fmin(t1y, t2y)), fmin(t1z, t2z));
    tf = fmin(fmin(fmax(t1x, t2x), fmax(t1y, t2y)), fmax(t1z, t2z));
    const double k = 1.0e-9;
    tn = tn - fabs(tn) * k;
    tf = tf + fabs(tf) * k;
}

// The conservative FP32 view of a ray against boxes stored relative to `centre` (BLAS nodes, TLAS nodes): the origin
// is advanced in FP64 to its entry into the exact bounds [lo, hi] before it is rounded, so the FP32 error stays
// relative to the box size; (plane - o) / d is one FMA, plane * inv + (-o * inv), with |inv| clamped to 1e30 so
// that no inf - inf arises.  The stored boxes' padding covers every rounding here.
struct RayF {
    float ix, iy, iz, nx, ny, nz, tmin, tmax;
    double t0;
    bool hits;     // the exact bounds are reached within [tLow, tHigh] (widened FP64 test)
};
__device__ __forceinline__ RayF makeRayF(const double* lo, const double* hi, const double* centre, V3 o, V3 d, V3 ia,
                                         double tLow, double tHigh) {
    RayF r;
    const V3 oc = o - ld3(centre);
    const double l[3] = {lo[0] - centre[0], lo[1] - centre[1], lo[2] - centre[2]};
    const double h[3] = {hi[0] - centre[0], hi[1] - centre[1], hi[2] - centre[2]};
    double tn, tf;
    slabApprox(l, h, oc, ia, tn, tf);
    r.hits = !(tn > tf || tf < tLow || tn > tHigh);
    double t0 = tn > 0.0 ? tn : 0.0;
    if (!(t0 < 1e300)) t0 = 0.0;
    r.t0 = t0;
    const V3 o2 = oc + d * t0;
    const float ox = (float)o2.x, oy = (float)o2.y, oz = (float)o2.z;
    r.ix = fminf(fmaxf(1.0f / (float)d.x, -1e30f), 1e30f);
    r.iy = fminf(fmaxf(1.0f / (float)d.y, -1e30f), 1e30f);
    r.iz = fminf(fmaxf(1.0f / (float)d.z, -1e30f), 1e30f);
    r.nx = -(ox * r.ix); r.ny = -(oy * r.iy); r.nz = -(oz * r.iz);
    r.tmin = __double2float_rd(tLow - t0);
    r.tmax = __double2float_ru(tHigh - t0);
    return r;
}
// both children of a node record against the ray: entry distances n0, n1 (relative to t0) and hit flags
__device__ __forceinline__ void twoBoxes(const GNode* node, const RayF& r, float& n0, bool& h0, float& n1, bool& h1, int2& child) {
    const float4* np = reinterpret_cast<const float4*>(node);
    const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2);
    const int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
    const float a0 = __fmaf_rn(q0.x, r.ix, r.nx), b0 = __fmaf_rn(q0.w, r.ix, r.nx);
    const float a1 = __fmaf_rn(q0.y, r.iy, r.ny), b1 = __fmaf_rn(q1.x, r.iy, r.ny);
    const float a2 = __fmaf_rn(q0.z, r.iz, r.nz), b2 = __fmaf_rn(q1.y, r.iz, r.nz);
    n0 = fmaxf(fmaxf(fminf(a0, b0), fminf(a1, b1)), fmaxf(fminf(a2, b2), r.tmin));
    const float f0 = fminf(fminf(fmaxf(a0, b0), fmaxf(a1, b1)), fminf(fmaxf(a2, b2), r.tmax));
    const float c0 = __fmaf_rn(q1.z, r.ix, r.nx), d0 = __fmaf_rn(q2.y, r.ix, r.nx);
    const float c1 = __fmaf_rn(q1.w, r.iy, r.ny), d1 = __fmaf_rn(q2.z, r.iy, r.ny);
    const float c2 = __fmaf_rn(q2.x, r.iz, r.nz), d2 = __fmaf_rn(q2.w, r.iz, r.nz);
    n1 = fmaxf(fmaxf(fminf(c0, d0), fminf(c1, d1)), fmaxf(fminf(c2, d2), r.tmin));
    const float f1 = fminf(fminf(fmaxf(c0, d0), fmaxf(c1, d1)), fminf(fmaxf(c2, d2), r.tmax));
    h0 = n0 <= f0; h1 = n1 <= f1;
    child = make_int2(ch.x, ch.y);
}

// 1/d to ~1e-13 relative without an FP64 division: FP32 reciprocal + one Newton step.
__device__ __forceinline__ double refinedInv1(double d) {
    double r = (double)(1.0f / (float)d);
    double r2 = r * (2.0 - d * r);
    return (isfinite(r) && r != 0.0) ? r2 : r;
}
__device__ __forceinline__ V3 refinedInv(V3 d) { return mk(refinedInv1(d.x), refinedInv1(d.y), refinedInv1(d.z)); }

// Moeller-Trumbore in the reference's operation order (RTContext.swift:479-494 / :832-847).
// The divisions-free early outs reject exactly the cases the reference rejects:
//   sign(num) != sign(det)      -> u (or v) < 0
//   |num| > |det| (1 + 1e-15)   -> u > 1 (or u + v > 1)
// so invDet = 1/det is only evaluated for (near-)hits.  Returns true with t,u,v when
// u, v pass; the caller applies the t window.
__device__ __forceinline__ bool mtTest(V3 v0, V3 v1, V3 v2, V3 org, V3 dir, double eps, double& t, double& u, double& v) {
    V3 e1 = v1 - v0, e2 = v2 - v0;
    V3 pvec = cross(dir, e2);
    double det = dot(e1, pvec);
    double adet = fabs(det);
    if (adet < eps) return false;
    V3 tvec = org - v0;
    double nu = dot(tvec, pvec);
    const double tiny = 1e-290, big = 1e290;
    bool safe = adet < big;
    double lim = adet * (1.0 + 1e-15);
    if (safe && fabs(nu) > tiny) {
        if ((nu < 0.0) != (det < 0.0)) return false;
        if (fabs(nu) > lim) return false;
    }
    V3 q = cross(tvec, e1);
    double nv = dot(dir, q);
    if (safe && fabs(nv) > tiny) {
        if ((nv < 0.0) != (det < 0.0)) return false;
        if (fabs(nu) + fabs(nv) > lim * (1.0 + 1e-15) && fabs(nu) > tiny) return false;
    }
    double invDet = ddiv(1.0, det);
    u = nu * invDet;
    if (u < 0.0 || u > 1.0) return false;
    v = nv * invDet;
    if (v < 0.0 || u + v > 1.0) return false;
    t = dot(e2, q) * invDet;
    return true;
}

// t window: closest-hit rejects `t <= max(eps,tMin) || t >= hit.t` (RTContext.swift:494,523,534); any-hit accepts
// `t > max(eps,tMin) && t < tMax` (:847,861,869).  The two forms differ only for NaN, kept as in the reference.
__device__ __forceinline__ bool inWindow(double t, double tLow, double tHigh, bool anyhit) {
    return anyhit ? (t > tLow && t < tHigh) : !(t <= tLow || t >= tHigh);
}

// RTContext.swift:512-523 (closest) and :851-861 (shadow: second-root rule uses max(eps,tMin))
__device__ __forceinline__ bool sphereT(const double* sp, V3 o, V3 d, double eps, double rootLimit, double& t) {
    V3 oc = o - ld3(sp);
    double a = dot(d, d);
    double b = 2.0 * dot(oc, d);
    double c = dot(oc, oc) - sp[3] * sp[3];
    double disc = b * b - 4 * a * c;
    if (disc < 0) return false;
    double sd = dsqrt(disc);
    t = ddiv(-b - sd, 2 * a);
    if (t < rootLimit) t = ddiv(-b + sd, 2 * a);
    return true;
}

// RTContext.swift:530-534 / :865-869
__device__ __forceinline__ bool planeT(const double* pl, V3 o, V3 d, double eps, double& t) {
    V3 n = ld3(pl + 3);
    double denom = dot(n, d);
    if (fabs(denom) < eps) return false;
    t = ddiv(dot(ld3(pl) - o, n), denom);
    return true;
}

__device__ __forceinline__ void loadTri(const DeviceScene& S, bool f64, int slot, V3& v0, V3& v1, V3& v2) {
    if (f64) {
        const GTriD* p = S.trisD + slot;
        v0 = ld3(p->v0); v1 = ld3(p->v1); v2 = ld3(p->v2);
    } else {
        const float4* p = reinterpret_cast<const float4*>(S.tris + slot);
        float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
        v0 = mk((double)a.x, (double)a.y, (double)a.z);
        v1 = mk((double)b.x, (double)b.y, (double)b.z);
        v2 = mk((double)c.x, (double)c.y, (double)c.z);
    }
}

// One BLAS (mesh instance) traversal.  anyhit (run-time flag: ONE copy of the loop serves closest-hit and
// shadow queries, which halves the instruction-cache footprint): return true at the first accepted primitive.
// oL/dL: local-space FP64 ray (bit-identical to the reference's).  tLow = max(eps, tMin).
// hit.t is the current closest t (closest) or tMax (any-hit); updated on accept.
// MOTION (general kernel, scenes with motion blur): the triangles of a mesh with Mesh.motionBlur are displaced by
// motionBlur * ray.time, which the reference applies to the ray origin of the triangle test only
// (RTContext.swift:480-481,833-834); the node boxes were grown by the motion vector at build time (:283-289) and are
// tested with the undisplaced ray, like the reference's.
// BUDGET (the tile probe of a first frame only): *budget node visits at most — the walk stops when they are spent and
// whatever it found stands (a cost estimate needs no more than "this ray is expensive").
template <bool COUNT, bool MOTION = false, bool BUDGET = false>
__device__ __forceinline__ bool traverseBlas(const DeviceScene& S, const GInstance& I, int instIdx, V3 oL, V3 dL, double eps,
                                             double tLow, HitRec& hit, bool anyhit, int32_t* sstack, int sstackN, Counters& C,
                                             double time = 0.0, int* budget = nullptr) {
    const V3 oTri = MOTION ? oL - ld3(I.meshMotion) * time : oL;
    // Advance the origin to the BLAS in FP64 so that its FP32 copy is accurate relative to the
    // box size: t0 = entry distance into the (exact) root box, from a widened FP64 slab test.
    V3 ia = refinedInv(dL);
    V3 oc = oL - ld3(I.centre);
    double lo[3] = {I.rootLo[0] - I.centre[0], I.rootLo[1] - I.centre[1], I.rootLo[2] - I.centre[2]};
    double hi[3] = {I.rootHi[0] - I.centre[0], I.rootHi[1] - I.centre[1], I.rootHi[2] - I.centre[2]};
    double tn, tf;
    slabApprox(lo, hi, oc, ia, tn, tf);
    if (COUNT) C.box++;
    if (tn > tf || tf < tLow || tn > hit.t) return false;  // widened interval: conservative (NaN -> keep)
    double t0 = tn > 0.0 ? tn : 0.0;
    if (!(t0 < 1e300)) t0 = 0.0;
    V3 o2 = oc + dL * t0;
    const float ox = (float)o2.x, oy = (float)o2.y, oz = (float)o2.z;
    // |1/d| clamped to 1e30 (a zero direction component becomes +-1e-30: moves < pad over any finite t) so that
    // plane * inv - o * inv never produces inf - inf
    const float ix = fminf(fmaxf(1.0f / (float)dL.x, -1e30f), 1e30f);
    const float iy = fminf(fmaxf(1.0f / (float)dL.y, -1e30f), 1e30f);
    const float iz = fminf(fmaxf(1.0f / (float)dL.z, -1e30f), 1e30f);
    const float nx = -(ox * ix), ny = -(oy * iy), nz = -(oz * iz);
    float tminF = __double2float_rd(tLow - t0);
    float tmaxF = __double2float_ru(hit.t - t0);

    const GNode* nodes = S.nodes;        // child references are absolute (gpu_layout.h)
    // stack: the first sstackN (<= RT_SSTACK) entries in shared memory (per-lane column), deeper ones in local memory
    int32_t lstack[RT_STACK];
    int sp = 0;
    int32_t cur = I.rootRef;
    const int tid = threadIdx.x;
    for (;;) {
        while ((uint32_t)cur < 0x7fffffffu) {  // inner node
            if (BUDGET && --*budget < 0) return false;
            const float4* np = reinterpret_cast<const float4*>(nodes + cur);
            const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2);
            const int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
            if (COUNT) C.box += 2;
            // child 0: lo = q0.xyz, hi = (q0.w, q1.x, q1.y); child 1: lo = (q1.z, q1.w, q2.x), hi = q2.yzw
            // (plane - o) * inv as one explicit FMA: plane * inv + (-o * inv); intrinsics are immune to -fmad=false
            float a0 = __fmaf_rn(q0.x, ix, nx), b0 = __fmaf_rn(q0.w, ix, nx);
            float a1 = __fmaf_rn(q0.y, iy, ny), b1 = __fmaf_rn(q1.x, iy, ny);
            float a2 = __fmaf_rn(q0.z, iz, nz), b2 = __fmaf_rn(q1.y, iz, nz);
            float n0 = fmaxf(fmaxf(fminf(a0, b0), fminf(a1, b1)), fmaxf(fminf(a2, b2), tminF));
            float f0 = fminf(fminf(fmaxf(a0, b0), fmaxf(a1, b1)), fminf(fmaxf(a2, b2), tmaxF));
            float c0 = __fmaf_rn(q1.z, ix, nx), d0 = __fmaf_rn(q2.y, ix, nx);
            float c1 = __fmaf_rn(q1.w, iy, ny), d1 = __fmaf_rn(q2.z, iy, ny);
            float c2 = __fmaf_rn(q2.x, iz, nz), d2 = __fmaf_rn(q2.w, iz, nz);
            float n1 = fmaxf(fmaxf(fminf(c0, d0), fminf(c1, d1)), fmaxf(fminf(c2, d2), tminF));
            float f1 = fminf(fminf(fmaxf(c0, d0), fmaxf(c1, d1)), fminf(fmaxf(c2, d2), tmaxF));
            // the padding of the stored boxes covers every FP32 rounding here
            const bool h0 = n0 <= f0, h1 = n1 <= f1;
            if (h0 && h1) {
                int32_t nearC = ch.x, farC = ch.y;
                if (n1 < n0) { nearC = ch.y; farC = ch.x; }
                if (sp < sstackN) sstack[sp * RT_BLOCK + tid] = farC; else lstack[sp - sstackN] = farC;
                sp++;
                if (COUNT && sp > C.maxStack) C.maxStack = sp;
                cur = nearC;
            } else if (h0) {
                cur = ch.x;
            } else if (h1) {
                cur = ch.y;
            } else {
                if (sp == 0) return false;
                sp--;
                cur = sp < sstackN ? sstack[sp * RT_BLOCK + tid] : lstack[sp - sstackN];
            }
        }
        if (cur != kEmptyRef) {  // leaf
            const uint32_t ref = ~(uint32_t)cur;
            const int first = (int)(ref >> 4), count = (int)(ref & 7u) + 1;
            const bool f64 = (ref & 8u) != 0;
            for (int k = 0; k < count; k++) {
                V3 v0, v1, v2;
                loadTri(S, f64, first + k, v0, v1, v2);
                if (COUNT) C.tri++;
                double t, u, v;
                if (!mtTest(v0, v1, v2, oTri, dL, eps, t, u, v)) continue;
                if (!inWindow(t, tLow, hit.t, anyhit)) continue;
                if (anyhit) return true;
                hit.t = t; hit.u = u; hit.v = v; hit.inst = instIdx; hit.slot = first + k;
                tmaxF = __double2float_ru(t - t0);
            }
        }
        if (sp == 0) return false;
        sp--;
        cur = sp < sstackN ? sstack[sp * RT_BLOCK + tid] : lstack[sp - sstackN];
    }
}

}  // namespace rtg
