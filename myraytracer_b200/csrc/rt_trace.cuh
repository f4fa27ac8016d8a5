// rt_trace.cuh — TLAS walk, instance entry, hit resolution (device).
// See rt_traverse.cuh for the reference lines each piece replaces.
#pragma once
#include "rt_traverse.cuh"

namespace rtg {

#define RT_TLAS_STACK 40

// world -> local ray (RTContext.swift:657-669).  MOTION: the instance is displaced by instanceMotion * ray.time,
// applied to the ray origin before the transform (:658-661, :756-758).
template <bool MOTION = false>
__device__ __forceinline__ void toLocal(const GInstance& I, V3 o, V3 d, V3& oL, V3& dL, double time = 0.0) {
    if (MOTION) o = o - ld3(I.motion) * time;
    if (I.identity) { oL = o; dL = d; }
    else { oL = xfmPoint(I.w2l, o); dL = xfmVector(I.w2l, d); }
}

// Analytic instance (sphere / plane / lone triangle): the reference's 1-primitive BLAS root test, then the
// primitive, bit-exact FP64.  oL/dL: local-space ray.  Returns true for an any-hit.
template <bool COUNT>
__device__ __forceinline__ bool analyticInstance(const DeviceScene& S, const GInstance& I, int idx, V3 oL, V3 dL, double eps,
                                                 double tLow, HitRec& hit, bool anyhit, Counters& C) {
    const int kind = I.kind;
    V3 inv = mk(ddiv(1.0, dL.x), ddiv(1.0, dL.y), ddiv(1.0, dL.z));
    if (COUNT) C.box++;
    if (!hitAABBExact(I.rootLo, I.rootHi, oL, inv, eps)) return false;
    double t, u = 0, v = 0;
    if (kind == INST_SPHERE) {
        if (COUNT) C.sphere++;
        // closest: second root if t < eps (RTContext.swift:522); shadow: if t < max(eps,tMin) (:860)
        if (!sphereT(I.prim, oL, dL, eps, anyhit ? tLow : eps, t)) return false;
    } else if (kind == INST_PLANE) {
        if (COUNT) C.plane++;
        if (!planeT(I.prim, oL, dL, eps, t)) return false;
    } else {
        if (COUNT) C.tri++;
        V3 v0, v1, v2;
        loadTri(S, I.triF64 != 0, I.triBase, v0, v1, v2);
        if (!mtTest(v0, v1, v2, oL, dL, eps, t, u, v)) return false;
    }
    if (!inWindow(t, tLow, hit.t, anyhit)) return false;
    if (anyhit) return true;
    hit.t = t; hit.u = u; hit.v = v; hit.inst = idx; hit.slot = I.triBase;
    return false;
}

// Closest-hit / any-hit test of one instance.  tLow = max(eps, tMin); hit.t = current closest t or tMax.
template <bool COUNT, bool MOTION = false, bool BUDGET = false>
__device__ __forceinline__ bool enterInstance(const DeviceScene& S, int idx, V3 o, V3 d, double eps, double tMin,
                                              HitRec& hit, bool anyhit, int32_t* sstack, int sstackN, Counters& C, double time = 0.0,
                                              int* budget = nullptr) {
    const GInstance& I = S.instances[idx];
    V3 oL, dL;
    toLocal<MOTION>(I, o, d, oL, dL, time);
    if (COUNT) C.inst++;
    const double tLow = dmax(eps, tMin);
    if (I.kind == INST_MESH) return traverseBlas<COUNT, MOTION, BUDGET>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, sstack, sstackN, C, time, budget);
    return analyticInstance<COUNT>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, C);
}

// Scenes with instanceMotion: the reference's instance boxes (Instance.worldBounds, RTContext.swift:437-455) do NOT
// follow the motion, and its TLAS walk enters an instance iff the UNDISPLACED ray passes hitAABB on every node from
// the root to the instance's leaf (:632-640, :745-747) — so which displaced instances a ray can reach depends on the
// topology of the reference's own TLAS.  For those scenes the host rebuilds that TLAS (bvh_build.cpp:
// buildReferenceTlas = BVHBuilder(maxLeaf 2, SAH, 12 bins) over the instances' bounds) and this walk reproduces the
// reference's visit rule exactly in FP64: pop, test the node's own box, leaf -> its instances in primitiveIdx order,
// inner -> both children tested, nearer pushed last (:600-606,710-716).
__device__ __forceinline__ double hitRefBox(const RefNode& n, V3 o, V3 inv, double eps) {
    const double t1x = (n.lo[0] - o.x) * inv.x, t1y = (n.lo[1] - o.y) * inv.y, t1z = (n.lo[2] - o.z) * inv.z;
    const double t2x = (n.hi[0] - o.x) * inv.x, t2y = (n.hi[1] - o.y) * inv.y, t2z = (n.hi[2] - o.z) * inv.z;
    const double tmin = dmax(dmax(dmin(t1x, t2x), dmin(t1y, t2y)), dmin(t1z, t2z));
    const double tmax = dmin(dmax(t1x, t2x), dmin(dmax(t1y, t2y), dmax(t1z, t2z)));
    return (tmax >= dmax(tmin, eps)) ? tmin : CUDART_INF;
}
template <bool COUNT>
__device__ __forceinline__ bool traceSceneRef(const DeviceScene& S, V3 o, V3 d, double tMin, HitRec& hit, bool anyhit,
                                              int32_t* sstack, int sstackN, Counters& C, double time) {
    const double eps = S.eps;
    const V3 inv = mk(ddiv(1.0, d.x), ddiv(1.0, d.y), ddiv(1.0, d.z));       // Ray.invDir
    int32_t stack[64];                                          // the reference's fixed 64 slots (:626)
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const RefNode& node = S.refTlas[stack[--sp]];
        if (COUNT) C.box++;
        if (hitRefBox(node, o, inv, eps) == CUDART_INF) continue;
        if (node.count > 0) {
            for (int k = 0; k < node.count; k++)
                if (enterInstance<COUNT, true>(S, S.refTlasPrim[node.leftFirst + k], o, d, eps, tMin, hit, anyhit, sstack, sstackN, C, time))
                    return true;
            continue;
        }
        int32_t L = node.leftFirst, R = L + 1;
        if (anyhit) { if (sp + 2 <= 64) { stack[sp++] = R; stack[sp++] = L; } continue; }      // :775-777
        double d1 = hitRefBox(S.refTlas[L], o, inv, eps), d2 = hitRefBox(S.refTlas[R], o, inv, eps);
        if (COUNT) C.box += 2;
        if (d1 > d2) { const double td = d1; d1 = d2; d2 = td; const int32_t tn = L; L = R; R = tn; }
        if (d2 != CUDART_INF && sp < 64) stack[sp++] = R;
        if (d1 != CUDART_INF && sp < 64) stack[sp++] = L;
    }
    return false;
}

// TLAS walk: conservative FP32 boxes (unions of Instance.worldBounds) like the BLAS nodes, near child first; every
// instance reached is then tested exactly.
template <bool COUNT, bool MOTION = false, bool BUDGET = false>
__device__ __forceinline__ bool traceScene(const DeviceScene& S, V3 o, V3 d, double tMin, HitRec& hit, bool anyhit,
                                           int32_t* sstack, int sstackN, Counters& C, double time = 0.0, int* budget = nullptr) {
    const double eps = S.eps;
    int32_t cur = S.tlasRootRef;
    if (cur == kEmptyRef) return false;
    if (MOTION && S.refTlasNodes > 0) return traceSceneRef<COUNT>(S, o, d, tMin, hit, anyhit, sstack, sstackN, C, time);
    if (cur < 0) return enterInstance<COUNT, MOTION, BUDGET>(S, ~cur, o, d, eps, tMin, hit, anyhit, sstack, sstackN, C, time, budget);
    const double tLow = dmax(eps, tMin);
    RayF rf = makeRayF(S.tlasLo, S.tlasHi, S.tlasCentre, o, d, refinedInv(d), tLow, hit.t);
    if (COUNT) C.box++;
    if (!rf.hits) return false;
    int32_t tstack[RT_TLAS_STACK];
    int sp = 0;
    for (;;) {
        if (cur >= 0) {
            float n0, n1;
            bool h0, h1;
            int2 ch;
            // reference: enter iff tmax >= max(tmin, eps) (no culling against hit.t in the box test,
            // RTContext.swift:632-640); culling against hit.t (rf.tmax) cannot change the closest hit.
            twoBoxes(S.tlas + cur, rf, n0, h0, n1, h1, ch);
            if (COUNT) C.box += 2;
            int32_t c0 = ch.x, c1 = ch.y;
            if (h0 && h1) {
                if (n1 < n0) { int32_t t = c0; c0 = c1; c1 = t; }
                if (sp < RT_TLAS_STACK) tstack[sp++] = c1;
                cur = c0;
                continue;
            } else if (h0) { cur = c0; continue; }
            else if (h1) { cur = c1; continue; }
        } else {
            if (enterInstance<COUNT, MOTION, BUDGET>(S, ~cur, o, d, eps, tMin, hit, anyhit, sstack, sstackN, C, time, budget)) return true;
            if (BUDGET && *budget < 0) return false;
            rf.tmax = __double2float_ru(hit.t - rf.t0);       // a closer hit prunes the boxes that follow
        }
        if (sp == 0) return false;
        cur = tstack[--sp];
    }
}

// The same query as traceScene with the two levels in ONE loop (the general kernel's walk).  Nested — a TLAS loop whose
// leaves call a BLAS loop — the lanes of a warp enter a mesh at different iterations of the outer loop and walk it a
// few at a time: ncu shows 2.3 to 4.2 of 32 lanes active in the node loop of the mesh walk on C3 with glossy materials,
// half of that kernel's instructions.  TLAS and BLAS nodes share one record format (GNode), so here a lane is simply at
// "a node": the node step is one piece of code for both levels, every lane that has a ray executes it whatever level
// it is on, and entering an instance is a leaf action — transform the ray, root test, swap the FP32 ray view and
// the node array, push a return marker under the BLAS root; popping the marker swaps back.  Same visiting order as
// the nested loops (an instance is a subtree), so the same hits and the same tie-breaking (every emulated and GPU
// parity case passes with it).
// MEASURED AND NOT USED (-DRT_GENERAL_WALK=traceSceneFlat builds it; B200, ms per frame, flat / nested, profiles/r2ac):
// C3 glossy 11.65 / 12.11, but C3 with area lights 22.86 / 20.03, C3 through the general kernel 10.30 / 9.47, and the
// mesh scenes forced through the general kernel C4 37.9 / 27.7, C5 410 / 284, C2 0.75 / 0.50.  More lanes are
// active per instruction, yet every iteration goes through a four-way dispatch (inner node, return marker, triangles,
// instance entry) and pays for every kind of step some lane is at, where the nested BLAS loop is 35 instructions that
// stay in the instruction cache; a lane that enters an instance (150 FP64 instructions) stalls 31 node steps.
static const int32_t kReturnRef = 0x7ffffffe;      // stack marker: the entries above it belonged to an instance's BLAS
#ifndef RT_FLAT_STACK
#define RT_FLAT_STACK (RT_STACK + RT_TLAS_STACK + 2)
#endif
template <bool COUNT, bool MOTION = false>
__device__ __forceinline__ bool traceSceneFlat(const DeviceScene& S, V3 o, V3 d, double tMin, HitRec& hit, bool anyhit,
                                               int32_t* sstack, int sstackN, Counters& C, double time = 0.0) {
    const double eps = S.eps;
    int32_t cur = S.tlasRootRef;
    if (cur == kEmptyRef) return false;
    if (MOTION && S.refTlasNodes > 0) return traceSceneRef<COUNT>(S, o, d, tMin, hit, anyhit, sstack, sstackN, C, time);
    const double tLow = dmax(eps, tMin);
    RayF top;                                   // the ray against the TLAS boxes (unused when the scene is one instance)
    top.ix = top.iy = top.iz = top.nx = top.ny = top.nz = top.tmin = top.tmax = 0.0f; top.t0 = 0.0; top.hits = true;
    if (cur >= 0) {
        top = makeRayF(S.tlasLo, S.tlasHi, S.tlasCentre, o, d, refinedInv(d), tLow, hit.t);
        if (COUNT) C.box++;
        if (!top.hits) return false;
    }
    RayF a = top;                               // the view the node step uses: `top`, or the current instance's
    const GNode* base = S.tlas;
    bool inBlas = false;
    int instIdx = -1;
    V3 oL = o, dL = d, oTri = o;
    int32_t lstack[RT_FLAT_STACK];
    int sp = 0;
    const int tid = threadIdx.x;
#define RT_FLAT_PUSH(v) do { if (sp < sstackN) sstack[sp * RT_BLOCK + tid] = (v); else lstack[sp - sstackN] = (v); sp++; } while (0)
    for (;;) {
        if ((uint32_t)cur < (uint32_t)kReturnRef) {            // an inner node of either level
            float n0, n1;
            bool h0, h1;
            int2 ch;
            twoBoxes(base + cur, a, n0, h0, n1, h1, ch);
            if (COUNT) C.box += 2;
            if (h0 && h1) {
                int32_t nearC = ch.x, farC = ch.y;
                if (n1 < n0) { nearC = ch.y; farC = ch.x; }
                RT_FLAT_PUSH(farC);
                if (COUNT && sp > C.maxStack) C.maxStack = sp;
                cur = nearC;
                continue;
            }
            if (h0) { cur = ch.x; continue; }
            if (h1) { cur = ch.y; continue; }
        } else if (cur == kReturnRef) {                       // the instance is done: back to the TLAS
            inBlas = false;
            base = S.tlas;
            top.tmax = __double2float_ru(hit.t - top.t0);       // a closer hit prunes the boxes that follow
            a = top;
        } else if (cur != kEmptyRef) {
            if (inBlas) {                                       // triangles of a leaf
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
                    a.tmax = __double2float_ru(t - a.t0);
                }
            } else {                                            // an instance (RTContext.swift:651-704)
                const int idx = ~cur;
                const GInstance& I = S.instances[idx];
                toLocal<MOTION>(I, o, d, oL, dL, time);
                if (COUNT) C.inst++;
                if (I.kind == INST_MESH) {
                    const RayF b = makeRayF(I.rootLo, I.rootHi, I.centre, oL, dL, refinedInv(dL), tLow, hit.t);
                    if (COUNT) C.box++;
                    if (b.hits) {
                        oTri = MOTION ? oL - ld3(I.meshMotion) * time : oL;
                        instIdx = idx;
                        a = b;
                        base = S.nodes;
                        inBlas = true;
                        RT_FLAT_PUSH(kReturnRef);
                        cur = I.rootRef;
                        continue;
                    }
                } else {
                    if (analyticInstance<COUNT>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, C)) return true;
                    top.tmax = __double2float_ru(hit.t - top.t0);
                    a = top;
                }
            }
        }
        if (sp == 0) return false;
        sp--;
        cur = sp < sstackN ? sstack[sp * RT_BLOCK + tid] : lstack[sp - sstackN];
    }
#undef RT_FLAT_PUSH
}

#ifdef RT_EMU
#define RT_NOINLINE __attribute__((noinline))
#else
#define RT_NOINLINE __noinline__
#endif

struct Surface {
    V3 p, n;          // world-space hit point and geometric normal (Hit.p, Hit.n)
    int64_t mat;      // Hit.mat
    int64_t prim;     // primitiveIndex of the hit primitive
    int32_t kind;     // RT_HIT_*
};

// Hit.p / Hit.n / Hit.mat exactly as committed at RTContext.swift:496-509,524-526,535-537
// and mapped to world space at :679-704.
template <bool MOTION = false>
__device__ __forceinline__ void resolveHit(const DeviceScene& S, const HitRec& hit, V3 o, V3 d, Surface& s, double time = 0.0) {
    const GInstance& I = S.instances[hit.inst];
    V3 pL, nL;
    const int kind = I.kind;
    if (kind == INST_MESH || kind == INST_TRI) {
        V3 v0, v1, v2;
        loadTri(S, I.triF64 != 0, hit.slot, v0, v1, v2);
        int64_t triIndex = I.triF64 ? S.trisD[hit.slot].triIndex : (int64_t)S.tris[hit.slot].triIndex;
        V3 e1 = v1 - v0, e2 = v2 - v0;
        const double u = hit.u, v = hit.v;
        if (I.smooth) {
            const double* nn = S.triNormals + triIndex * 9;
            const double w = 1.0 - u - v;
            nL = normalize((w * ld3(nn) + u * ld3(nn + 3)) + v * ld3(nn + 6));
        } else {
            nL = normalize(cross(e1, e2));
        }
        pL = (v0 + u * e1) + v * e2;
        s.kind = 1;  // RT_HIT_TRIANGLE
        s.prim = triIndex;
    } else {
        V3 oL, dL;
        toLocal<MOTION>(I, o, d, oL, dL, time);
        pL = oL + dL * hit.t;
        if (kind == INST_SPHERE) { nL = normalize(pL - ld3(I.prim)); s.kind = 2; }
        else { nL = normalize(ld3(I.prim + 3)); s.kind = 3; }
        s.prim = I.primIndex;
    }
    s.p = I.identity ? pL : xfmPoint(I.l2w, pL);
    if (MOTION) s.p = s.p + ld3(I.motion) * time;       // wp + instOffset (:681-701); the mesh's own motionBlur is not added (:504)
    V3 wn = normalize(mul3(I.nmat, nL));
    if (I.flipNormal) wn = -wn;
    s.n = wn;
    s.mat = I.matId;
}

}  // namespace rtg
