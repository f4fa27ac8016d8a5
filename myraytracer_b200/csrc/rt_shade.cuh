// rt_shade.cuh — the trace closure of the reference as an explicit depth-first loop (device).
//
//   trace (shading + recursion)            Extensions/Object+Extension.swift:96-283
//   Fresnel / reflect / Beer helpers       :467-505, orthonormalBasis :530-552
// The pieces of `trace` (surface set-up, one point / area light, child-ray spawning) are shared by the two kernels: the
// general kernel walks every lane's path tree depth-first with an explicit per-thread stack of (ray, throughput)
// entries, reflection before refraction, which keeps the reference's RNG draw order (scalarPathLoop); the fast kernel
// sends child rays through the warp's ray queue (WarpQueue).
#pragma once
#include <math_constants.h>

#include "rt_kernels.h"
#include "rt_packet.cuh"

namespace rtg {

#define PATH_FIELDS 11

struct PathEntry {
    V3 o, d, w;
    long long beerMat;   // Hit.mat of the dielectric surface this refraction ray entered
    int depth;
    int slot;            // fast kernel: the lane (sample of the work unit) this path's light is added to
    int beer;            // bit 0: apply Beer attenuation at the next hit (Object+Extension.swift:245-248)
                         // bit 1 (PATH_DEAD): an ancestor's Lo was non-finite — the reference still walks this subtree
                         // (rays, RNG draws, jitter index) and then throws the level away (:277-280): nothing is added
};
#define PATH_BEER 1
#define PATH_DEAD 2

__device__ __forceinline__ void pathStore(double* base, size_t stride, size_t gtid, int level, const PathEntry& e) {
    double* p = base + ((size_t)level * PATH_FIELDS) * stride + gtid;
    p[0 * stride] = e.o.x; p[1 * stride] = e.o.y; p[2 * stride] = e.o.z;
    p[3 * stride] = e.d.x; p[4 * stride] = e.d.y; p[5 * stride] = e.d.z;
    p[6 * stride] = e.w.x; p[7 * stride] = e.w.y; p[8 * stride] = e.w.z;
    p[9 * stride] = __longlong_as_double(e.beerMat);
    p[10 * stride] = __longlong_as_double((long long)(e.depth & 0xffff) | ((long long)(e.slot & 0xffff) << 16) | ((long long)e.beer << 32));
}
__device__ __forceinline__ void pathLoad(const double* base, size_t stride, size_t gtid, int level, PathEntry& e) {
    const double* p = base + ((size_t)level * PATH_FIELDS) * stride + gtid;
    e.o = mk(p[0 * stride], p[1 * stride], p[2 * stride]);
    e.d = mk(p[3 * stride], p[4 * stride], p[5 * stride]);
    e.w = mk(p[6 * stride], p[7 * stride], p[8 * stride]);
    e.beerMat = __double_as_longlong(p[9 * stride]);
    long long m = __double_as_longlong(p[10 * stride]);
    e.depth = (int)(m & 0xffffLL);
    e.slot = (int)((m >> 16) & 0xffffLL);
    e.beer = (int)(m >> 32);
}

__device__ __forceinline__ V3 reflectDir(V3 d, V3 n) { return d - 2.0 * dot(d, n) * n; }  // :467

// :530-552
__device__ __forceinline__ void orthonormalBasis(V3 n, V3& tangent, V3& bitangent) {
    double sign = n.z >= 0 ? 1.0 : -1.0;
    double a = ddiv(-1.0, sign + n.z);
    double b = n.x * n.y * a;
    tangent = normalize(mk(1.0 + sign * n.x * n.x * a, sign * b, -sign * n.x));
    bitangent = normalize(mk(b, sign + n.y * n.y * a, -n.y));
}

// :476-490
__device__ __forceinline__ void fresnelDielectric(double n1, double n2, double cosI, double& R, bool& hasCosT, double& cosT,
                                                  double& sin2T) {
    double cosTheta = dmax(0.0, dmin(1.0, fabs(cosI)));
    double eta = ddiv(n1, n2);
    sin2T = eta * eta * dmax(0.0, 1.0 - cosTheta * cosTheta);
    if (sin2T > 1.0) { R = 1.0; hasCosT = false; cosT = 0; return; }
    double cosPhi = dsqrt(dmax(0.0, 1.0 - sin2T));
    double Rs = ddiv(n2 * cosTheta - n1 * cosPhi, n2 * cosTheta + n1 * cosPhi);
    double Rp = ddiv(n1 * cosTheta - n2 * cosPhi, n1 * cosTheta + n2 * cosPhi);
    R = 0.5 * (Rs * Rs + Rp * Rp);
    hasCosT = true; cosT = cosPhi;
}

// :492-505
__device__ __forceinline__ double fresnelConductor(double eta, double k, double cosI_) {
    double cosI = dmax(0.0, dmin(1.0, fabs(cosI_)));
    double cos2 = cosI * cosI;
    double eta2k2 = eta * eta + k * k;
    double twoEtaCos = 2.0 * eta * cosI;
    double Rs = ddiv(eta2k2 - twoEtaCos + cos2, eta2k2 + twoEtaCos + cos2);
    double Rp = ddiv((eta2k2 * cos2) - twoEtaCos + 1.0, (eta2k2 * cos2) + twoEtaCos + 1.0);
    return 0.5 * (Rs + Rp);
}

__device__ __forceinline__ unsigned char quantize(double c) {  // RayTracer.swift:189-193
    double x = dmin(dmax(c, 0.0), 255.0);
    return (unsigned char)(int)x;
}


struct RayCounts {
    unsigned long long primary, shadow, secondary;
};

struct TraceScratch {
    int32_t* sstack;   // per-lane traversal stacks (shared memory, sstackN x RT_BLOCK; deeper entries in local memory)
    int sstackN;
    int2* wstack;      // this warp's packet stack (shared memory, RT_WARP_STACK_INT2)
    double* pathStack; // per-thread recursion stack (global)
    size_t stride, gtid;
};

// What `trace` derives from a hit before lighting it (Object+Extension.swift:105-117).
struct ShadePoint {
    V3 p, N;               // hit point, face-forwarded normal
    V3 w;                  // throughput of this path vertex (after Beer attenuation)
    const GMaterial* mat;
    long long matId;       // Hit.mat
    bool frontFacing, computeDirect;
};

__device__ __forceinline__ void setupShade(const DeviceScene& S, const Surface& sf, const PathEntry& cur, double hitT, ShadePoint& sp) {
    long long mi = sf.mat - 1;                                                          // :105
    if (mi > (long long)S.nMaterials - 1) mi = (long long)S.nMaterials - 1;
    if (mi < 0) mi = 0;
    const GMaterial& mat = S.materials[mi];
    sp.mat = &mat;
    sp.matId = sf.mat;
    V3 w = cur.w;
    if ((cur.beer & PATH_BEER) && sf.mat == cur.beerMat) {                                            // :245-248
        const double dist = dmax(hitT, 0.0);
        if (isfinite(dist) && dist > 0)
            w = w * mk(exp(-mat.absorption[0] * dist), exp(-mat.absorption[1] * dist), exp(-mat.absorption[2] * dist));
    }
    sp.w = w;
    sp.p = sf.p;
    sp.frontFacing = dot(cur.d, sf.n) < 0;                                               // :109
    sp.N = sp.frontFacing ? sf.n : -sf.n;
    sp.computeDirect = !(mat.ior > 0) || sp.frontFacing;                                // :112-115
}

// Shadow ray towards one point light (:120-124): direction, distance (= tMax), offset origin.
__device__ __forceinline__ void shadowRay(const GPointLight& pl, V3 p, double shadowEps, V3& org, V3& wi, double& dist) {
    wi = ld3(pl.position) - p;
    dist = length(wi);
    wi = normalize(wi);
    org = p + wi * shadowEps;
}

// Blinn-Phong term of one visible point light (:126-140).
__device__ __forceinline__ V3 pointLightTerm(const GMaterial& mat, const GPointLight& pl, V3 N, V3 rayDir, V3 wi, double dist) {
    const double NdotL = dmax(0.0, dot(N, wi));
    if (!(NdotL > 0)) return mk(0, 0, 0);
    const double shininess = dmax(1.0, mat.phong);
    const V3 Ld = ld3(mat.diffuse) * NdotL;
    const V3 view = normalize(-rayDir);
    const V3 h = normalize(wi + view);
    const double NdotH = dmax(0.0, dot(N, h));
    const V3 Ls = ld3(mat.specular) * dpow(NdotH, shininess);
    const V3 atten = ld3(pl.intensity) / dmax(dist * dist, 1e-12);
    return (Ld + Ls) * atten;
}

// Area lights (:145-186).  The light sample comes from the 10x10 jitter table through the chunk-wide running index
// (see RenderTargets::areaIndex); countOnly: the index is advanced and nothing else happens (pass 1).
struct AreaState {
    unsigned jitterIndex;     // kept mod 100: only `jitterIndex % 100` is ever used (:152-153)
    bool countOnly;
};

// The sample point of one area light for jitter entry k, and everything of its contribution that does not need
// the shadow ray: returns false when the light is behind the surface or edge-on (:166-171).
__device__ __forceinline__ bool areaLightSample(const DeviceScene& S, const GAreaLight& al, unsigned k, V3 p, V3 N, V3& nL, V3& wi,
                                                double& dist, double& dist2, double& NdotL, double& Ln) {
    nL = normalize(ld3(al.normal));
    V3 t, b;
    orthonormalBasis(nL, t, b);
    const double size = al.size;
    const double r1 = ddiv(S.jitter[k], 10.0) - 0.5;
    const double r2 = ddiv(S.jitter[100 + k], 10.0) - 0.5;
    const V3 samplePos = (ld3(al.position) + t * (r1 * size)) + b * (r2 * size);
    wi = samplePos - p;
    dist2 = dot(wi, wi);
    dist = dsqrt(dist2);
    wi = wi / dist;
    NdotL = dot(N, wi);
    if (NdotL <= 0) return false;
    Ln = fabs(dot(nL, -wi));
    if (Ln <= 0) return false;
    return true;
}
__device__ __forceinline__ V3 areaLightTerm(const GMaterial& mat, const GAreaLight& al, V3 N, V3 rayDir, V3 wi, double dist2,
                                            double NdotL, double Ln) {
    const double area = al.size * al.size;
    const V3 view = normalize(-rayDir);
    const V3 h = normalize(wi + view);
    const V3 Ld = ld3(mat.diffuse) * NdotL;
    const V3 Ls = ld3(mat.specular) * dpow(dmax(0.0, dot(N, h)), mat.phong);       // raw phong exponent (:179)
    const V3 brdf = Ld + Ls;
    return ((brdf * ld3(al.radiance)) * (Ln / dist2)) * area;
}

// Recursive branches of `trace` (:189-275): fills c1 (always) and c2 (refraction) and returns the child count.
// GLOSSY = false compiles the roughness perturbations (and their RNG draws) out: the fast kernel runs only on
// scenes where every roughness is 0, so its RNG state is dead after primary-ray generation.
template <bool GLOSSY>
__device__ __forceinline__ int spawnChildren(const DeviceScene& S, const ShadePoint& sp, const PathEntry& cur, PCG32& rng,
                                             PathEntry& c1, PathEntry& c2) {
    const GMaterial& mat = *sp.mat;
    if (!(cur.depth < S.maxDepth) || mat.type == 0) return 0;
    const double shadowEps = S.shadowEps;
    const V3 p = sp.p, N = sp.N, w = sp.w;
    int count = 1;
    c1.depth = c2.depth = cur.depth + 1;
    c1.slot = c2.slot = cur.slot;
    c1.beer = c2.beer = 0; c1.beerMat = c2.beerMat = 0;
    if (mat.type == 1) {                                                                // mirror :189-206
        V3 rd = normalize(reflectDir(cur.d, N));
        if (GLOSSY && mat.roughness != 0) {
            V3 t, b;
            orthonormalBasis(rd, t, b);
            const double rand1 = rng.nextFloat() - 0.5;
            const double rand2 = rng.nextFloat() - 0.5;
            rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
            rd = normalize(rd);
        }
        c1.o = p + N * shadowEps; c1.d = rd; c1.w = w * ld3(mat.mirror);
    } else if (mat.type == 2) {                                                         // dielectric :207-251
        V3 mN = N;
        if (GLOSSY && mat.roughness != 0) {
            V3 t, b;
            orthonormalBasis(N, t, b);
            const double r1 = (rng.nextFloat() * 2.0) - 1.0;
            const double r2 = (rng.nextFloat() * 2.0) - 1.0;
            mN = normalize(N + (t * r1 * mat.roughness) + (b * r2 * mat.roughness));
        }
        const bool entering = sp.frontFacing;
        const double n1 = entering ? 1.0 : mat.ior;
        const double n2 = entering ? mat.ior : 1.0;
        const double cosI = -dot(cur.d, mN);
        double Rf, cosT, sin2T;
        bool hasCosT;
        fresnelDielectric(n1, n2, cosI, Rf, hasCosT, cosT, sin2T);
        const V3 rd = normalize(reflectDir(cur.d, mN));
        c1.o = p + rd * shadowEps; c1.d = rd;
        if (!hasCosT || sin2T > 1) {
            c1.w = w;                                                                   // total internal reflection :235-236
        } else {
            c1.w = w * Rf;
            const double eta = ddiv(n1, n2);
            const V3 td = normalize((cur.d * eta) + (mN * (eta * cosI - cosT)));
            c2.o = p + td * shadowEps; c2.d = td; c2.w = w * (1.0 - Rf);
            c2.beer = (entering && mat.hasAbsorption) ? 1 : 0;
            c2.beerMat = sp.matId;
            count = 2;
        }
    } else {                                                                            // conductor :252-275
        const double cosI = dmax(0.0, -dot(cur.d, N));
        const double Rf = fresnelConductor(mat.ior, mat.absorptionIndex, cosI);
        V3 rd = normalize(reflectDir(cur.d, N));
        if (GLOSSY && mat.roughness != 0) {
            V3 t, b;
            orthonormalBasis(rd, t, b);
            const double rand1 = rng.nextFloat() - 0.5;
            const double rand2 = rng.nextFloat() - 0.5;
            rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
            rd = normalize(rd);
        }
        c1.o = p + N * shadowEps; c1.d = rd; c1.w = w * (mk(Rf, Rf, Rf) * ld3(mat.mirror));
    }
    return count;
}

// the general kernel's scene walk: the nested loops of rt_trace.cuh (-DRT_GENERAL_WALK=traceSceneFlat: both levels in
// one loop — measured and dropped, see the comment there)
#ifndef RT_GENERAL_WALK
#define RT_GENERAL_WALK traceScene
#endif

// Depth-first processing of path entries with the per-lane traversal: `cur` is the entry to start with and
// `level` the number of entries already waiting on this thread's path stack.  firstIsPrimary: `cur` is the
// camera ray (tMin = near-plane clip :341-344, hit id recorded).  Returns `color` plus the w * Lo of all
// processed entries, added in processing order.
// AREA: the scene has area lights; `area` carries the chunk's jitter index through the path (general kernel only).
// MOTION: motion blur; `time` is the sample's ray time, inherited by every ray of the path (:123,172,198,242).
template <bool GLOSSY, bool COUNT, bool AREA = false, bool MOTION = false>
__device__ __forceinline__ V3 scalarPathLoop(const DeviceScene& S, const TraceScratch& T, PathEntry cur, int level,
                                             bool firstIsPrimary, double tMinPrimary, HitId* idSlot, PCG32& rng, Counters& C,
                                             RayCounts& R, V3 color, AreaState* area = nullptr, double time = 0.0) {
    const V3 ambient = ld3(S.ambient), background = ld3(S.background);
    bool primary = firstIsPrimary;
    for (;;) {
        HitRec hit;
        hit.t = CUDART_INF; hit.u = hit.v = 0; hit.inst = -1; hit.slot = 0;
        RT_GENERAL_WALK<COUNT, MOTION>(S, cur.o, cur.d, primary ? tMinPrimary : 0.0, hit, false, T.sstack, T.sstackN, C, time);
        Surface sf;
        if (hit.inst >= 0) {
            resolveHit<MOTION>(S, hit, cur.o, cur.d, sf, time);
            if (COUNT && sf.kind == 1 && S.instances[hit.inst].smooth) C.smooth++;
        }
        if (primary && idSlot) {
            idSlot->inst = hit.inst;
            idSlot->kind = hit.inst >= 0 ? sf.kind : 0;
            idSlot->prim = hit.inst >= 0 ? sf.prim : -1;
            idSlot->t = hit.t;
        }
        primary = false;
        int spawned = 0;
        if (hit.inst < 0) {
            if (!(cur.beer & PATH_DEAD)) color = color + cur.w * background;            // :101-103
        } else {
            ShadePoint sp;
            setupShade(S, sf, cur, hit.t, sp);
            V3 Lo = sp.computeDirect ? ambient * ld3(sp.mat->ambient) : mk(0, 0, 0);
            const bool shade = !(AREA && area->countOnly);      // pass 1 of an area-light frame only follows the paths
            if (sp.computeDirect && shade) {
                for (int l = 0; l < S.nPointLights; l++) {                             // :118-143
                    const GPointLight& pl = S.pointLights[l];
                    V3 so, wi;
                    double dist;
                    shadowRay(pl, sp.p, S.shadowEps, so, wi, dist);
                    HitRec sh;
                    sh.t = dist; sh.u = sh.v = 0; sh.inst = -1; sh.slot = 0;             // tMax = dist (:124)
                    R.shadow++;
                    const bool occluded = RT_GENERAL_WALK<COUNT, MOTION>(S, so, wi, 0.0, sh, true, T.sstack, T.sstackN, C, time);
                    if (!occluded)
                        Lo = Lo + pointLightTerm(*sp.mat, pl, sp.N, cur.d, wi, dist);
                }
            }
            if (AREA && sp.computeDirect) {
                for (int l = 0; l < S.nAreaLights; l++) {                               // :145-186
                    const unsigned k = area->jitterIndex;
                    area->jitterIndex = k + 1 >= 100u ? 0u : k + 1;                     // jitterIndex &+= 1, kept mod 100
                    if (!shade) continue;
                    const GAreaLight& al = S.areaLights[l];
                    V3 nL, wi;
                    double dist, dist2, NdotL, Ln;
                    if (!areaLightSample(S, al, k, sp.p, sp.N, nL, wi, dist, dist2, NdotL, Ln)) continue;
                    HitRec sh;
                    sh.t = dist - S.shadowEps; sh.u = sh.v = 0; sh.inst = -1; sh.slot = 0;   // tMax = dist - shadowEps (:173)
                    R.shadow++;
                    const V3 so = sp.p + wi * S.shadowEps;
                    const bool occluded = RT_GENERAL_WALK<COUNT, MOTION>(S, so, wi, 0.0, sh, true, T.sstack, T.sstackN, C, time);
                    if (!occluded) Lo = Lo + areaLightTerm(*sp.mat, al, sp.N, cur.d, wi, dist2, NdotL, Ln);
                }
            }
            // NaN guard (:277-280): the reference traces the children first (their rays, RNG draws and jitter-index
            // steps happen) and then returns black for a level whose Lo is not finite.  A child's own level returns a
            // finite value or black, so the level's total is non-finite exactly when its direct term is: such a level
            // and everything below it is walked with PATH_DEAD set and adds nothing.
            const bool dead = (cur.beer & PATH_DEAD) != 0 || !finite3(Lo);
            if (!dead) color = color + sp.w * Lo;
            PathEntry c1, c2;
            const int nc = spawnChildren<GLOSSY>(S, sp, cur, rng, c1, c2);
            if (dead) { c1.beer |= PATH_DEAD; c2.beer |= PATH_DEAD; }
            if (nc == 2) { pathStore(T.pathStack, T.stride, T.gtid, level, c2); level++; }
            if (nc > 0) { cur = c1; spawned = 1; R.secondary += nc; }
        }
        if (!spawned) {
            if (level == 0) break;
            level--;
            pathLoad(T.pathStack, T.stride, T.gtid, level, cur);
        }
    }
    return color;
}


// ---- secondary rays of the fast kernel: a wavefront inside the warp ---------------------------------------------------
// The recursion of `trace` (Object+Extension.swift:189-275) as a ray queue shared by the 32 lanes of a warp.  Run per
// lane (scalarPathLoop: every lane walks its own path tree depth-first with the per-lane traversal) the child rays kept
// 2.5 of 32 lanes busy in the mesh walk of the recursive scene C3, through a second copy of every traversal routine
// that thrashed the instruction cache.  Here the children of a batch of samples are compacted into ONE last-in-first-out
// queue per warp (ballot + prefix popcount), and the warp takes them back 32 at a time: a popped batch is processed
// by exactly the code that processes a pass of primary rays — closest-hit warp packet, shading, one any-hit packet per
// light, children back into the queue (rt_render_kernel.cuh) — so the kernel holds one traversal, the rays of a batch
// meet every instance together (the packet's TLAS walk is warp-uniform: "sorted by instance" for free), and neighbouring
// lanes' children, pushed side by side, come back as one fairly coherent packet.
//   * Each entry carries the sample (`slot`) it belongs to; light is added to col[slot] in shared memory.  Lanes that
//     add to one slot in the same round are ordered by lane (match_any + rank), so the sum is deterministic; it is
//     formed in queue order, not in the reference's depth-first order: the addends are the same FP64 values, the
//     reordering moves a sample by a few ulp, i.e. an 8-bit pixel by at most 1 LSB and only if it lies within ~1e-13
//     of a truncation boundary.
//   * No non-finite guard here (:277-280): with finite lights and materials every term is finite (a NaN normal or
//     light direction fails `NdotL > 0`), and the host sends scenes with a non-finite constant to the general kernel.
//   * Queue memory: the warp's 32 columns of the path stack, levels [0, D + RT_QUEUE_EXTRA): a batch starts with at
//     most 64 entries per pass, then at most 32 entries leave and 64 enter per round, deepest first.  Should it ever
//     be full, the round's children go to the lanes' private stacks (the levels behind), which a lane drains before it
//     takes shared entries: bounded by D + 2 per lane like any depth-first walk.
#define RT_QUEUE_EXTRA (3 + 2 * (RT_BATCH_PASSES - 1))   // initial children of a batch: 64 per pass
#define RT_PRIV_EXTRA 2

__device__ __forceinline__ double* queueField(double* qbase, size_t stride, int e, int f) {
    return qbase + ((size_t)(e >> 5) * PATH_FIELDS + (size_t)f) * stride + (size_t)(e & 31);
}
__device__ __forceinline__ void queueStore(double* qbase, size_t stride, int e, const PathEntry& v) {
    *queueField(qbase, stride, e, 0) = v.o.x; *queueField(qbase, stride, e, 1) = v.o.y; *queueField(qbase, stride, e, 2) = v.o.z;
    *queueField(qbase, stride, e, 3) = v.d.x; *queueField(qbase, stride, e, 4) = v.d.y; *queueField(qbase, stride, e, 5) = v.d.z;
    *queueField(qbase, stride, e, 6) = v.w.x; *queueField(qbase, stride, e, 7) = v.w.y; *queueField(qbase, stride, e, 8) = v.w.z;
    *queueField(qbase, stride, e, 9) = __longlong_as_double(v.beerMat);
    *queueField(qbase, stride, e, 10) = __longlong_as_double((long long)(v.depth & 0xffff) | ((long long)(v.slot & 0xffff) << 16) | ((long long)v.beer << 32));
}
__device__ __forceinline__ void queueLoad(double* qbase, size_t stride, int e, PathEntry& v) {
    v.o = mk(*queueField(qbase, stride, e, 0), *queueField(qbase, stride, e, 1), *queueField(qbase, stride, e, 2));
    v.d = mk(*queueField(qbase, stride, e, 3), *queueField(qbase, stride, e, 4), *queueField(qbase, stride, e, 5));
    v.w = mk(*queueField(qbase, stride, e, 6), *queueField(qbase, stride, e, 7), *queueField(qbase, stride, e, 8));
    v.beerMat = __double_as_longlong(*queueField(qbase, stride, e, 9));
    const long long m = __double_as_longlong(*queueField(qbase, stride, e, 10));
    v.depth = (int)(m & 0xffffLL); v.slot = (int)((m >> 16) & 0xffffLL); v.beer = (int)(m >> 32);
}

// col[slot] += v for the lanes with `has`, in lane order where several lanes name one slot.  All 32 lanes call it.
__device__ __forceinline__ void warpAccumulate(double* col, int slot, bool has, V3 v) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned any = __ballot_sync(RT_FULL, has);
    if (!any) return;
    const unsigned peers = __match_any_sync(RT_FULL, has ? slot : -1 - (int)lane);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    for (int r = 0;; r++) {
        if (__ballot_sync(RT_FULL, has && rank >= r) == 0u) break;
        if (has && rank == r) {
            double* c = col + slot * 3;
            c[0] = c[0] + v.x; c[1] = c[1] + v.y; c[2] = c[2] + v.z;
        }
        __syncwarp();
    }
}

// Only `top` and `priv` live across a round; where the queue lies is re-derived from the launch geometry at every push
// and pop (uniform arithmetic), so the packet walk in between does not carry the queue in registers.
struct WarpQueue {
    int top;             // entries in the shared queue (warp-uniform)
    int priv;            // entries on this lane's private stack

    __device__ __forceinline__ void init() { top = 0; priv = 0; }

    // Offer n (0..2) entries per lane; all 32 lanes call it.  Compaction: ballot + prefix popcount.
    __device__ __forceinline__ void push(double* pathStack, int maxDepth, int n, const PathEntry& a, const PathEntry& b) {
        const unsigned m1 = __ballot_sync(RT_FULL, n >= 1), m2 = __ballot_sync(RT_FULL, n == 2);
        const int total = __popc(m1) + __popc(m2);
        if (total == 0) return;
        const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
        const size_t stride = (size_t)gridDim.x * blockDim.x, gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (top + total <= 32 * (maxDepth + RT_QUEUE_EXTRA)) {
            double* qbase = pathStack + (gtid - lane);
            const int at = top + __popc(m1 & lt) + __popc(m2 & lt);
            if (n == 2) queueStore(qbase, stride, at, b);            // refraction below, reflection on top of it
            if (n >= 1) queueStore(qbase, stride, at + (n == 2 ? 1 : 0), a);
            top += total;
        } else {
            const int privBase = maxDepth + RT_QUEUE_EXTRA;
            if (n == 2) { pathStore(pathStack, stride, gtid, privBase + priv, b); priv++; }
            if (n >= 1) { pathStore(pathStack, stride, gtid, privBase + priv, a); priv++; }
        }
        __syncwarp();
    }

    // Up to 32 entries off the top, one per lane (a lane's own spill-over first); all 32 lanes call it.  Returns false
    // when nothing is left anywhere in the warp.
    __device__ __forceinline__ bool pop(double* pathStack, int maxDepth, PathEntry& cur, bool& valid) {
        valid = false;
        if (top == 0 && __ballot_sync(RT_FULL, priv > 0) == 0u) return false;
        const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
        const size_t stride = (size_t)gridDim.x * blockDim.x, gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (priv > 0) { priv--; pathLoad(pathStack, stride, gtid, maxDepth + RT_QUEUE_EXTRA + priv, cur); valid = true; }
        const unsigned need = __ballot_sync(RT_FULL, !valid);
        if (need != 0u && top > 0) {
            const int want = __popc(need), k = want < top ? want : top;
            const int r = __popc(need & lt);
            if (!valid && r < k) { queueLoad(pathStack + (gtid - lane), stride, top - 1 - r, cur); valid = true; }
            top -= k;
        }
        return __ballot_sync(RT_FULL, valid) != 0u;
    }
};

}  // namespace rtg
