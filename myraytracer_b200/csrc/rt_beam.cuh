// rt_beam.cuh — beam walk of a wide BVH: a coherent packet's inner-node tests with ONE LANE PER CHILD BOX.
//
// Same queries as packetBlas (rt_packet.cuh; intersectBLAS / occludedBLAS, Models/RTContext.swift:544-610,786-829)
// for the packets that are coherent — the 32 primary rays of one or two neighbouring pixels, or their shadow rays
// towards one light.  Such a packet walks a single root-to-leaf path of the binary BVH (C4: 20 node visits per
// packet, the depth of the tree), and in the packet walk every one of those visits costs the warp ~50 instructions
// in which 32 lanes test the same two boxes.  Here the parallelism is turned round:
//   * the packet is bounded ONCE by a beam — dominant axis K, travel coordinate u = sigma * x_K measured from the
//     rearmost ray origin, and for the two other axes the interval of the rays' positions as a linear function of u,
//     [cmin + rmin u, cmax + rmax u] (u >= 0).  All of it is warp-uniform FP32, rounded outward;
//   * the tree is the binary tree collapsed to kWide children per node (gpu_layout.h: WNode, bvh_build.cpp:
//     collapseWide), and lane r tests child perm[K, sigma][r] against the beam: the slab of the box along K clipped to
//     the beam's range gives [sa, sb], and the beam's extent over [sa, sb] must overlap the box on the other two axes.
//     One visit decides kWide children — a 16-ary tree over a million triangles is 5 to 6 levels deep;
//   * the children that pass are pushed in front-to-back order (the permutation sorted them by their near face along
//     the travel direction) with one ballot + popcount; inner entries carry the travel coordinate of their near face
//     and are skipped when popped behind the farthest closest hit of the packet;
//   * a LEAF entry is tested per ray, exactly as before: the conservative FP32 slab test of rt_packet.cuh against the
//     leaf's box gives the lane mask, then the FP64 Moeller-Trumbore tests in the reference's operation order.  What
//     decides a result is unchanged; the beam only culls, and it culls conservatively (it contains every ray of the
//     packet, directed rounding everywhere), so no hit the reference finds is lost and t, u, v stay bit-identical.
// A packet that cannot be bounded this way (dominant direction components of both signs, slopes above kBeamSlope,
// non-finite values) returns formed = false and takes packetBlas; so do incoherent packets (child rays off the ray
// queue) and meshes without a wide tree.
//
// Every function here must be entered by all 32 lanes of a converged warp.
#pragma once
#if defined(RT_BEAM_STATS) && defined(RT_EMU)
#include <cstdio>
#endif

namespace rtg {

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7fffffff); }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }
// min / max of a float over the warp: one REDUX on an order-preserving integer image
__device__ __forceinline__ float warpMinF(float v) { return ord2f(__reduce_min_sync(RT_FULL, f2ord(v))); }
__device__ __forceinline__ float warpMaxF(float v) { return ord2f(__reduce_max_sync(RT_FULL, f2ord(v))); }

static const int kWideFloats = (int)(sizeof(WNode) / 4);     // node stride in floats (leaf entries address their box by float index)
#define RT_BEAM_SLOPE 4.0     // |d_other| <= slope * |d_K| for every ray of a beam
// A beam is the hull of its rays: a packet whose rays split into bundles far apart (primary hits on both sides of a
// silhouette, then their shadow rays) bounds the space between the bundles too.  The walk therefore carries a budget
// of node visits + leaf entries; a packet that spends it gives up (formed = false with what it found so far — closest
// hits stand, occluded lanes are returned) and the caller finishes with the packet walk, whose lane masks follow
// the bundles separately.  A coherent packet needs ~10 visits and ~5 leaves on the million-triangle terrain.
#ifndef RT_BEAM_BUDGET
#define RT_BEAM_BUDGET 96
#endif

#if defined(RT_BEAM_STATS) && defined(RT_EMU)
struct BeamStats {
    unsigned long long calls = 0, notFormed = 0, bails = 0, visits = 0, leaves = 0, maxVisits = 0;
    ~BeamStats() { fprintf(stderr, "BEAMSTATS calls %llu notFormed %llu bails %llu visits %llu leaves %llu\n", calls, notFormed, bails, visits, leaves); }
};
static BeamStats g_beamStats;
#define RT_BS(x) do { if (lane == 0) { g_beamStats.x; } } while (0)
#else
#define RT_BS(x) do { } while (0)
#endif

template <bool COUNT>
__device__ __forceinline__ uint32_t beamBlas(const DeviceScene& S, const GInstance& I, int instIdx, V3 oL, V3 dL, double eps,
                                             double tLow, HitRec& hit, bool anyhit, uint32_t inMask, int2* wstack, Counters& C,
                                             bool& formed) {
    const int lane = threadIdx.x & 31;
    const uint32_t bit = 1u << lane;
    formed = true;
    // ---- per-lane root test and FP32 view of the ray: exactly packetBlas
    const V3 ia = refinedInv(dL);
    const V3 oc = oL - ld3(I.centre);
    const double lo[3] = {I.rootLo[0] - I.centre[0], I.rootLo[1] - I.centre[1], I.rootLo[2] - I.centre[2]};
    const double hi[3] = {I.rootHi[0] - I.centre[0], I.rootHi[1] - I.centre[1], I.rootHi[2] - I.centre[2]};
    double tn, tf;
    slabApprox(lo, hi, oc, ia, tn, tf);
    const bool in = (inMask & bit) != 0;
    const bool alive = in && !(tn > tf || tf < tLow || tn > hit.t);
    uint32_t live = __ballot_sync(RT_FULL, alive);
    if (!live) { if (COUNT && in) C.box++; return 0u; }

    // ---- the beam's frame: K = dominant axis of the first ray, A and B the two others (cyclic)
    const int leader = __ffs((int)live) - 1;
    const double ax = fabs(dL.x), ay = fabs(dL.y), az = fabs(dL.z);
    const int kMine = ax >= ay ? (ax >= az ? 0 : 2) : (ay >= az ? 1 : 2);
    const int k = __shfl_sync(RT_FULL, kMine, leader);
    const double dK = k == 0 ? dL.x : k == 1 ? dL.y : dL.z;
    const double dA = k == 0 ? dL.y : k == 1 ? dL.z : dL.x;
    const double dB = k == 0 ? dL.z : k == 1 ? dL.x : dL.y;
    const double oK = k == 0 ? oc.x : k == 1 ? oc.y : oc.z;
    const double oA = k == 0 ? oc.y : k == 1 ? oc.z : oc.x;
    const double oB = k == 0 ? oc.z : k == 1 ? oc.x : oc.y;
    const int neg = __shfl_sync(RT_FULL, dK < 0.0 ? 1 : 0, leader);
    const double dKa = neg ? -dK : dK;
    const bool okLane = !alive || (dKa > 0.0 && fabs(dA) <= RT_BEAM_SLOPE * dKa && fabs(dB) <= RT_BEAM_SLOPE * dKa);
    RT_BS(calls++);
    if (__ballot_sync(RT_FULL, !okLane)) { RT_BS(notFormed++); formed = false; return 0u; }

    // ---- per ray: position on A, B as a function of the travel coordinate u = sigma * x_K - refU (u >= u0 >= 0):
    //      p(u) = c + r u.  The beam is the rays' extent on two planes — u = 0 (the rearmost origin) and u = U1 (the
    //      far end: the farthest closest hit / shadow-ray end, never beyond the exact root box) — joined by straight
    //      lines: a ray lies between the lines through the minima and through the maxima for every u in [0, U1], and
    //      the hull is tight for bundles that spread out (primary rays) as well as for bundles that converge (shadow
    //      rays towards a point light, which independent intervals of c and r would let fan out instead).
    //      FP64 here (errors ~1e-13 relative, widened by kRel), FP32 rounded outward below.
    const float INF = CUDART_INF_F;
    const double kRel = 1e-12;
    const double inv = refinedInv1(dKa);
    const double rA = dA * inv, rB = dB * inv;
    const double xs = neg ? -oK : oK;
    const float refU = warpMinF(alive ? __double2float_rd(xs) : INF);     // rearmost origin: u0 >= 0 for every ray
    const double u0 = xs - (double)refU;
    const float umin = warpMinF(alive ? __double2float_rd((u0 + dKa * tLow) * (1.0 - kRel)) : INF);
    const double farRoot = neg ? -(k == 0 ? lo[0] : k == 1 ? lo[1] : lo[2]) : (k == 0 ? hi[0] : k == 1 ? hi[1] : hi[2]);
    const float uClip = __double2float_ru((farRoot - (double)refU) + fabs(farRoot) * kRel);
    float umax = fminf(uClip, warpMaxF(alive ? __double2float_ru((u0 + dKa * hit.t) * (1.0 + kRel)) : -INF));
    const float U1 = umax;
    if (!(U1 > 1e-30f && U1 < 1e37f && fabsf(refU) < 1e37f)) { formed = false; return 0u; }      // warp-uniform
    const double cA = oA - rA * u0, cB = oB - rB * u0;
    const double fA = cA + rA * (double)U1, fB = cB + rB * (double)U1;
    const double eA0 = (fabs(oA) + fabs(rA * u0)) * kRel, eB0 = (fabs(oB) + fabs(rB * u0)) * kRel;
    const double eA1 = eA0 + fabs(rA * (double)U1) * kRel, eB1 = eB0 + fabs(rB * (double)U1) * kRel;
    const float cAmin = warpMinF(alive ? __double2float_rd(cA - eA0) : INF), cAmax = warpMaxF(alive ? __double2float_ru(cA + eA0) : -INF);
    const float cBmin = warpMinF(alive ? __double2float_rd(cB - eB0) : INF), cBmax = warpMaxF(alive ? __double2float_ru(cB + eB0) : -INF);
    const float fAmin = warpMinF(alive ? __double2float_rd(fA - eA1) : INF), fAmax = warpMaxF(alive ? __double2float_ru(fA + eA1) : -INF);
    const float fBmin = warpMinF(alive ? __double2float_rd(fB - eB1) : INF), fBmax = warpMaxF(alive ? __double2float_ru(fB + eB1) : -INF);
    // slopes of the lower / upper lines, rounded so that the lines stay outside the hull on all of [0, U1]
    const double invU = refinedInv1((double)U1);
    const double gAl = ((double)fAmin - (double)cAmin) * invU, gAu = ((double)fAmax - (double)cAmax) * invU;
    const double gBl = ((double)fBmin - (double)cBmin) * invU, gBu = ((double)fBmax - (double)cBmax) * invU;
    const float rAmin = __double2float_rd(gAl - fabs(gAl) * kRel), rAmax = __double2float_ru(gAu + fabs(gAu) * kRel);
    const float rBmin = __double2float_rd(gBl - fabs(gBl) * kRel), rBmax = __double2float_ru(gBu + fabs(gBu) * kRel);
    {
        const float chk = ((fabsf(rAmin) + fabsf(rAmax)) + (fabsf(rBmin) + fabsf(rBmax))) + ((fabsf(cAmin) + fabsf(cAmax)) + (fabsf(cBmin) + fabsf(cBmax))) + fabsf(umin);
        if (!(chk < 1e37f)) { formed = false; return 0u; }     // warp-uniform: a value overflowed FP32 (or is NaN)
    }
    if (COUNT && in) C.box++;

    // the lane's own ray against a leaf box: the conservative FP32 slab test of packetBlas
    double t0 = tn > 0.0 ? tn : 0.0;
    if (!(t0 < 1e300)) t0 = 0.0;
    const V3 o2 = oc + dL * t0;
    const float ox = (float)o2.x, oy = (float)o2.y, oz = (float)o2.z;
    const float ix = fminf(fmaxf(1.0f / (float)dL.x, -1e30f), 1e30f);
    const float iy = fminf(fmaxf(1.0f / (float)dL.y, -1e30f), 1e30f);
    const float iz = fminf(fmaxf(1.0f / (float)dL.z, -1e30f), 1e30f);
    const float nx = -(ox * ix), ny = -(oy * iy), nz = -(oz * iz);
    const float tminF = __double2float_rd(tLow - t0);
    float tmaxF = alive ? __double2float_ru(hit.t - t0) : -INF;

    // ---- byte offsets of the fields this beam reads in a node (warp-uniform)
    const int aA = k == 2 ? 0 : k + 1, aB = k == 0 ? 2 : k - 1;       // (k + 1) % 3, (k + 2) % 3
    const int fieldBytes = kWide * 4;
    const int nearOff = (neg ? 3 + k : k) * fieldBytes, farOff = (neg ? k : 3 + k) * fieldBytes;
    const int loAOff = aA * fieldBytes, hiAOff = (3 + aA) * fieldBytes, loBOff = aB * fieldBytes, hiBOff = (3 + aB) * fieldBytes;
    const int childOff = 6 * fieldBytes, permOff = 7 * fieldBytes + (2 * k + neg) * kWide;
    const float sgF = neg ? -1.0f : 1.0f, mref = -refU;
    const bool childLane = lane < kWide;
    const uint32_t below = bit - 1u;

    uint32_t occluded = 0u;
    const uint32_t spBase = (uint32_t)__cvta_generic_to_shared(wstack);
    uint32_t sp = spBase + 8u;
    stsPair(spBase, kEmptyRef, 0);            // sentinel: popping it ends the walk
    int32_t cur = I.wideRoot;
    int budget = RT_BEAM_BUDGET;
    for (;;) {
        if (--budget < 0) { RT_BS(bails++); formed = false; return occluded; }
        RT_BS(visits++);
        // ---- one wide node: lane r tests the r-th child in front-to-back order
        const char* base = reinterpret_cast<const char*>(S.wnodes + cur);
        const int slot = childLane ? (int)__ldg(reinterpret_cast<const unsigned char*>(base + permOff + lane)) : 0;
        const char* p = base + slot * 4;
        const float nearK = __ldg(reinterpret_cast<const float*>(p + nearOff)), farK = __ldg(reinterpret_cast<const float*>(p + farOff));
        const float loA = __ldg(reinterpret_cast<const float*>(p + loAOff)), hiA = __ldg(reinterpret_cast<const float*>(p + hiAOff));
        const float loB = __ldg(reinterpret_cast<const float*>(p + loBOff)), hiB = __ldg(reinterpret_cast<const float*>(p + hiBOff));
        const int32_t ref = __ldg(reinterpret_cast<const int32_t*>(p + childOff));
        const float uN = __fmaf_rd(sgF, nearK, mref), uF = __fmaf_ru(sgF, farK, mref);
        const float sa = fmaxf(uN, umin), sb = fminf(uF, umax);
        const float aL = fminf(__fmaf_rd(rAmin, sa, cAmin), __fmaf_rd(rAmin, sb, cAmin));
        const float aU = fmaxf(__fmaf_ru(rAmax, sa, cAmax), __fmaf_ru(rAmax, sb, cAmax));
        const float bL = fminf(__fmaf_rd(rBmin, sa, cBmin), __fmaf_rd(rBmin, sb, cBmin));
        const float bU = fmaxf(__fmaf_ru(rBmax, sa, cBmax), __fmaf_ru(rBmax, sb, cBmax));
        const bool ok = childLane && sa <= sb && aL <= hiA && aU >= loA && bL <= hiB && bU >= loB;
        if (COUNT && childLane && ref != kEmptyRef) C.box++;
#ifdef RT_BEAM_STATS
        if (COUNT && lane == 0) C.sphere++;           // warp-level: wide-node visits
#endif
        const uint32_t m = __ballot_sync(RT_FULL, ok);
        if (m) {
            const int n = __popc(m);
            // nearest child on top; an inner entry carries its near face (travel coordinate), a leaf entry its box
            if (ok) stsPair(sp + 8u * (uint32_t)(n - 1 - __popc(m & below)), ref, ref >= 0 ? __float_as_int(uN) : cur * kWideFloats + slot);
            sp += 8u * (uint32_t)n;
            if (COUNT && (int)((sp - spBase) >> 3) - 1 > C.maxStack) C.maxStack = (int)((sp - spBase) >> 3) - 1;
            __syncwarp();
        }
        // ---- pop: inner node -> next visit; leaf -> per-ray box test, then the triangles
        for (;;) {
            sp -= 8u;
            const int2 e = ldsPair(sp);
            if (e.x == kEmptyRef) return occluded;
            if (e.x >= 0) {
                if (__int_as_float(e.y) > umax) continue;       // behind every ray's closest hit
                cur = e.x;
                break;
            }
            const float* bx = reinterpret_cast<const float*>(S.wnodes) + e.y;
            const float l0 = __ldg(bx), l1 = __ldg(bx + kWide), l2 = __ldg(bx + 2 * kWide);
            const float h0 = __ldg(bx + 3 * kWide), h1 = __ldg(bx + 4 * kWide), h2 = __ldg(bx + 5 * kWide);
            const float a0 = __fmaf_rn(l0, ix, nx), b0 = __fmaf_rn(h0, ix, nx);
            const float a1 = __fmaf_rn(l1, iy, ny), b1 = __fmaf_rn(h1, iy, ny);
            const float a2 = __fmaf_rn(l2, iz, nz), b2 = __fmaf_rn(h2, iz, nz);
            const float nn = fmaxf(fmaxf(fminf(a0, b0), fminf(a1, b1)), fmaxf(fminf(a2, b2), tminF));
            const float ff = fminf(fminf(fmaxf(a0, b0), fmaxf(a1, b1)), fminf(fmaxf(a2, b2), tmaxF));
            const bool inl = nn <= ff;                 // lanes that are not (or no longer) part of the packet carry tmaxF = -inf
            if (COUNT && (live & bit) != 0) C.box++;
            const uint32_t lm = __ballot_sync(RT_FULL, inl);
#ifdef RT_BEAM_STATS
            if (COUNT && lane == 0) C.plane++;              // warp-level: leaf entries popped
            if (COUNT && lane == 1 && lm) C.inst++;         // ... of which some ray enters the box
#endif
            if (!lm) continue;
            budget--;
            RT_BS(leaves++);
            const uint32_t lref = ~(uint32_t)e.x;
            const int first = (int)(lref >> 4), count = (int)(lref & 7u) + 1;
            const bool f64 = (lref & 8u) != 0;
            bool occ = false, closer = false;
            for (int j = 0; j < count; j++) {
                V3 v0, v1, v2;
                loadTri(S, f64, first + j, v0, v1, v2);
                if (inl && !occ) {
                    if (COUNT) C.tri++;
                    double t, u, v;
                    if (mtTest(v0, v1, v2, oL, dL, eps, t, u, v) && inWindow(t, tLow, hit.t, anyhit)) {
                        if (anyhit) {
                            occ = true;
                        } else {
                            hit.t = t; hit.u = u; hit.v = v; hit.inst = instIdx; hit.slot = first + j;
                            tmaxF = __double2float_ru(t - t0);
                            closer = true;
                        }
                    }
                }
            }
            if (anyhit) {
                const uint32_t o = __ballot_sync(RT_FULL, occ);
                occluded |= o;
                live &= ~o;
                if (!live) return occluded;
                if (occ) tmaxF = -INF;                 // this lane is done: it fails every leaf box from here on
            } else if (__ballot_sync(RT_FULL, closer)) {
                umax = fminf(uClip, warpMaxF((live & bit) != 0 ? __double2float_ru((u0 + dKa * hit.t) * (1.0 + kRel)) : -INF));
            }
        }
    }
}

}  // namespace rtg
