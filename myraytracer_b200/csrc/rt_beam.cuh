// rt_beam.cuh — beam walk of a wide BVH: a coherent packet's inner-node tests with ONE LANE PER CHILD BOX.
//
// Same queries as packetBlas (rt_packet.cuh; intersectBLAS / occludedBLAS, Models/RTContext.swift:544-610,786-829)
// for the packets that are coherent — the 32 primary rays of one or two neighbouring pixels, or their shadow rays
// towards one light.  Such a packet walks a single root-to-leaf path of the binary BVH (C4: 20 node visits per
// packet, the depth of the tree), and in the packet walk every one of those visits costs the warp ~50 instructions
// in which 32 lanes test the same two boxes.  Here the parallelism is turned round:
//   * the packet is bounded ONCE by a beam — dominant axis K, travel coordinate u = sigma * x_K measured from the
//     rearmost ray origin, and for the two other axes the rays' extent on the planes u = 0 and u = U1 (the far end)
//     joined by straight lines: a two-plane hull, tight for bundles that spread out and for bundles that converge.
//     All of it is warp-uniform FP32 with an explicit widening (GInstance::beamSlack) above every rounding error;
//   * the tree is the binary tree collapsed to kWide = 16 children per node (gpu_layout.h: WNode, bvh_build.cpp:
//     collapseWide), and lanes s and s + 16 test child s against the beam: the slab of the box along K clipped to the
//     beam's range gives [sa, sb], and the beam's extent over [sa, sb] must overlap the box on axis A (lower half of
//     the warp) and on axis B (upper half).  One visit decides 16 children — a 16-ary tree over a million
//     triangles is 5 to 6 levels deep;
//   * the children that pass are pushed in front-to-back order (they are stored sorted along the node's widest axis; a
//     beam travelling down that axis takes them in reverse) with one ballot + popcount; inner entries carry the travel
//     coordinate of their near face and are skipped when popped behind the farthest closest hit of the packet;
//   * a LEAF entry is tested per ray, exactly as before: the conservative FP32 slab test of rt_packet.cuh against the
//     leaf's box gives the lane mask, then the FP64 Moeller-Trumbore tests in the reference's operation order.  What
//     decides a result is unchanged; the beam only culls, and it culls conservatively (it contains every ray of the
//     packet), so no hit the reference finds is lost and t, u, v stay bit-identical.
// A packet that cannot be bounded this way (dominant direction components of both signs, slopes above RT_BEAM_SLOPE,
// non-finite values) or that spends its visit budget is finished by the packet walk; incoherent packets (child rays
// off the ray queue) and meshes without a wide tree take the packet walk from the start.
//
// Every function here must be entered by all 32 lanes of a converged warp.
#pragma once

namespace rtg {

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7fffffff); }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }
// min / max of a float over the warp: one REDUX on an order-preserving integer image
__device__ __forceinline__ float warpMinF(float v) { return ord2f(__reduce_min_sync(RT_FULL, f2ord(v))); }
__device__ __forceinline__ float warpMaxF(float v) { return ord2f(__reduce_max_sync(RT_FULL, f2ord(v))); }

#define RT_BEAM_SLOPE 4.0     // |d_other| <= slope * |d_K| for every ray of a beam
// A beam is the hull of its rays: a packet whose rays split into bundles far apart (primary hits on both sides of a
// silhouette, then their shadow rays) bounds the space between the bundles too.  The walk therefore carries a budget
// of node visits + leaf entries; a packet that spends it gives up (formed = false with what it found so far — closest
// hits stand, occluded lanes are returned) and the caller finishes with the packet walk, whose lane masks follow
// the bundles separately.  A coherent packet needs ~10 visits and ~5 leaves on the million-triangle terrain.
#ifndef RT_BEAM_BUDGET
#define RT_BEAM_BUDGET 96
#endif

// The lane's FP32 view of its ray inside one BLAS (set up by packetBlas: origin advanced to the root box in FP64, then
// rounded; t relative to that point) — what the conservative per-ray slab test reads, and what the beam is built from.
struct LaneRayF {
    float ox, oy, oz;        // origin relative to the BLAS centre, advanced by t0
    float ix, iy, iz;        // clamped reciprocal direction
    float nx, ny, nz;        // -o * i
    float tminF, tmaxF;      // window relative to t0; tmaxF = -inf: the lane is not (or no longer) part of the packet
    double t0;
};

// Returns true when the walk is complete; false when no beam could be formed or its budget is spent — the caller then
// runs the packet walk with what this one left: `live` (lanes still searching), `occluded`, hit / L.tmaxF.
template <bool COUNT>
__device__ __forceinline__ bool beamWalk(const DeviceScene& S, const GInstance& I, int instIdx, V3 oL, V3 dL, double eps,
                                         double tLow, HitRec& hit, bool anyhit, uint32_t& live, uint32_t& occluded, LaneRayF& L,
                                         int2* wstack, Counters& C) {
    const int lane = threadIdx.x & 31;
    const uint32_t bit = 1u << lane;
    const bool alive = (live & bit) != 0;
    const float INF = CUDART_INF_F;

    // ---- the beam's frame: K = dominant axis of the first ray, A and B the two others (cyclic)
    const float dx = (float)dL.x, dy = (float)dL.y, dz = (float)dL.z;
    const int leader = __ffs((int)live) - 1;
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    const int kMine = ax >= ay ? (ax >= az ? 0 : 2) : (ay >= az ? 1 : 2);
    const int signs = (dx < 0.0f ? 1 : 0) | (dy < 0.0f ? 2 : 0) | (dz < 0.0f ? 4 : 0);
    const int lead = __shfl_sync(RT_FULL, kMine | (signs << 2), leader);
    const int k = lead & 3, downBits = lead >> 2;         // downBits: axes along which the first ray travels downwards
    const bool k0 = k == 0, k1 = k == 1;
    const bool neg = ((downBits >> k) & 1) != 0;
    const float dK = k0 ? dx : k1 ? dy : dz, dA = k0 ? dy : k1 ? dz : dx, dB = k0 ? dz : k1 ? dx : dy;
    const float oK = k0 ? L.ox : k1 ? L.oy : L.oz, oA = k0 ? L.oy : k1 ? L.oz : L.ox, oB = k0 ? L.oz : k1 ? L.ox : L.oy;
    const float dKa = neg ? -dK : dK;
    const bool okLane = !alive || (dKa > 0.0f && fabsf(dA) <= (float)RT_BEAM_SLOPE * dKa && fabsf(dB) <= (float)RT_BEAM_SLOPE * dKa);
    if (__ballot_sync(RT_FULL, !okLane)) return false;

    // ---- per ray: position on A, B as a function of the travel coordinate u = sigma * x_K - ref (u >= u0 >= 0):
    //      p(u) = c + r u.  The beam is the rays' extent on two planes — u = 0 (the rearmost origin) and u = U1 (the far
    //      end: the farthest closest hit / shadow-ray end, never beyond the root box) — joined by straight lines: a ray
    //      lies between the line through the two minima and the line through the two maxima for every u in [0, U1], and
    //      the hull is tight for bundles that spread out (primary rays) as well as for bundles that converge (shadow
    //      rays towards a point light).  Plain FP32: every rounding below is a few 2^-24 of values bounded by 17 M
    //      (M = largest |coordinate - centre| of the BLAS; slopes <= RT_BEAM_SLOPE, u <= 2 M), together < 2^-17 M (the reciprocals are 2-ulp MUFU results), and
    //      the extents are widened by I.beamSlack = 2^-15 M (3 % of a triangle edge of the million-triangle terrain).
    const float E = (float)I.beamSlack;
    const float inv = __fdividef(1.0f, dKa);
    const float rA = dA * inv, rB = dB * inv;
    const float xs = neg ? -oK : oK;
    const float ref = warpMinF(alive ? xs : INF);
    const float u0 = xs - ref;                              // >= 0
    // the part of the ray before its entry into the root box (t < t0) meets no geometry: the window starts at max(tmin, 0)
    const float umin = warpMinF(alive ? __fmaf_rn(dKa, fmaxf(L.tminF, 0.0f), u0) : INF) - E;
    const float rootFar = (float)(neg ? -(k0 ? I.rootLo[0] - I.centre[0] : k1 ? I.rootLo[1] - I.centre[1] : I.rootLo[2] - I.centre[2])
                                      : (k0 ? I.rootHi[0] - I.centre[0] : k1 ? I.rootHi[1] - I.centre[1] : I.rootHi[2] - I.centre[2]));
    float umax = fminf(rootFar - ref, warpMaxF(alive ? __fmaf_rn(dKa, L.tmaxF, u0) : -INF)) + E;
    const float U1 = umax;
    if (!(U1 > 1e-30f && U1 < 1e30f && fabsf(ref) < 1e30f && E > 0.0f)) return false;      // warp-uniform
    const float cA = __fmaf_rn(-rA, u0, oA), cB = __fmaf_rn(-rB, u0, oB);
    const float fA = __fmaf_rn(rA, U1, cA), fB = __fmaf_rn(rB, U1, cB);
    const float cAl = warpMinF(alive ? cA : INF) - E, cAu = warpMaxF(alive ? cA : -INF) + E;
    const float cBl = warpMinF(alive ? cB : INF) - E, cBu = warpMaxF(alive ? cB : -INF) + E;
    const float fAl = warpMinF(alive ? fA : INF) - E, fAu = warpMaxF(alive ? fA : -INF) + E;
    const float fBl = warpMinF(alive ? fB : INF) - E, fBu = warpMaxF(alive ? fB : -INF) + E;
    const float invU = __fdividef(1.0f, U1);
    const float gAl = (fAl - cAl) * invU, gAu = (fAu - cAu) * invU, gBl = (fBl - cBl) * invU, gBu = (fBu - cBu) * invU;
    if (!(((fabsf(gAl) + fabsf(gAu)) + (fabsf(gBl) + fabsf(gBu))) + ((fabsf(cAl) + fabsf(cAu)) + (fabsf(cBl) + fabsf(cBu))) + fabsf(umin) < 1e30f)) {
        return false;                                       // warp-uniform: a value overflowed FP32 (or is NaN)
    }

    static_assert(kWide == 16, "the two halves of a warp share the kWide children of a node");
    const float sgF = neg ? -1.0f : 1.0f, mref = -ref;
    const uint32_t below = (1u << (lane & (kWide - 1))) - 1u, self = 1u << (lane & (kWide - 1));
    // Lane s and lane s + kWide both hold child s: the lower half of the warp tests the box against the beam's extent on
    // axis A, the upper half on axis B (each lane keeps the two lines of ITS axis only), and a child passes when both do.
    const bool upper = lane >= kWide;
    const float gl = upper ? gBl : gAl, cl = upper ? cBl : cAl, gu = upper ? gBu : gAu, cu = upper ? cBu : cAu;
    const int kJ = upper ? (k == 0 ? 2 : k - 1) : (k == 2 ? 0 : k + 1);      // this lane's axis: B = (k + 2) % 3, A = (k + 1) % 3
    const bool j0 = kJ == 0, j1 = kJ == 1;
    // this lane's child record in node 0; node n is n * sizeof(WNode) further
    const char* laneBase = reinterpret_cast<const char*>(S.wnodes) + (lane & (kWide - 1)) * (int)sizeof(WChild);

    const uint32_t spBase = (uint32_t)__cvta_generic_to_shared(wstack);
    uint32_t sp = spBase + 8u;
    stsPair(spBase, kEmptyRef, 0);            // sentinel: popping it ends the walk
    int32_t cur = I.wideRoot;
    int budget = RT_BEAM_BUDGET;
    for (;;) {
        if (--budget < 0) return false;
        // ---- one wide node: lanes s and s + kWide test child s
        const float4* cp = reinterpret_cast<const float4*>(laneBase + (size_t)(uint32_t)cur * sizeof(WNode));
        const float4 q0 = __ldg(cp), q1 = __ldg(cp + 1);         // lo.x lo.y lo.z hi.x | hi.y hi.z ref aux
        const int32_t ref_ = __float_as_int(q1.z);
        const float loK = k0 ? q0.x : k1 ? q0.y : q0.z, hiK = k0 ? q0.w : k1 ? q1.x : q1.y;
        const float loJ = j0 ? q0.x : j1 ? q0.y : q0.z, hiJ = j0 ? q0.w : j1 ? q1.x : q1.y;
        const float u1 = __fmaf_rn(sgF, loK, mref), u2 = __fmaf_rn(sgF, hiK, mref);
        const float uN = fminf(u1, u2), uF = fmaxf(u1, u2);
        const float sa = fmaxf(uN, umin), sb = fminf(uF, umax);
        const float jL = fminf(__fmaf_rn(gl, sa, cl), __fmaf_rn(gl, sb, cl));
        const float jU = fmaxf(__fmaf_rn(gu, sa, cu), __fmaf_rn(gu, sb, cu));
        const bool okJ = sa <= sb && jL <= hiJ && jU >= loJ;
        if (COUNT && !upper && ref_ != kEmptyRef) C.box++;
        const uint32_t mj = __ballot_sync(RT_FULL, okJ);
        const uint32_t m = mj & (mj >> kWide);
        if (m) {
            // front to back: the children are sorted up the node's sort axis; a beam travelling down it takes them in reverse
            const int axisS = __shfl_sync(RT_FULL, __float_as_int(q1.w), 0);
            const bool rev = ((downBits >> axisS) & 1) != 0;
            const int n = __popc(m);
            const int before = rev ? __popc(m & ~(below | self)) : __popc(m & below);      // hit children visited before this one
            // nearest child on top; an inner entry carries its near face (travel coordinate), a leaf entry its box
            if (!upper && (m & self) != 0)
                stsPair(sp + 8u * (uint32_t)(n - 1 - before), ref_, ref_ >= 0 ? __float_as_int(uN) : cur * kWide + lane);
            sp += 8u * (uint32_t)n;
            if (COUNT && (int)((sp - spBase) >> 3) - 1 > C.maxStack) C.maxStack = (int)((sp - spBase) >> 3) - 1;
            __syncwarp();
        }
        // ---- pop: inner node -> next visit; leaf -> per-ray box test, then the triangles
        for (;;) {
            sp -= 8u;
            const int2 e = ldsPair(sp);
            if (e.x == kEmptyRef) return true;
            if (e.x >= 0) {
                if (__int_as_float(e.y) > umax) continue;       // behind every ray's closest hit
                cur = e.x;
                break;
            }
            const float4* bp = reinterpret_cast<const float4*>(S.wnodes) + 2 * (size_t)(uint32_t)e.y;
            const float4 b0_ = __ldg(bp), b1_ = __ldg(bp + 1);
            const float a0 = __fmaf_rn(b0_.x, L.ix, L.nx), b0 = __fmaf_rn(b0_.w, L.ix, L.nx);
            const float a1 = __fmaf_rn(b0_.y, L.iy, L.ny), b1 = __fmaf_rn(b1_.x, L.iy, L.ny);
            const float a2 = __fmaf_rn(b0_.z, L.iz, L.nz), b2 = __fmaf_rn(b1_.y, L.iz, L.nz);
            const float nn = fmaxf(fmaxf(fminf(a0, b0), fminf(a1, b1)), fmaxf(fminf(a2, b2), L.tminF));
            const float ff = fminf(fminf(fmaxf(a0, b0), fmaxf(a1, b1)), fminf(fmaxf(a2, b2), L.tmaxF));
            const bool inl = nn <= ff;                 // lanes that are not (or no longer) part of the packet carry tmaxF = -inf
            if (COUNT && (live & bit) != 0) C.box++;
            const uint32_t lm = __ballot_sync(RT_FULL, inl);
            if (!lm) continue;
            budget--;
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
                            L.tmaxF = __double2float_ru(t - L.t0);
                            closer = true;
                        }
                    }
                }
            }
            if (anyhit) {
                const uint32_t o = __ballot_sync(RT_FULL, occ);
                occluded |= o;
                live &= ~o;
                if (!live) return true;
                if (occ) L.tmaxF = -INF;               // this lane is done: it fails every leaf box from here on
            } else if (__ballot_sync(RT_FULL, closer)) {
                umax = fminf(U1, warpMaxF((live & bit) != 0 ? __fmaf_rn(dKa, L.tmaxF, u0) : -INF) + E);
            }
        }
    }
}

}  // namespace rtg
