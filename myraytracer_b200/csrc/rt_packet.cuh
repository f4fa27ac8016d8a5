// rt_packet.cuh — warp-packet traversal: the 32 rays of a warp walk the TLAS/BLAS together.
//
// Same queries as rt_traverse.cuh / rt_trace.cuh (intersectTLAS/BLAS, occludedTLAS/BLAS,
// Models/RTContext.swift:544-829) for COHERENT ray bundles: the 32 lanes of a warp hold the samples
// of one or two neighbouring pixels (primary rays) or their shadow rays towards one light, so their
// BVH paths overlap almost completely.  Instead of 32 private stacks and 32 divergent node fetches
// (measured: 16 of 32 lanes active in the while-while loop), the warp keeps ONE stack of
// (node, lane-mask) entries in shared memory, fetches each node once (a warp-uniform, broadcast
// load) and every lane tests its own ray against it; control flow is decided by ballots only, so
// the warp never diverges inside the traversal.  A lane whose ray misses a box is simply cleared
// from the mask that travels with the node.
// The per-lane arithmetic (conservative FP32 slab test, exact FP64 Moeller-Trumbore in the
// reference's operation order) is unchanged, so results are identical to the per-lane path except
// for the visiting order, which only matters at equal-t ties.
//
// Every function here must be entered by all 32 lanes of a converged warp.
#pragma once
#include "rt_trace.cuh"

namespace rtg {

#define RT_FULL 0xffffffffu

// Packet stack entries through 32-bit shared-window addresses: the stack pointer IS the address, so the loop
// never rebuilds the shared base (three uniform-datapath instructions per push and per pop otherwise).
#ifndef RT_EMU
__device__ __forceinline__ void stsPair(uint32_t addr, int a, int b) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ int2 ldsPair(uint32_t addr) {
    int2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
    return v;
}
#else   // host emulation of the kernels (tests/emu, test infrastructure only): the shared window starts at an anchor
__device__ __forceinline__ void stsPair(uint32_t addr, int a, int b) {
    *reinterpret_cast<int2*>(rtemu::g_sharedBase + (intptr_t)(int32_t)addr) = make_int2(a, b);
}
__device__ __forceinline__ int2 ldsPair(uint32_t addr) { return *reinterpret_cast<int2*>(rtemu::g_sharedBase + (intptr_t)(int32_t)addr); }
#endif
// RT_PREFETCH=1: prefetch.global.L1 of both children before the box tests decide between them; =2: into L2 only
#if defined(RT_PREFETCH) && !defined(RT_EMU)
__device__ __forceinline__ void prefetchL1(const void* p) {
#if RT_PREFETCH == 2
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}
#else
__device__ __forceinline__ void prefetchL1(const void*) {}
#endif
#ifndef RT_WSTACK
#define RT_WSTACK (RT_BSTACK + 1)   // BLAS entries per warp: the packet walk needs the tree depth (checked against RT_STACK at
#endif                              // build time, <= 64), the beam walk of the wide tree RT_BSTACK entries + its sentinel
#define RT_WTSTACK 40   // TLAS entries per warp
#ifdef RT_WARPTIMES
#define RT_WT_EXTRA 2   // profiling build: (node visits, lanes), (leaf visits, lane-triangle tests) per warp
#else
#define RT_WT_EXTRA 0
#endif
#define RT_WARP_STACK_INT2 (RT_WSTACK + RT_WTSTACK + RT_WT_EXTRA)

}  // namespace rtg
#include "rt_beam.cuh"
namespace rtg {

// Packet walk of one mesh BLAS.  inMask (warp-uniform): lanes whose ray takes part.
// anyhit: run-time, warp-uniform (one copy of the loop for closest-hit and shadow packets).
// Returns (any-hit only) the warp-uniform mask of lanes found occluded.
// tryBeam (warp-uniform): the packet is coherent and the mesh has a wide tree — the beam walk (rt_beam.cuh) goes first
// and this walk only finishes what it gives up.
template <bool COUNT>
__device__ __forceinline__ uint32_t packetBlas(const DeviceScene& S, const GInstance& I, int instIdx, V3 oL, V3 dL,
                                               double eps, double tLow, HitRec& hit, bool anyhit, uint32_t inMask,
                                               int2* wstack, Counters& C, bool tryBeam = false) {
    const uint32_t bit = 1u << (threadIdx.x & 31);
    // per-lane root test + origin advance, exactly as traverseBlas
    V3 ia = refinedInv(dL);
    V3 oc = oL - ld3(I.centre);
    double lo[3] = {I.rootLo[0] - I.centre[0], I.rootLo[1] - I.centre[1], I.rootLo[2] - I.centre[2]};
    double hi[3] = {I.rootHi[0] - I.centre[0], I.rootHi[1] - I.centre[1], I.rootHi[2] - I.centre[2]};
    double tn, tf;
    slabApprox(lo, hi, oc, ia, tn, tf);
    const bool in = (inMask & bit) != 0;
    if (COUNT && in) C.box++;
    const bool alive = in && !(tn > tf || tf < tLow || tn > hit.t);
    uint32_t live = __ballot_sync(RT_FULL, alive);
    if (!live) return 0u;
    LaneRayF L;
    L.t0 = tn > 0.0 ? tn : 0.0;
    if (!(L.t0 < 1e300)) L.t0 = 0.0;
    V3 o2 = oc + dL * L.t0;
    L.ox = (float)o2.x; L.oy = (float)o2.y; L.oz = (float)o2.z;
    L.ix = fminf(fmaxf(1.0f / (float)dL.x, -1e30f), 1e30f);
    L.iy = fminf(fmaxf(1.0f / (float)dL.y, -1e30f), 1e30f);
    L.iz = fminf(fmaxf(1.0f / (float)dL.z, -1e30f), 1e30f);
    L.nx = -(L.ox * L.ix); L.ny = -(L.oy * L.iy); L.nz = -(L.oz * L.iz);
    L.tminF = __double2float_rd(tLow - L.t0);
    L.tmaxF = alive ? __double2float_ru(hit.t - L.t0) : -CUDART_INF_F;
    uint32_t occluded = 0u;
    if (tryBeam && beamWalk<COUNT>(S, I, instIdx, oL, dL, eps, tLow, hit, anyhit, live, occluded, L, wstack, C)) return occluded;
    const float ix = L.ix, iy = L.iy, iz = L.iz, nx = L.nx, ny = L.ny, nz = L.nz, tminF = L.tminF;
    float tmaxF = L.tmaxF;
    const double t0 = L.t0;

    const GNode* nodes = S.nodes;        // child references are absolute (gpu_layout.h)
    // sp: shared-window address of the next free entry.  Entry 0 is a sentinel (kEmptyRef, all lanes): popping it
    // ends the walk, so the loop needs neither the stack base nor a depth counter.
    const uint32_t spBase = (uint32_t)__cvta_generic_to_shared(wstack);
    uint32_t sp = spBase + 8u;
    stsPair(spBase, kEmptyRef, (int)RT_FULL);
    int32_t cur = I.rootRef;
    uint32_t curMask = live;

#define RT_PACKET_POP()                                  \
    for (;;) {                                           \
        sp -= 8u;                                        \
        const int2 e_ = ldsPair(sp);                     \
        cur = e_.x;                                      \
        curMask = (uint32_t)e_.y & live;                 \
        if (curMask) break;                              \
    }                                                    \
    if (cur == kEmptyRef) return occluded;

    for (;;) {
        while ((uint32_t)cur < 0x7fffffffu) {  // inner node, fetched once for the warp
            const float4* np = reinterpret_cast<const float4*>(nodes + cur);
            const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2);
            const int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
#ifdef RT_PREFETCH
            // both children are requested before the box tests decide between them: the next visit's node
            // (an L1 miss most of the time) is already on its way while this one is tested
            if ((uint32_t)ch.x < 0x7fffffffu) prefetchL1(nodes + ch.x);
            if ((uint32_t)ch.y < 0x7fffffffu) prefetchL1(nodes + ch.y);
#endif
            if (COUNT && (curMask & bit) != 0) C.box += 2;
#ifdef RT_WARPTIMES
            if (bit == 1u) { wstack[RT_WSTACK + RT_WTSTACK].x++; wstack[RT_WSTACK + RT_WTSTACK].y += __popc(curMask); }
#endif
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
            // No test of curMask here: a lane whose ray missed an ancestor's box misses the boxes inside it too, and
            // lanes that are not part of the packet (any more) carry tmaxF = -inf, which fails every box.
            const bool h0 = n0 <= f0, h1 = n1 <= f1;
            const uint32_t m0 = __ballot_sync(RT_FULL, h0), m1 = __ballot_sync(RT_FULL, h1);
            if (m0 && m1) {
                const uint32_t v1 = __ballot_sync(RT_FULL, h1 && (!h0 || n1 < n0));   // lanes preferring child 1
                const bool swap = 2 * __popc(v1) > __popc(m0 | m1);
                stsPair(sp, swap ? ch.x : ch.y, (int)(swap ? m0 : m1));
                sp += 8u;
                if (COUNT && (int)((sp - spBase) >> 3) - 1 > C.maxStack) C.maxStack = (int)((sp - spBase) >> 3) - 1;
                cur = swap ? ch.y : ch.x;
                curMask = swap ? m1 : m0;
            } else if (m0) {
                cur = ch.x; curMask = m0;
            } else if (m1) {
                cur = ch.y; curMask = m1;
            } else {
                RT_PACKET_POP();
            }
        }
        if (cur != kEmptyRef) {  // leaf: triangles fetched once for the warp, tested by the lanes in curMask
            const uint32_t ref = ~(uint32_t)cur;
            const int first = (int)(ref >> 4), count = (int)(ref & 7u) + 1;
            const bool f64 = (ref & 8u) != 0;
            const bool inl = (curMask & bit) != 0;
            bool occ = false;
#ifdef RT_WARPTIMES
            if (bit == 1u) { wstack[RT_WSTACK + RT_WTSTACK + 1].x++; wstack[RT_WSTACK + RT_WTSTACK + 1].y += __popc(curMask) * count; }
#endif
            for (int k = 0; k < count; k++) {
                V3 v0, v1, v2;
                loadTri(S, f64, first + k, v0, v1, v2);
                if (inl && !occ) {
                    if (COUNT) C.tri++;
                    double t, u, v;
                    if (mtTest(v0, v1, v2, oL, dL, eps, t, u, v) && inWindow(t, tLow, hit.t, anyhit)) {
                        if (anyhit) {
                            occ = true;
                        } else {
                            hit.t = t; hit.u = u; hit.v = v; hit.inst = instIdx; hit.slot = first + k;
                            tmaxF = __double2float_ru(t - t0);
                        }
                    }
                }
            }
            if (anyhit) {
                const uint32_t o = __ballot_sync(RT_FULL, occ);
                occluded |= o;
                live &= ~o;
                if (!live) return occluded;
                if (occ) tmaxF = -CUDART_INF_F;      // this lane is done: it fails every box from here on
            }
        }
        RT_PACKET_POP();
    }
#undef RT_PACKET_POP
}

// One instance for the lanes in inMask (warp-uniform instance index): RTContext.swift:657-705.
// INCOHERENT (compile time: the kernel instantiation with the ray queue) + incoherent (run time, warp-uniform: the rays
// are child rays off the queue, not primary rays or their shadow rays): a mesh of more than kBigBlasTriangles
// triangles is walked per lane — the lanes of the mask enter together and each follows its own path (while-while,
// rt_traverse.cuh) — because a packet visits the UNION of its rays' paths, which for 32 reflected rays in a
// million-triangle BVH is ten times one ray's path; small meshes and analytic primitives stay with the packet.
template <bool COUNT, bool INCOHERENT = false>
__device__ __forceinline__ uint32_t packetInstance(const DeviceScene& S, int idx, V3 o, V3 d, double eps, double tMin,
                                                   HitRec& hit, bool anyhit, uint32_t inMask, int2* wstack, Counters& C,
                                                   bool incoherent = false) {
    const GInstance& I = S.instances[idx];
    const bool in = (inMask >> (threadIdx.x & 31)) & 1u;
    V3 oL, dL;
    toLocal(I, o, d, oL, dL);
    if (COUNT && in) C.inst++;
    const double tLow = dmax(eps, tMin);
    if (I.kind == INST_MESH) {
        if (INCOHERENT && incoherent && I.bigBlas) {
            bool o1 = false;
            if (in) o1 = traverseBlas<COUNT>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, nullptr, 0, C);
            return __ballot_sync(RT_FULL, o1);
        }
        // coherent packet, mesh with a wide tree: the beam walk first
        const bool tryBeam = !(INCOHERENT && incoherent) && I.wideRoot != kEmptyRef &&
                             (S.wideMode == 1 || (S.wideMode == 2 && !anyhit) || (S.wideMode == 3 && anyhit));
        return packetBlas<COUNT>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, inMask, wstack, C, tryBeam);
    }
    bool occ = false;
    if (in) occ = analyticInstance<COUNT>(S, I, idx, oL, dL, eps, tLow, hit, anyhit, C);
    return __ballot_sync(RT_FULL, occ);   // always 0 for closest-hit queries
}

// Packet TLAS walk.  `active`: this lane carries a ray.  Closest hit: updates hit; any-hit (hit.t = tMax):
// returns true for the lanes whose ray is occluded.
template <bool COUNT, bool INCOHERENT = false>
__device__ __forceinline__ bool packetScene(const DeviceScene& S, V3 o, V3 d, double tMin, HitRec& hit, bool active,
                                            bool anyhit, int2* wstack, Counters& C, bool incoherent = false) {
    const uint32_t bit = 1u << (threadIdx.x & 31);
    uint32_t live = __ballot_sync(RT_FULL, active);
    if (!live) return false;
    const double eps = S.eps;
    int32_t cur = S.tlasRootRef;
    if (cur == kEmptyRef) return false;
    if (cur < 0) return (packetInstance<COUNT, INCOHERENT>(S, ~cur, o, d, eps, tMin, hit, anyhit, live, wstack, C, incoherent) & bit) != 0;
    const double tLow = dmax(eps, tMin);
    RayF rf = makeRayF(S.tlasLo, S.tlasHi, S.tlasCentre, o, d, refinedInv(d), tLow, hit.t);
    if (COUNT && active) C.box++;
    live = __ballot_sync(RT_FULL, active && rf.hits);
    if (!live) return false;
    if (!(active && rf.hits)) rf.tmax = -CUDART_INF_F;          // not part of the packet: fails every box
    int2* tstack = wstack + RT_WSTACK;
    uint32_t occluded = 0u, curMask = live;
    int sp = 0;
    for (;;) {
        bool descended = false;
        if (cur >= 0) {
            float n0, n1;
            bool h0, h1;
            int2 ch;
            twoBoxes(S.tlas + cur, rf, n0, h0, n1, h1, ch);       // conservative FP32, like the BLAS nodes
            const bool inl = (curMask & bit) != 0;
            if (COUNT && inl) C.box += 2;
            h0 = h0 && inl; h1 = h1 && inl;
            const uint32_t m0 = __ballot_sync(RT_FULL, h0), m1 = __ballot_sync(RT_FULL, h1);
            if (m0 && m1) {
                const uint32_t v1 = __ballot_sync(RT_FULL, h1 && (!h0 || n1 < n0));
                const bool swap = 2 * __popc(v1) > __popc(m0 | m1);
                if (sp < RT_WTSTACK) { tstack[sp] = swap ? make_int2(ch.x, (int)m0) : make_int2(ch.y, (int)m1); sp++; }
                cur = swap ? ch.y : ch.x;
                curMask = swap ? m1 : m0;
                descended = true;
            } else if (m0) { cur = ch.x; curMask = m0; descended = true; }
            else if (m1) { cur = ch.y; curMask = m1; descended = true; }
        } else {
            const uint32_t o_ = packetInstance<COUNT, INCOHERENT>(S, ~cur, o, d, eps, tMin, hit, anyhit, curMask, wstack, C, incoherent);
            if (anyhit) {
                occluded |= o_;
                live &= ~o_;
                if (!live) return (occluded & bit) != 0;
                if (o_ & bit) rf.tmax = -CUDART_INF_F;
            } else {
                rf.tmax = fminf(rf.tmax, __double2float_ru(hit.t - rf.t0));   // a closer hit prunes the boxes that follow
            }
        }
        if (descended) continue;
        for (;;) {
            if (sp == 0) return (occluded & bit) != 0;
            --sp;
            const int2 e = tstack[sp];
            cur = e.x;
            curMask = (uint32_t)e.y & live;
            if (curMask) break;
        }
    }
}

}  // namespace rtg
