// rt_render_kernel.cuh — the persistent render kernels for sm_100a.   COMPILE WITH -fmad=false.
//
// One kernel replaces the reference's whole per-frame loop:
//   Renderer.render pixel/sample loops     Extensions/Object+Extension.swift:285-361
//   primary-ray generation                 :292-344      (PCG32 :556-589)
//   trace closure                          :96-283       (rt_shade.cuh)
//   sample average                         :354
//   RGBA8 conversion                       RayTracer.swift:186-195
// Work distribution (both kernels): warps pull 8x4-pixel tiles of the rank's 8-row chunks (the
// reference's chunk unit, :75-82) from a global atomic counter (persistent threads).
//
//   renderKernelFast    lane = one (pixel, sample) item; the tile's 32*n*n items are walked 32 at a
//                       time with the samples of a pixel in adjacent lanes (maximum ray coherence).
//                       Depth 0 — the primary ray and its shadow rays — runs as a warp packet
//                       (rt_packet.cuh) through ONE traversal call site; spawned children go to the
//                       per-lane depth-first loop (rt_shade.cuh).  Sample s of a pixel starts its PCG32 stream 3 (5 with
//                       DOF) draws after sample s-1 (:307-340), reached by the LCG skip-ahead; this is
//                       exact as long as no draw happens inside `trace`, i.e. every material has
//                       roughness == 0 (:191,210,259) — the host checks that.  The per-pixel sum is
//                       formed in sample order (shuffles), so the FP64 result equals the sequential one.
//   renderKernelGeneral lane = one pixel looping over its samples with one sequential PCG32 stream and
//                       the per-lane traversal: the general path (glossy materials draw inside the recursion).
#pragma once
#include <cuda_runtime.h>

#include "rt_shade.cuh"

namespace rtg {

// Resident blocks per SM the kernels are compiled for (register cap = 65536 / (threads * blocks)).  Measured on
// B200 (C4, profiles/): the traversal is latency-bound, so more resident warps beat fewer spills.
// The fast kernel runs ONE WARP PER BLOCK, 32 blocks per SM (64 registers): nothing is shared between the warps
// of a block, and with a single warp the packet stack sits at a fixed shared-memory address — the compiler no
// longer rebuilds it from threadIdx inside the node loop (3.5 % of the frame, measured).
#ifndef RT_FAST_BLOCK
#define RT_FAST_BLOCK 32
#endif
#ifndef RT_FAST_BLOCKS
#define RT_FAST_BLOCKS 32
#endif
#ifndef RT_SKIPTAB
#define RT_SKIPTAB 1
#endif
// The general kernel wants registers, not warps: its per-lane FP64 path state (ray, throughput, shading point, light
// sums, RNG, hit record) does not fit 96 registers, and every spill sits on the path between two dependent scene walks.
// Measured on B200 (C3 glossy / C3 with area lights, ms per frame): 5 blocks of 128 threads per SM (96 registers) 18.6 /
// 28.9, 3 blocks (168) 15.8 / 25.9, 2 blocks (255 registers, 8 warps per SM) 13.4 / 21.8.
#ifndef RT_GEN_BLOCKS
#define RT_GEN_BLOCKS 2
#endif

struct PrimaryRay {
    V3 o, d;
    double tMin;
    double time;      // Ray.time in [0, 1): only motion blur reads it
};

// Object+Extension.swift:307-344 for sample (sx, sy) of pixel (i, j); draws xi1, xi2, [lens u, v], time.
__device__ __forceinline__ PrimaryRay primaryRay(const CameraParams& cam, int i, int j, int sx, int sy, PCG32& rng) {
    const V3 eye = ld3(cam.eye), cu = ld3(cam.u), cv = ld3(cam.v), cw = ld3(cam.w), q00 = ld3(cam.q00), cm = ld3(cam.m);
    const int n = cam.n;
    const double xi1 = rng.nextFloat();
    const double xi2 = rng.nextFloat();
    const double iOffset = ddiv((double)sx + xi1, (double)n);
    const double jOffset = ddiv((double)sy + xi2, (double)n);
    const double currentI = (double)i + iOffset;
    const double currentJ = (double)j + jOffset;
    const V3 vOff = cv * (currentJ * cam.dv);
    const V3 rowTopLeft = q00 - vOff;
    const V3 uOff = cu * (currentI * cam.du);
    const V3 s = rowTopLeft + uOff;
    const V3 dir0 = normalize(s - eye);
    PrimaryRay r;
    r.d = dir0; r.o = eye;
    if (cam.dof) {  // :325-338
        const V3 forward = -cw;
        const double denom = dot(dir0, forward);
        const double tFocus = fabs(denom) < 1e-6 ? cam.focus : ddiv(cam.focus, denom);
        const V3 pFocus = eye + dir0 * tFocus;
        const double uRand = rng.nextFloat() - 0.5;
        const double vRand = rng.nextFloat() - 0.5;
        const V3 lensOffset = (uRand * cu + vRand * cv) * cam.aperture;
        const V3 a = eye + lensOffset;
        r.d = normalize(pFocus - a);
        r.o = a;
    }
    r.time = rng.nextFloat();  // ray time: always drawn (:340)
    const double denomW = dot(r.d, cw);
    const double tImg = ddiv(dot(cm - r.o, cw), denomW == 0.0 ? 4.9406564584124654e-324 : denomW);  // :342
    r.tMin = dmax(tImg, 0.0);
    return r;
}

__device__ __forceinline__ void writePixel(const RenderTargets& out, size_t pix, V3 c) {
    if (out.linear) { out.linear[pix * 3] = c.x; out.linear[pix * 3 + 1] = c.y; out.linear[pix * 3 + 2] = c.z; }
    if (out.rgba) {
        uchar4 q;
        q.x = quantize(c.x); q.y = quantize(c.y); q.z = quantize(c.z); q.w = 255;
        reinterpret_cast<uchar4*>(out.rgba)[pix] = q;
    }
}

// a finished pixel: its colour — or, in the counting pass of an area-light frame, its number of jitter-table draws
template <bool AREACOUNT>
__device__ __forceinline__ void emitPixel(const RenderTargets& out, size_t pix, V3 c) {
    if (AREACOUNT) out.areaIndex[pix] = (unsigned)c.x;
    else writePixel(out, pix, c);
}

template <bool COUNT>
__device__ __forceinline__ void flushCounters(const RenderTargets& out, Counters& C, RayCounts& R) {
    for (int off = 16; off > 0; off >>= 1) {
        R.primary += __shfl_down_sync(0xffffffffu, R.primary, off);
        R.shadow += __shfl_down_sync(0xffffffffu, R.shadow, off);
        R.secondary += __shfl_down_sync(0xffffffffu, R.secondary, off);
        if (COUNT) {
            C.box += __shfl_down_sync(0xffffffffu, C.box, off);
            C.tri += __shfl_down_sync(0xffffffffu, C.tri, off);
            C.sphere += __shfl_down_sync(0xffffffffu, C.sphere, off);
            C.plane += __shfl_down_sync(0xffffffffu, C.plane, off);
            C.inst += __shfl_down_sync(0xffffffffu, C.inst, off);
            C.smooth += __shfl_down_sync(0xffffffffu, C.smooth, off);
            int m = __shfl_down_sync(0xffffffffu, C.maxStack, off);
            if (m > C.maxStack) C.maxStack = m;
        }
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out.counters[0], R.primary);
        atomicAdd(&out.counters[1], R.shadow);
        atomicAdd(&out.counters[2], R.secondary);
        if (COUNT) {
            atomicAdd(&out.counters[3], C.box);
            atomicAdd(&out.counters[4], C.tri);
            atomicAdd(&out.counters[5], C.sphere);
            atomicAdd(&out.counters[6], C.plane);
            atomicAdd(&out.counters[7], C.inst);
            atomicMax(&out.counters[8], (unsigned long long)C.maxStack);
            atomicAdd(&out.counters[9], C.smooth);
        }
    }
}

#ifdef RT_WARPTIMES
__device__ __forceinline__ unsigned long long globalTimerNs() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#endif

__device__ __forceinline__ double shflD(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// RECURSE: the scene can spawn child rays (maxRecursionDepth > 0 and a mirror / dielectric / conductor material).  Such
// scenes run the instantiation that carries the warp's shared ray queue (rt_shade.cuh: WarpQueue) and batches up to
// RT_BATCH_PASSES passes (128 samples) of depth 0 before it drains the queue, so that the secondary rays of a whole
// 8x4 tile at 4 spp are traced 32 at a time; scenes without recursion run the lean instantiation (no queue code, no
// colour buffer in shared memory: the packet walk keeps its L1 and its instruction cache).
// BLOCKS: resident one-warp blocks per SM the instantiation is compiled for.  32 (64 registers) is what a scene of big
// meshes wants — its node walks are dependent chains of cache misses and resident warps hide them (C5: 178 ms at 32,
// 189 at 24, 209 at 16).  A scene of analytic primitives and small meshes with recursion (C3) spends its time in
// shading code that streams through the instruction caches, and 12 warps of 168 registers — no spills, a third as many
// warps fetching — run it 12 % faster (6.49 -> 5.71 ms; rt_kernel_variant.cu: RT_FAST_BLOCKS_SMALL): RenderTargets::recurse
// == 2 takes that instantiation.
// AREACOUNT: pass 1 of an area-light frame (RenderTargets::areaPass == 1) for scenes this kernel can follow — no glossy
// material, no motion blur.  The reference draws an area-light sample for every area light at every shading point that
// computes direct light (:145-186), through one running index per 8-row chunk; what pass 2 (general kernel) needs is each
// pixel's number of such draws.  This instantiation walks the same path trees — closest-hit packets, children through
// the ray queue — without shadow rays or light terms, adds nAreaLights per shading point to the sample's slot (small
// integers: exact in FP64 in any order) and leaves the pixel's total in areaIndex.  It replaces the general kernel's
// per-lane counting walk, which cost 40 % of such a frame.
// DEFER (with RECURSE = false; RenderTargets::deferPhase == 1): the first of two launches on a scene with recursion whose
// child rays are rare (big meshes, a few reflective objects: C5 reflects 0.65 % of its rays).  The queue instantiation
// costs every ray 12 % whether it spawns children or not (C4 forced through it: 12.07 against 10.81 ms), so depth 0
// runs here, in the lean code, and a unit in which some primary hit lands on a mirror / dielectric / conductor is not
// finished but appended to RenderTargets::deferList; the second launch — the queue instantiation with deferPhase == 2 —
// draws its units from that list and traces them whole (depth 0 again, then the queue).  A deferred unit's rays are
// not counted here and its pixels are written there.
template <bool COUNT, bool RECURSE, int BLOCKS = RT_FAST_BLOCKS, bool AREACOUNT = false, bool DEFER = false>
__global__ void __launch_bounds__(RT_FAST_BLOCK, BLOCKS) renderKernelFast(DeviceScene S, CameraParams cam, RenderTargets out) {
    static_assert(RT_FAST_BLOCK == 32, "one warp per block: the packet stack and the colour buffer are the warp's");
    __shared__ int2 wstacks[RT_WARP_STACK_INT2];
    __shared__ double colBuf[RECURSE ? RT_BATCH_PASSES * 96 : 1];     // per-sample sums of the batch in flight
    const int lane = threadIdx.x & 31;
    TraceScratch T;
    T.sstack = nullptr; T.sstackN = 0;   // secondary rays keep their private stack in local memory: shared memory stays L1
    const int warpInBlock = 0;
    T.wstack = wstacks;
    T.pathStack = out.pathStack;
    T.gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    T.stride = (size_t)gridDim.x * blockDim.x;

    Counters C;
    C.box = C.tri = C.sphere = C.plane = C.inst = C.smooth = 0; C.maxStack = 0;
    RayCounts R;
    R.primary = R.shadow = R.secondary = 0;

    const int W = cam.width, H = cam.height;
    const ShardTiles tiles(cam.width, cam.height, cam.shardIndex, cam.shardCount, cam.shardTiles);
    const int n = cam.n, ns = n * n;                 // samples traced per pixel (:298-306)
    const int drawsPerSample = cam.dof ? 5 : 3;      // xi1, xi2, [lens u, v], time
    const double samplesD = AREACOUNT ? 1.0 : (double)cam.samples;     // the sum is divided by `samples` (:354); a count is not
    const bool segmented = ns < 32 && (32 % ns) == 0;
    const V3 ambient = ld3(S.ambient), background = ld3(S.background);
    // sample s of a lane is the same in every pass when n*n divides 32: its skip-ahead factors are computed once
    __shared__ ulonglong2 skipTab[RT_FAST_BLOCK];
    {
        uint64_t ak = 1, gk = 0;
        if (segmented) PCG32::skipFactors((uint64_t)((lane % ns) * drawsPerSample), ak, gk);
        skipTab[threadIdx.x] = make_ulonglong2(ak, gk);
        __syncwarp();
    }
    // Work unit (gpu_layout.h: FastUnits): passes of 32 (pixel, sample) items that are complete in themselves.
    const FastUnits units(ns, RECURSE);
    const int unitsPerTile = units.perTile;
    const int totalUnits = tiles.count * unitsPerTile;
    __shared__ uint2 unitNote[RT_FAST_BLOCK / 32];               // (clock at unit start, tile): read back when the unit ends
    WarpQueue Q;
    Q.init();
#ifdef RT_WARPTIMES
    unsigned long long wt0 = globalTimerNs(), wt1 = 0, wt2 = 0;
    int wtUnits = 0, wtUnit = 0;
#endif

    for (;;) {
#ifdef RT_WARPTIMES
        if (wtUnits == 1) wt1 = globalTimerNs();
        if (wtUnits > 0 && lane == 0 && out.warpTimes) {    // [4 x warps] then (start, end, visits, tests) per unit
            unsigned long long* u = out.warpTimes + 4 * (T.stride >> 5) + 4 * (size_t)wtUnit;
            const int2 a = T.wstack[RT_WSTACK + RT_WTSTACK], b = T.wstack[RT_WSTACK + RT_WTSTACK + 1];
            u[1] = globalTimerNs();
            u[2] = ((unsigned long long)(unsigned)a.x << 32) | (unsigned)a.y;
            u[3] = ((unsigned long long)(unsigned)b.x << 32) | (unsigned)b.y;
        }
        if (lane == 0) T.wstack[RT_WSTACK + RT_WTSTACK] = T.wstack[RT_WSTACK + RT_WTSTACK + 1] = make_int2(0, 0);
        __syncwarp();
        wt2 = globalTimerNs();
        wtUnits++;
#endif
        // Small units keep the tail short when a frame is split over 8 GPUs (a few thousand warps share ~8k tiles
        // per rank).  One unit per draw: consecutive units go to warps running at the same time, which share tiles,
        // BVH nodes and triangles in L2 (drawing 8 at a time measured 10 % slower).
        int unit = 0;
        if (lane == 0) unit = atomicAdd(out.tileCounter, 1);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        const bool fromList = RECURSE && out.deferPhase == 2;           // second launch: the units the first one deferred
        if (unit >= (fromList ? *out.deferCount : totalUnits)) break;
        if (fromList) unit = out.deferList[unit];                       // (tile, pass) as the first launch resolved it
        // Slot -> tile through the schedule of this frame: tiles in descending order of the cost measured on the
        // previous frame (longest first), so the frame ends on its cheapest units instead of a few slow ones.
        const int slot = unit / unitsPerTile;
        const int tile = (out.tileOrder && !fromList) ? __ldg(out.tileOrder + slot) : slot;
        const RayCounts R0 = R;
        const Counters C0 = C;
        bool deferred = false;
        if (lane == 0) unitNote[warpInBlock] = make_uint2((unsigned)clock(), (unsigned)tile);
#ifdef RT_WARPTIMES
        wtUnit = tile * unitsPerTile + (unit - slot * unitsPerTile);     // records are indexed by (tile, pass)
        if (lane == 0 && out.warpTimes) out.warpTimes[4 * (T.stride >> 5) + 4 * (size_t)wtUnit] = wt2;
#endif
        const int sub = unit - slot * unitsPerTile;
        const int itBegin = sub * units.passes;
        const int itEnd = itBegin + units.passes;
        int x0, y0;
        tiles.origin(tile, x0, y0);
        int openP = -1;               // ordered accumulation state of the generic path (warp-uniform)
        V3 acc = mk(0, 0, 0);
#pragma unroll 1
        for (int b0 = itBegin; b0 < itEnd; b0 += (RECURSE ? RT_BATCH_PASSES : 1)) {
            const int nb = RECURSE ? (itEnd - b0 < RT_BATCH_PASSES ? itEnd - b0 : RT_BATCH_PASSES) : 1;
            V3 c = mk(0, 0, 0);
            // ---- one round = 32 rays through ONE copy of the code: the primary rays of a pass (ray generation :292-344), or —
            //      scenes with recursion — 32 child rays off the warp's queue; closest-hit packet, shading, one any-hit packet
            //      per light (:118-143), children into the queue (rt_shade.cuh: WarpQueue)
            int pass = 0;
#pragma unroll 1
            for (;;) {
                const bool primary = pass < nb;                  // warp-uniform
                PathEntry cur0;
                bool active = false;
                double tMin0 = 0.0;
                int idI = 0, idJ = 0, idS = 0;
                PCG32 rng;
                rng.state = 0; rng.inc = 1;
                if (primary) {
                    const int it = b0 + pass;
                    const int g = it * 32 + lane;
                    const int p = g / ns, s = g - p * ns;
                    // pixel p of the tile in Z-order: neighbouring lanes' pixels form compact blocks
                    const int px = (p & 1) | ((p >> 1) & 2) | ((p >> 2) & 4), py = ((p >> 1) & 1) | ((p >> 2) & 2);
                    const int i = x0 + px, j = y0 + py;
                    active = i < W && j < H;
                    const int sy = s / n, sx = s - sy * n;
                    rng.seed((((uint64_t)j << 32) ^ (uint64_t)i) + 0x9E3779B97F4A7C15ULL);  // :294 (Swift precedence)
                    if (RT_SKIPTAB && segmented) rng.skip(skipTab[threadIdx.x].x, skipTab[threadIdx.x].y);   // s is fixed per lane
                    else rng.advance((uint64_t)(s * drawsPerSample));
                    const PrimaryRay pr = primaryRay(cam, i, j, sx, sy, rng);
                    if (active) R.primary++;
                    cur0.o = pr.o; cur0.d = pr.d; cur0.w = mk(1, 1, 1); cur0.beerMat = 0; cur0.depth = 0; cur0.beer = 0;
                    cur0.slot = pass * 32 + lane;
                    tMin0 = pr.tMin;
                    idI = i; idJ = j; idS = s;
                } else {
                    if (!RECURSE) break;
                    cur0.o = cur0.d = cur0.w = mk(0, 0, 0); cur0.beerMat = 0; cur0.depth = 0; cur0.beer = 0; cur0.slot = 0;
                    if (!Q.pop(out.pathStack, S.maxDepth, cur0, active)) break;
                }

                // ---- the rays of the round as a warp packet: one traversal call site serves the closest-hit query
                //      (phase 0) and the shadow ray towards light phase-1
                V3 ro = cur0.o, rd = cur0.d, wi = mk(0, 0, 1), Lo = mk(0, 0, 0);
                double rtMin = tMin0, dist = 0;
                HitRec hit;
                hit.t = CUDART_INF; hit.u = hit.v = 0; hit.inst = -1; hit.slot = 0;
                ShadePoint sp;
                sp.p = sp.N = sp.w = mk(0, 0, 0); sp.mat = S.materials; sp.matId = 0; sp.frontFacing = sp.computeDirect = false;
                bool rayOn = active, anyhit = false, isHit = false;
                int phase = 0;
                __syncwarp();
#pragma unroll 1
                for (;;) {
                    const bool occ = packetScene<COUNT, RECURSE>(S, ro, rd, rtMin, hit, rayOn, anyhit, T.wstack, C, !primary);
                    if (phase == 0) {
                        isHit = active && hit.inst >= 0;
                        Surface sf;
                        sf.kind = 0; sf.prim = -1;
                        if (isHit) {
                            resolveHit(S, hit, cur0.o, cur0.d, sf);
                            if (COUNT && sf.kind == 1 && S.instances[hit.inst].smooth) C.smooth++;
                            setupShade(S, sf, cur0, hit.t, sp);
                            if (sp.computeDirect) Lo = ambient * ld3(sp.mat->ambient);
                        }
                        if (!AREACOUNT && primary && out.ids && active) {
                            HitId& h = out.ids[((size_t)idJ * W + idI) * (size_t)ns + (size_t)idS];
                            h.inst = hit.inst; h.kind = sf.kind; h.prim = sf.prim; h.t = hit.t;
                        }
                        if (AREACOUNT) break;                                       // counting pass: no shadow rays
                        if (DEFER && __ballot_sync(0xffffffffu, isHit && sp.mat->type != 0) != 0u) { deferred = true; break; }    // type 0 = plain (rt_shade.cuh: spawnChildren)
                        if (__ballot_sync(0xffffffffu, sp.computeDirect) == 0u) break;
                    } else if (rayOn && !occ) {
                        Lo = Lo + pointLightTerm(*sp.mat, S.pointLights[phase - 1], sp.N, cur0.d, wi, dist);
                    }
                    if (phase >= S.nPointLights) break;
                    shadowRay(S.pointLights[phase], sp.p, S.shadowEps, ro, wi, dist);
                    rd = wi; rtMin = 0.0;
                    hit.t = dist;                                   // tMax = dist (:124)
                    rayOn = sp.computeDirect; anyhit = true;
                    if (rayOn) R.shadow++;
                    phase++;
                }
                // ---- what the rays of this round add to their samples (:345): miss -> background (:101-103), else the
                //      direct light of the vertex; no non-finite guard here (see WarpQueue)
                c = mk(0, 0, 0);
                if (AREACOUNT) { if (active && isHit && sp.computeDirect) c.x = (double)S.nAreaLights; }
                else if (active) c = isHit ? sp.w * Lo : cur0.w * background;
                if (!RECURSE) break;
                PathEntry c1 = cur0, c2 = cur0;
                int nc = 0;
                if (isHit) nc = spawnChildren<false>(S, sp, cur0, rng, c1, c2);
                R.secondary += nc;
                __syncwarp();
                if (primary) {
                    double* col = colBuf + cur0.slot * 3;
                    col[0] = c.x; col[1] = c.y; col[2] = c.z;
                    __syncwarp();
                    pass++;
                } else {
                    warpAccumulate(colBuf, cur0.slot, active, c);
                }
                Q.push(out.pathStack, S.maxDepth, nc, c1, c2);
            }
            __syncwarp();
            if (DEFER && deferred) break;       // a primary hit of this unit spawns child rays: the second launch traces the unit
            // ---- pixelColor += trace(...) in sample order (:345), then / samples (:354), pass by pass
#pragma unroll 1
            for (int b = 0; b < nb; b++) {
                const int it = b0 + b;
                const int g = it * 32 + lane;
                const int p = g / ns, s = g - p * ns;
                const int px = (p & 1) | ((p >> 1) & 2) | ((p >> 2) & 4), py = ((p >> 1) & 1) | ((p >> 2) & 2);
                const int i = x0 + px, j = y0 + py;
                const bool active = i < W && j < H;
                static_assert(RT_WARP_STACK_INT2 * sizeof(int2) >= 96 * sizeof(double), "sample colours of a warp fit the packet stack");
                double* scratch = RECURSE ? colBuf + b * 96 : reinterpret_cast<double*>(T.wstack);   // the packet stack is idle here
                if (RECURSE) c = mk(scratch[lane * 3], scratch[lane * 3 + 1], scratch[lane * 3 + 2]);
                __syncwarp();
                if (segmented && ns >= 4) {
                    // The colours go through shared memory and lanes s = 0, 1, 2 of each pixel add up one component
                    // each, in sample order: 16 loads + adds instead of every lane shuffling all three components of
                    // all samples.
                    const int base = lane & ~(ns - 1);
                    if (!RECURSE) { scratch[lane * 3] = c.x; scratch[lane * 3 + 1] = c.y; scratch[lane * 3 + 2] = c.z; }
                    __syncwarp();
                    double comp = 0.0;
                    if (s < 3) {
                        const double* col = scratch + base * 3 + s;
#pragma unroll 4
                        for (int k = 0; k < ns; k++) comp = comp + col[k * 3];
                        comp = ddiv(comp, samplesD);
                    }
                    const V3 px3 = mk(shflD(comp, base), shflD(comp, base + 1), shflD(comp, base + 2));
                    if (active && s == 0) emitPixel<AREACOUNT>(out, (size_t)j * W + i, px3);
                    __syncwarp();
                } else if (segmented) {
                    const int base = lane & ~(ns - 1);
                    V3 sum = mk(0, 0, 0);
#pragma unroll 1
                    for (int k = 0; k < ns; k++) {
                        const V3 ck = mk(shflD(c.x, base + k), shflD(c.y, base + k), shflD(c.z, base + k));
                        sum = sum + ck;
                    }
                    if (active && s == 0) emitPixel<AREACOUNT>(out, (size_t)j * W + i, sum / samplesD);
                } else if (units.perPixel) {
                    // 32 divides ns: the 32 samples of this pass are consecutive samples of ONE pixel.  Lanes 0..2 add one
                    // component each, in sample order, carrying their sum from pass to pass (acc.x in all three).
                    if (!RECURSE) { scratch[lane * 3] = c.x; scratch[lane * 3 + 1] = c.y; scratch[lane * 3 + 2] = c.z; }
                    __syncwarp();
                    if (lane < 3) {
                        double comp = acc.x;
#pragma unroll 8
                        for (int k = 0; k < 32; k++) comp = comp + scratch[k * 3 + lane];
                        acc.x = comp;
                    }
                    __syncwarp();
                    if (it == itEnd - 1) {
                        const double comp = ddiv(acc.x, samplesD);
                        const V3 px3 = mk(shflD(comp, 0), shflD(comp, 1), shflD(comp, 2));
                        if (active && lane == 0) emitPixel<AREACOUNT>(out, (size_t)j * W + i, px3);
                    }
                } else {
#pragma unroll 1
                    for (int k = 0; k < 32; k++) {
                        const int pk = (it * 32 + k) / ns;
                        const V3 ck = mk(shflD(c.x, k), shflD(c.y, k), shflD(c.z, k));
                        if (pk != openP) {
                            if (openP >= 0 && lane == 0) {
                                const int qx = (openP & 1) | ((openP >> 1) & 2) | ((openP >> 2) & 4), qy = ((openP >> 1) & 1) | ((openP >> 2) & 2);
                                if (x0 + qx < W && y0 + qy < H) emitPixel<AREACOUNT>(out, (size_t)(y0 + qy) * W + (x0 + qx), acc / samplesD);
                            }
                            openP = pk;
                            acc = mk(0, 0, 0);
                        }
                        acc = acc + ck;
                    }
                }
            }
        }
        if (DEFER && deferred) {                // not finished here: its rays are not counted, its unit goes on the list
            R = R0; C = C0;
            if (lane == 0) out.deferList[atomicAdd(out.deferCount, 1)] = tile * unitsPerTile + sub;
        } else if (!segmented && !units.perPixel && openP >= 0 && lane == 0) {
            const int qx = (openP & 1) | ((openP >> 1) & 2) | ((openP >> 2) & 4), qy = ((openP >> 1) & 1) | ((openP >> 2) & 2);
            if (x0 + qx < W && y0 + qy < H) emitPixel<AREACOUNT>(out, (size_t)(y0 + qy) * W + (x0 + qx), acc / samplesD);
        }
        if (out.tileCost && lane == 0) {       // cost of the unit in clocks/64, summed per tile for the next frame's schedule
            const uint2 note = unitNote[warpInBlock];
            atomicAdd(out.tileCost + note.y, ((unsigned)clock() - note.x) >> 6);
        }
    }
#ifdef RT_WARPTIMES
    if (lane == 0 && out.warpTimes) {
        unsigned long long* w = out.warpTimes + 4 * (T.gtid >> 5);
        w[0] = wt0; w[1] = wt1; w[2] = wt2; w[3] = globalTimerNs();
    }
#endif
    if (AREACOUNT) return;                     // the counting pass reports no rays: pass 2 traces them all again
    flushCounters<COUNT>(out, C, R);
}

// EXT: scenes with area lights (RenderTargets::areaIndex / areaPass: pass 1 leaves each pixel's number of
// jitter-table draws, pass 2 starts each pixel at its index within the 8-row chunk) and / or motion blur (ray time
// carried down the path).  A second instantiation, so that scenes without them run exactly the code they always ran.
template <bool COUNT, bool EXT>
__global__ void __launch_bounds__(RT_BLOCK, RT_GEN_BLOCKS) renderKernelGeneral(DeviceScene S, CameraParams cam, RenderTargets out) {
    __shared__ int32_t sstack[RT_SSTACK * RT_BLOCK];
    const int lane = threadIdx.x & 31;
    TraceScratch T;
    T.sstack = sstack; T.sstackN = RT_SSTACK;
    T.wstack = nullptr;
    T.pathStack = out.pathStack;
    T.gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    T.stride = (size_t)gridDim.x * blockDim.x;

    Counters C;
    C.box = C.tri = C.sphere = C.plane = C.inst = C.smooth = 0; C.maxStack = 0;
    RayCounts R;
    R.primary = R.shadow = R.secondary = 0;

    const int W = cam.width, H = cam.height;
    const ShardTiles tiles(cam.width, cam.height, cam.shardIndex, cam.shardCount, cam.shardTiles);
    const int n = cam.n, samples = cam.samples;

    for (;;) {
        int slot = 0;
        if (lane == 0) slot = atomicAdd(out.tileCounter, 1);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= tiles.count) break;
        const int tile = out.tileOrder ? __ldg(out.tileOrder + slot) : slot;    // longest tiles of the previous frame first
        const unsigned c0 = (unsigned)clock();
        int x0, y0;
        tiles.origin(tile, x0, y0);
        const int i = x0 + (lane & 7), j = y0 + (lane >> 3);
        // lanes outside the image idle until the warp's barrier below (no `continue`: the full-mask shuffle at the
        // top of the loop must be reached by all 32 lanes together)
        if (i < W && j < H) {
            PCG32 rng;
            rng.seed((((uint64_t)j << 32) ^ (uint64_t)i) + 0x9E3779B97F4A7C15ULL);  // :294 (Swift precedence)
            V3 pixel = mk(0, 0, 0);
            int sampleIndex = 0;
            AreaState area;
            area.jitterIndex = 0; area.countOnly = false;
            if (EXT && out.areaPass != 0) {
                area.countOnly = out.areaPass == 1;
                if (!area.countOnly) area.jitterIndex = out.areaIndex[(size_t)j * W + i];
            }
            for (int sy = 0; sy < n && sampleIndex < samples; sy++) {
                for (int sx = 0; sx < n && sampleIndex < samples; sx++) {
                    const PrimaryRay pr = primaryRay(cam, i, j, sx, sy, rng);
                    R.primary++;
                    HitId* idSlot = (out.ids && !area.countOnly) ? out.ids + (((size_t)j * W + i) * (size_t)(n * n) + (size_t)(sy * n + sx)) : nullptr;
                    PathEntry cur0;
                    cur0.o = pr.o; cur0.d = pr.d; cur0.w = mk(1, 1, 1); cur0.beerMat = 0; cur0.depth = 0; cur0.beer = 0; cur0.slot = 0;
                    pixel = pixel + scalarPathLoop<true, COUNT, EXT, EXT>(S, T, cur0, 0, true, pr.tMin, idSlot, rng, C, R, mk(0, 0, 0),
                                                                                 EXT ? &area : nullptr, pr.time);
                    sampleIndex++;
                }
            }
            if (EXT && area.countOnly) out.areaIndex[(size_t)j * W + i] = area.jitterIndex;
            else writePixel(out, (size_t)j * W + i, pixel / (double)samples);                              // :354
        }
        __syncwarp();
        if (out.tileCost && lane == 0) atomicAdd(out.tileCost + tile, ((unsigned)clock() - c0) >> 6);
    }
    if (EXT && out.areaPass == 1) return;      // the counting pass reports no rays: pass 2 traces them all again
    flushCounters<COUNT>(out, C, R);
}

}  // namespace rtg
