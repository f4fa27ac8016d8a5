// rt_kernel_variant.cu — one render kernel per object file (build time): -DRT_V_FAST=1 compiles
// renderKernelFast, -DRT_V_FAST=0 renderKernelGeneral; rt_kernels.cu dispatches.   COMPILE WITH -fmad=false.
#include "rt_render_kernel.cuh"

#ifndef RT_V_FAST
#error "define RT_V_FAST"
#endif

namespace rtg {

#if RT_V_FAST
#define RT_KERNEL(c) renderKernelFast<c, false>
#define RT_KERNEL_RECURSE(c) renderKernelFast<c, true>
#ifndef RT_FAST_BLOCKS_SMALL
#define RT_FAST_BLOCKS_SMALL 12    // B200, C3: 8 warps 6.15 ms, 12 5.71, 16 6.07, 20 6.27, 32 6.49
#endif
#define RT_KERNEL_RECURSE_SMALL(c) renderKernelFast<c, true, RT_FAST_BLOCKS_SMALL>
#define RT_NAME(x) x##Fast
#define RT_THREADS RT_FAST_BLOCK
#else
#define RT_KERNEL(c) renderKernelGeneral<c, false>
#define RT_KERNEL_EXT(c) renderKernelGeneral<c, true>
#define RT_NAME(x) x##General
#define RT_THREADS RT_BLOCK
#endif

cudaError_t RT_NAME(launchVariant)(const DeviceScene& S, const CameraParams& cam, const RenderTargets& out, int blocks,
                                   bool counters, cudaStream_t stream) {
#ifdef RT_KERNEL_EXT
    if (out.areaPass != 0 || out.motion) {     // area lights / motion blur: the instantiation that carries them
        if (counters) RT_LAUNCH(RT_KERNEL_EXT(true), blocks, RT_THREADS, stream, S, cam, out);
        else RT_LAUNCH(RT_KERNEL_EXT(false), blocks, RT_THREADS, stream, S, cam, out);
        return cudaGetLastError();
    }
#endif
#ifdef RT_KERNEL_RECURSE
    if (out.areaPass == 1) {                   // counting pass of an area-light frame (no counters: it reports no rays)
        if (out.recurse == 2) RT_LAUNCH((renderKernelFast<false, true, RT_FAST_BLOCKS_SMALL, true>), blocks, RT_THREADS, stream, S, cam, out);
        else if (out.recurse) RT_LAUNCH((renderKernelFast<false, true, RT_FAST_BLOCKS, true>), blocks, RT_THREADS, stream, S, cam, out);
        else RT_LAUNCH((renderKernelFast<false, false, RT_FAST_BLOCKS, true>), blocks, RT_THREADS, stream, S, cam, out);
        return cudaGetLastError();
    }
    if (out.deferPhase == 1) {                 // first of two launches: depth 0 in the lean code, units with child rays deferred
        if (counters) RT_LAUNCH((renderKernelFast<true, false, RT_FAST_BLOCKS, false, true>), blocks, RT_THREADS, stream, S, cam, out);
        else RT_LAUNCH((renderKernelFast<false, false, RT_FAST_BLOCKS, false, true>), blocks, RT_THREADS, stream, S, cam, out);
        return cudaGetLastError();
    }
    if (out.recurse == 2) {                    // ... and no big mesh: RT_FAST_BLOCKS_SMALL resident warps per SM
        if (counters) RT_LAUNCH(RT_KERNEL_RECURSE_SMALL(true), blocks, RT_THREADS, stream, S, cam, out);
        else RT_LAUNCH(RT_KERNEL_RECURSE_SMALL(false), blocks, RT_THREADS, stream, S, cam, out);
        return cudaGetLastError();
    }
    if (out.recurse) {                         // child rays possible: the instantiation with the warp's ray queue
        if (counters) RT_LAUNCH(RT_KERNEL_RECURSE(true), blocks, RT_THREADS, stream, S, cam, out);
        else RT_LAUNCH(RT_KERNEL_RECURSE(false), blocks, RT_THREADS, stream, S, cam, out);
        return cudaGetLastError();
    }
#endif
    if (counters) RT_LAUNCH(RT_KERNEL(true), blocks, RT_THREADS, stream, S, cam, out);
    else RT_LAUNCH(RT_KERNEL(false), blocks, RT_THREADS, stream, S, cam, out);
    return cudaGetLastError();
}

// resident blocks per SM: the smallest over the instantiations a scene may take (one grid size serves them all)
cudaError_t RT_NAME(occupancyVariant)(int* blocksPerSm, bool counters) {
    int a = 0, b = 0;
    cudaError_t e = counters ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, RT_KERNEL(true), RT_THREADS, 0)
                             : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, RT_KERNEL(false), RT_THREADS, 0);
    if (e != cudaSuccess) return e;
#if defined(RT_KERNEL_RECURSE)
    e = counters ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, RT_KERNEL_RECURSE(true), RT_THREADS, 0)
                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, RT_KERNEL_RECURSE(false), RT_THREADS, 0);
#elif defined(RT_KERNEL_EXT)
    e = counters ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, RT_KERNEL_EXT(true), RT_THREADS, 0)
                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, RT_KERNEL_EXT(false), RT_THREADS, 0);
#endif
    if (e != cudaSuccess) return e;
    *blocksPerSm = a < b ? a : b;
    return cudaSuccess;
}

int RT_NAME(blockSizeVariant)() { return RT_THREADS; }

#if !RT_V_FAST
// First frame of a geometry: nothing has measured the tiles yet, and in the spatial order a frame split over 8 GPUs
// ends on whatever slow units happen to be drawn last.  The probe traces ONE ray per 8x4 tile —
// the pinhole ray through the tile's centre, no random numbers, per lane, with a budget of RT_PROBE_VISITS node visits
// (a lone warp pays 0.6 us per visit, and a grazing ray of the terrain makes hundreds: without the budget the probe
// took 0.4 ms, more than the order it produces gains) — and the visits it made, plus a constant when it hit something
// (shading and shadow rays follow) or ran out of budget, are the tile's cost estimate.  launchSchedule turns the
// estimates into the order of this very frame: grazing rays first, sky tiles last.  64,800 rays for a 1920x1080
// frame.  The frame itself then records the real costs for the frames that follow.
#ifndef RT_PROBE_VISITS
#define RT_PROBE_VISITS 48
#endif
__global__ void __launch_bounds__(128) tileProbeKernel(DeviceScene S, CameraParams cam, unsigned* tileCost, int nTiles) {
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= nTiles) return;
    const ShardTiles tiles(cam.width, cam.height, cam.shardIndex, cam.shardCount, cam.shardTiles);
    int x0, y0;
    tiles.origin(tile, x0, y0);
    unsigned cost = 1;
    if (x0 < cam.width && y0 < cam.height) {
        const V3 eye = ld3(cam.eye), cu = ld3(cam.u), cv = ld3(cam.v), cw = ld3(cam.w), q00 = ld3(cam.q00), cm = ld3(cam.m);
        const double ci = (double)x0 + 4.0, cj = (double)y0 + 2.0;                 // centre of the 8x4 block
        const V3 s = (q00 - cv * (cj * cam.dv)) + cu * (ci * cam.du);              // Object+Extension.swift:311-318
        const V3 d = normalize(s - eye);
        const double denomW = dot(d, cw);
        const double tImg = ddiv(dot(cm - eye, cw), denomW == 0.0 ? 4.9406564584124654e-324 : denomW);
        Counters C;
        C.box = C.tri = C.sphere = C.plane = C.inst = C.smooth = 0; C.maxStack = 0;
        HitRec hit;
        hit.t = CUDART_INF; hit.u = hit.v = 0; hit.inst = -1; hit.slot = 0;
        int budget = RT_PROBE_VISITS;
        traceScene<true, false, true>(S, eye, d, dmax(tImg, 0.0), hit, false, nullptr, 0, C, 0.0, &budget);
        const int visits = RT_PROBE_VISITS - (budget < 0 ? 0 : budget);
        cost += (unsigned)visits + 4u * (unsigned)C.inst;
        if (hit.inst >= 0 || budget < 0) cost += 2u * (unsigned)visits + 16u;     // shading and one shadow packet per light follow
    }
    tileCost[tile] = cost;
}

cudaError_t launchTileProbe(const DeviceScene& S, const CameraParams& cam, unsigned* tileCost, int nTiles, cudaStream_t stream) {
    if (nTiles > 0) RT_LAUNCH(tileProbeKernel, (nTiles + 127) / 128, 128, stream, S, cam, tileCost, nTiles);
    return cudaGetLastError();
}
#endif

}  // namespace rtg
