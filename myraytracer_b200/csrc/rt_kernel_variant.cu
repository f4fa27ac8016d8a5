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
// the pinhole ray through the tile's centre, no random numbers — and its shadow rays, per lane, with the counters on:
// box tests + weighted triangle tests are the tile's cost estimate, launchSchedule turns the estimates into the order
// of this very frame (grazing rays, whose packets visit hundreds of nodes, go first; sky tiles last).  65 k rays for
// a 1920x1080 frame: microseconds.  The frame itself then records the real costs for the frames that follow.
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
        const double tImg = dot(cm - eye, cw) / (denomW == 0.0 ? 4.9406564584124654e-324 : denomW);
        Counters C;
        C.box = C.tri = C.sphere = C.plane = C.inst = C.smooth = 0; C.maxStack = 0;
        HitRec hit;
        hit.t = CUDART_INF; hit.u = hit.v = 0; hit.inst = -1; hit.slot = 0;
        traceScene<true>(S, eye, d, dmax(tImg, 0.0), hit, false, nullptr, 0, C);
        if (hit.inst >= 0) {
            const V3 p = eye + d * hit.t;
            for (int l = 0; l < S.nPointLights; l++) {                              // shadow rays: :118-143 (from the hit point, no normal offset needed for a cost)
                const V3 toL = ld3(S.pointLights[l].position) - p;
                const double dist = length(toL);
                HitRec sh;
                sh.t = dist; sh.u = sh.v = 0; sh.inst = -1; sh.slot = 0;
                traceScene<true>(S, p, toL / dist, S.shadowEps, sh, true, nullptr, 0, C);
            }
            cost += 16;                                                             // shading of a hit
        }
        cost += (unsigned)C.box + 8u * (unsigned)(C.tri + C.sphere + C.plane) + 4u * (unsigned)C.inst;
    }
    tileCost[tile] = cost;
}

cudaError_t launchTileProbe(const DeviceScene& S, const CameraParams& cam, unsigned* tileCost, int nTiles, cudaStream_t stream) {
    if (nTiles > 0) RT_LAUNCH(tileProbeKernel, (nTiles + 127) / 128, 128, stream, S, cam, tileCost, nTiles);
    return cudaGetLastError();
}
#endif

}  // namespace rtg
