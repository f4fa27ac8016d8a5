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

}  // namespace rtg
