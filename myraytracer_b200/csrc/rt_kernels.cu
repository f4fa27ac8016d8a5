// rt_kernels.cu — dispatcher over the two render kernels (rt_kernel_variant.cu, rt_render_kernel.cuh).
//
//   KV_FAST     lane = (pixel, sample); depth 0 (primary + shadow rays) as warp packets; valid when no
//               material draws random numbers inside `trace` (roughness == 0): the default
//   KV_GENERAL  lane = pixel, sequential PCG32 stream, per-lane traversal: any scene
#include "rt_kernels.h"

namespace rtg {

cudaError_t launchVariantFast(const DeviceScene&, const CameraParams&, const RenderTargets&, int, bool, cudaStream_t);
cudaError_t launchVariantGeneral(const DeviceScene&, const CameraParams&, const RenderTargets&, int, bool, cudaStream_t);
cudaError_t occupancyVariantFast(int*, bool);
cudaError_t occupancyVariantGeneral(int*, bool);

cudaError_t launchRender(int variant, const DeviceScene& S, const CameraParams& cam, const RenderTargets& out, int blocks,
                         bool counters, cudaStream_t stream) {
    if (variant == KV_FAST) return launchVariantFast(S, cam, out, blocks, counters, stream);
    return launchVariantGeneral(S, cam, out, blocks, counters, stream);
}

cudaError_t renderOccupancy(int variant, int* blocksPerSm, bool counters) {
    if (variant == KV_FAST) return occupancyVariantFast(blocksPerSm, counters);
    return occupancyVariantGeneral(blocksPerSm, counters);
}

int renderBlockSize() { return RT_BLOCK; }
int renderStackCapacity() { return RT_STACK; }

}  // namespace rtg
