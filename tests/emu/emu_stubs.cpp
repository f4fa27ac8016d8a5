// tests/emu/emu_stubs.cpp — TEST INFRASTRUCTURE ONLY: parts of the product that the host emulation cannot run.
// The device BVH builder (csrc/bvh_build_gpu.cu) is built on cub's device-wide radix sort; the emulation build
// reports it as unsupported, so rt_scene_create falls back to the host SAH builder for every mesh.
#include <cuda_runtime.h>

#include "bvh_build.h"

namespace rtg {
cudaError_t buildBlasDevice(const double*, const BlasBuildInput&, BlasBuildOutput& out) {
    out.maxDepth = 1 << 30;     // "too deep": the caller's documented fallback to buildBlas
    return cudaSuccess;
}
}  // namespace rtg
