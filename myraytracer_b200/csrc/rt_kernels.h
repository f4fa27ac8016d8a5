// rt_kernels.h — host-visible interface of rt_kernels.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "gpu_layout.h"

namespace rtg {

struct HitId {  // same layout as rt_hit_id (include/rtcuda.h)
    int32_t inst, kind;
    int64_t prim;
    double t;
};

struct RenderTargets {
    uint8_t* rgba;                 // W*H*4 or null
    double* linear;                // W*H*3 or null
    HitId* ids;                    // W*H*n*n or null
    double* pathStack;             // (maxDepth+1) * PATH_FIELDS * totalThreads doubles
    int* tileCounter;              // zeroed before launch
    unsigned long long* counters;  // 16 x u64, zeroed before launch
    const int* tileOrder;          // schedule: slot -> tile of this shard (null = identity)
    unsigned* tileCost;            // per tile, clocks/64 spent on it this frame (null = not recorded); zeroed before launch
    // Area lights (general kernel only).  The reference draws the light sample of every shading point from a 100-entry
    // table through ONE index per 8-row chunk that counts up in pixel, sample and recursion order
    // (Object+Extension.swift:152-154,288).  areaPass 1 walks the paths without shading and leaves each pixel's number
    // of draws (mod 100) in areaIndex[pixel]; launchAreaPrefix turns them into each pixel's first index (exclusive
    // prefix over the chunk, row-major); areaPass 2 renders with it.  areaPass 0: scene without area lights.
    unsigned* areaIndex;           // W*H or null
    int areaPass;
    int motion;                    // 1: the scene has motion blur (general kernel, ray time carried down the path)
    int recurse;                   // 1: child rays are possible (fast kernel: the instantiation with the warp's ray queue);
                                   // 2: ... and the scene has no big mesh (the same, compiled for 12 warps of 168 registers per SM)
    // Two-launch frames of scenes with rare child rays (rt_render_kernel.cuh: DEFER): 1 = the lean instantiation, which
    // appends the units (tile * unitsPerTile + pass) whose primary hits spawn children to deferList; 2 = the queue
    // instantiation drawing its units from that list; 0 = one launch.
    int deferPhase;
    int* deferList;
    int* deferCount;
#ifdef RT_WARPTIMES
    unsigned long long* warpTimes; // profiling build only: 4 x u64 per warp (start, first unit done, last draw, exit; ns)
#endif
};

static const int kPathFields = 11;

// Kernel launch.  RT_EMU: the host emulation build of tests/emu (test infrastructure only, never part of
// librtcuda.so) runs the same kernel body on fibers.
#ifdef RT_EMU
#define RT_LAUNCH(kernel, grid, block, stream, ...) rtemu::launch(dim3((unsigned)(grid)), dim3((unsigned)(block)), [&]() { kernel(__VA_ARGS__); })
#else
#define RT_LAUNCH(kernel, grid, block, stream, ...) kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#endif

#ifndef RT_BLOCK
#define RT_BLOCK 128
#endif
#ifndef RT_STACK
#define RT_STACK 64   // same capacity as the reference's fixed stacks (RTContext.swift:550,626)
#endif
#ifndef RT_SSTACK
#define RT_SSTACK 24  // entries of the per-lane stack kept in shared memory; the rest spill to local memory
#endif

#ifndef RT_BSTACK
#define RT_BSTACK 127 // entries of the beam walk's stack (rt_beam.cuh) besides its sentinel; it shares the packet stack's memory
#endif

// kernel variants (rt_kernels.cu); values match rt_render_params.kernelVariant - 1
enum KernelVariant { KV_FAST = 0, KV_GENERAL = 1, KV_COUNT = 2 };

cudaError_t launchRender(int variant, const DeviceScene& S, const CameraParams& cam, const RenderTargets& out, int blocks,
                         bool counters, cudaStream_t stream);
// tileCost (clocks per tile of the frame just rendered) -> tileOrder for the next frame, longest first; clears tileCost
// areaIndex: per-pixel draw counts -> per-pixel first jitter index, for the 8-row chunks of this shard (rows mode)
cudaError_t launchAreaPrefix(unsigned* areaIndex, int width, int height, int shardIndex, int shardCount, cudaStream_t stream);
// Animation: bottom-up refit of the TLAS after rt_scene_update_instances (topology kept, like BVHBuilder.refit,
// Accelearion/BVH.swift:254-292): FP64 boxes in tlasD, conservative FP32 encoding (centre, padding P) in tlasOut.
// Instance.worldBounds of mesh instances on the device: the union over the BLAS's triangles of their transformed
// boxes, exactly the reference's makeInstance / refitTLAS loop (RTContext.swift:437-447,891-900; AABB.transformed,
// AABB.swift:71-92).  chunks: (instance, first triangle slot, count, f64) work items made at scene creation;
// boundsEnc: 6 order-preserving u64 per instance (encodeOrdered), pre-set to (+inf, -inf) for the instances to derive
// and left alone for the others (derive[i] == 0).
struct BoundsChunk { int32_t inst, first, count, f64; };
cudaError_t launchInstanceBounds(const DeviceScene& S, const BoundsChunk* chunks, int nChunks, const int32_t* derive,
                                 unsigned long long* boundsEnc, cudaStream_t stream);
inline unsigned long long encodeOrdered(double d) {
    unsigned long long b;
    memcpy(&b, &d, 8);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
inline double decodeOrdered(unsigned long long u) {
    const unsigned long long b = (u >> 63) ? (u & 0x7fffffffffffffffULL) : ~u;
    double d;
    memcpy(&d, &b, 8);
    return d;
}
cudaError_t launchTlasRefit(TNode* tlasD, GNode* tlasOut, const double* instBounds, const int* order, const int* levelStart,
                            int nLevels, const double* centre, double pad, cudaStream_t stream);
// a device-built BLAS into its slice of the scene's node array, references made absolute (rt_api.cu: scene assembly)
cudaError_t launchRebaseNodes(GNode* dst, const GNode* src, int n, int nodeBase, int triBase, int f64, cudaStream_t stream);
cudaError_t launchSchedule(unsigned* tileCost, int* tileOrder, unsigned* work, int nTiles, cudaStream_t stream);
// first frame of a geometry: per-tile cost estimates from one probe ray per tile (rt_kernel_variant.cu: tileProbeKernel)
cudaError_t launchTileProbe(const DeviceScene& S, const CameraParams& cam, unsigned* tileCost, int nTiles, cudaStream_t stream);
size_t scheduleWorkBytes(int nTiles);   // scratch of launchSchedule
int renderBlockSize(int variant);
int renderStackCapacity();
int beamStackCapacity();                // entries of the beam walk's stack (rt_beam.cuh)
cudaError_t renderOccupancy(int variant, int* blocksPerSm, bool counters);

}  // namespace rtg
