// rt_kernels.cu — dispatcher over the two render kernels (rt_kernel_variant.cu, rt_render_kernel.cuh).
//
//   KV_FAST     lane = (pixel, sample); depth 0 (primary + shadow rays) as warp packets; valid when no
//               material draws random numbers inside `trace` (roughness == 0): the default
//   KV_GENERAL  lane = pixel, sequential PCG32 stream, per-lane traversal: any scene
#include <math_constants.h>

#include "rt_kernels.h"

namespace rtg {

cudaError_t launchVariantFast(const DeviceScene&, const CameraParams&, const RenderTargets&, int, bool, cudaStream_t);
cudaError_t launchVariantGeneral(const DeviceScene&, const CameraParams&, const RenderTargets&, int, bool, cudaStream_t);
cudaError_t occupancyVariantFast(int*, bool);
cudaError_t occupancyVariantGeneral(int*, bool);
int blockSizeVariantFast();
int blockSizeVariantGeneral();

cudaError_t launchRender(int variant, const DeviceScene& S, const CameraParams& cam, const RenderTargets& out, int blocks,
                         bool counters, cudaStream_t stream) {
    if (variant == KV_FAST) return launchVariantFast(S, cam, out, blocks, counters, stream);
    return launchVariantGeneral(S, cam, out, blocks, counters, stream);
}

cudaError_t renderOccupancy(int variant, int* blocksPerSm, bool counters) {
    if (variant == KV_FAST) return occupancyVariantFast(blocksPerSm, counters);
    return occupancyVariantGeneral(blocksPerSm, counters);
}

// ---- tile schedule -------------------------------------------------------------------------------------
// The render kernels hand out tiles through a counter; a frame therefore ends when the LAST tiles drawn are
// finished, and a packet of grazing rays can take 50x the mean (measured on the terrain scene: mean 55 us,
// p99.9 380 us, max 990 us) — a ~0.3 ms tail that does not shrink when the frame is split across GPUs.
// Longest-processing-time-first removes it: every frame records the clocks spent per tile, and this kernel
// turns them (two small launches: count, scatter) into the order of the NEXT frame of the same geometry (descending cost class, tiles of one
// class roughly in their spatial order), then clears the costs.  Classes are quarter-octaves of the cost.
// Any permutation renders the same image; the order only decides when each tile is drawn.
#define RT_SCHED_CLASSES 128
#define RT_SCHED_THREADS 1024

__device__ __forceinline__ int costClass(unsigned c) {
    const int k = 31 - __clz((int)(c | 1u));
    const int sub = k >= 2 ? (int)((c >> (k - 2)) & 3u) : 0;
    return RT_SCHED_CLASSES - 1 - (k * 4 + sub);      // class 0 = the most expensive tiles
}

// A stable counting sort in two launches.  work[b * CLASSES + c] = tiles of class c in block b's 1024-tile chunk.
// Stability matters: tiles of one class are drawn in their spatial order, which keeps the BVH nodes and triangles
// of neighbouring tiles in L2 while they are needed (an unordered scatter gave the whole gain back).
__global__ void __launch_bounds__(RT_SCHED_THREADS) scheduleCount(const unsigned* __restrict__ cost, unsigned* __restrict__ work, int nTiles) {
    __shared__ unsigned hist[RT_SCHED_CLASSES];
    const int tid = threadIdx.x, i = blockIdx.x * RT_SCHED_THREADS + tid;
    if (tid < RT_SCHED_CLASSES) hist[tid] = 0;
    __syncthreads();
    // neighbouring tiles mostly share a class: one atomic per (warp, class), not per tile
    const int c = i < nTiles ? costClass(cost[i]) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, c);
    if (c >= 0 && (tid & 31) == __ffs((int)peers) - 1) atomicAdd(&hist[c], (unsigned)__popc(peers));
    __syncthreads();
    if (tid < RT_SCHED_CLASSES) work[blockIdx.x * RT_SCHED_CLASSES + tid] = hist[tid];
}

__global__ void __launch_bounds__(RT_SCHED_THREADS) scheduleScatter(unsigned* __restrict__ cost, int* __restrict__ order,
                                                                    const unsigned* __restrict__ work, int nTiles) {
    __shared__ unsigned warpBase[RT_SCHED_THREADS / 32][RT_SCHED_CLASSES];   // per (warp, class): count, then first slot
    __shared__ unsigned classTotal[RT_SCHED_CLASSES], classBefore[RT_SCHED_CLASSES];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, i = blockIdx.x * RT_SCHED_THREADS + tid;
    for (int k = tid; k < (RT_SCHED_THREADS / 32) * RT_SCHED_CLASSES; k += RT_SCHED_THREADS) (&warpBase[0][0])[k] = 0;
    if (tid < RT_SCHED_CLASSES) {      // tiles of this class in all chunks / in the chunks before this block's
        unsigned total = 0, before = 0;
        for (int b = 0; b < (int)gridDim.x; b++) {
            const unsigned n = work[b * RT_SCHED_CLASSES + tid];
            total += n;
            if (b < (int)blockIdx.x) before += n;
        }
        classTotal[tid] = total; classBefore[tid] = before;
    }
    __syncthreads();
    const int c = i < nTiles ? costClass(cost[i]) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, c);
    if (c >= 0 && lane == __ffs((int)peers) - 1) warpBase[warp][c] = (unsigned)__popc(peers);
    if (tid < 32) {                    // exclusive prefix of the class sizes: first slot of each class
        unsigned run = 0;
        for (int c0 = 0; c0 < RT_SCHED_CLASSES; c0 += 32) {
            const unsigned n = classTotal[c0 + tid];
            unsigned incl = n;
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned up = __shfl_up_sync(0xffffffffu, incl, d);
                if (tid >= d) incl += up;
            }
            classTotal[c0 + tid] = run + incl - n;
            run += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    __syncthreads();
    if (tid < RT_SCHED_CLASSES) {      // first slot of (warp, class) = class start + earlier chunks + earlier warps
        unsigned run = classTotal[tid] + classBefore[tid];
        for (int w = 0; w < RT_SCHED_THREADS / 32; w++) {
            const unsigned n = warpBase[w][tid];
            warpBase[w][tid] = run;
            run += n;
        }
    }
    __syncthreads();
    if (c >= 0) {
        order[warpBase[warp][c] + __popc(peers & ((1u << lane) - 1u))] = i;
        cost[i] = 0;
    }
}

// ---- area lights: per-pixel draw counts -> first jitter index of every pixel ------------------------------------
// One block per 8-row chunk of this shard (the reference's task unit, Object+Extension.swift:75-82,288): exclusive
// prefix sum of the chunk's pixels in row-major order — the order in which the reference's chunk task advances its
// jitterIndex — kept mod 100 like every use of the index (:152-153).
#define RT_AREA_THREADS 256
__global__ void __launch_bounds__(RT_AREA_THREADS) areaPrefixKernel(unsigned* __restrict__ areaIndex, int width, int height,
                                                                    int shardIndex, int shardCount) {
    __shared__ unsigned part[RT_AREA_THREADS];
    const int chunk = (int)blockIdx.x * shardCount + shardIndex;
    const int row0 = chunk * 8;
    const int rows = height - row0 < 8 ? height - row0 : 8;
    const int n = rows * width;
    unsigned* a = areaIndex + (size_t)row0 * width;
    const int per = (n + RT_AREA_THREADS - 1) / RT_AREA_THREADS;
    const int tid = threadIdx.x;
    const int begin = tid * per, end = begin + per < n ? begin + per : n;
    unsigned sum = 0;
    for (int k = begin; k < end; k++) sum += a[k];
    part[tid] = sum % 100u;
    __syncthreads();
    if (tid == 0) {                    // 256 partial sums: a serial scan is a fraction of a microsecond
        unsigned run = 0;
        for (int t = 0; t < RT_AREA_THREADS; t++) { const unsigned v = part[t]; part[t] = run; run = (run + v) % 100u; }
    }
    __syncthreads();
    unsigned run = part[tid];
    for (int k = begin; k < end; k++) { const unsigned v = a[k]; a[k] = run; run = (run + v) % 100u; }
}

cudaError_t launchAreaPrefix(unsigned* areaIndex, int width, int height, int shardIndex, int shardCount, cudaStream_t stream) {
    const int nChunks = (height + 7) / 8;
    const int mine = (nChunks - shardIndex + shardCount - 1) / shardCount;
    if (mine > 0) RT_LAUNCH(areaPrefixKernel, mine, RT_AREA_THREADS, stream, areaIndex, width, height, shardIndex, shardCount);
    return cudaGetLastError();
}

// ---- animation: world bounds of the mesh instances ------------------------------------------------------------------
#define RT_BOUNDS_THREADS 256
__device__ __forceinline__ unsigned long long encOrdered(double d) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__global__ void __launch_bounds__(RT_BOUNDS_THREADS) instanceBoundsKernel(DeviceScene S, const BoundsChunk* __restrict__ chunks,
                                                                          const int32_t* __restrict__ derive,
                                                                          unsigned long long* __restrict__ boundsEnc) {
    __shared__ double red[6][RT_BOUNDS_THREADS / 32];
    const BoundsChunk ck = chunks[blockIdx.x];
    if (!derive[ck.inst]) return;
    const GInstance& I = S.instances[ck.inst];
    double M[12];
    for (int k = 0; k < 12; k++) M[k] = I.l2w[k];                 // column-major 3x4: M[c * 3 + r]
    const double mb[3] = {I.meshMotion[0], I.meshMotion[1], I.meshMotion[2]};
    double lo[3] = {CUDART_INF, CUDART_INF, CUDART_INF}, hi[3] = {-CUDART_INF, -CUDART_INF, -CUDART_INF};
    for (int k = (int)threadIdx.x; k < ck.count; k += RT_BOUNDS_THREADS) {
        double v[3][3];
        if (ck.f64) {
            const GTriD& t = S.trisD[ck.first + k];
            for (int a = 0; a < 3; a++) { v[0][a] = t.v0[a]; v[1][a] = t.v1[a]; v[2][a] = t.v2[a]; }
        } else {
            const GTri& t = S.tris[ck.first + k];
            for (int a = 0; a < 3; a++) { v[0][a] = (double)t.v0[a]; v[1][a] = (double)t.v1[a]; v[2][a] = (double)t.v2[a]; }
        }
        double pl[3], ph[3];                                       // the prim's local box, grown by the mesh motion (:283-289)
        for (int a = 0; a < 3; a++) {
            pl[a] = fmin(v[0][a], fmin(v[1][a], v[2][a]));
            ph[a] = fmax(v[0][a], fmax(v[1][a], v[2][a]));
            pl[a] = fmin(pl[a], pl[a] + mb[a]);
            ph[a] = fmax(ph[a], ph[a] + mb[a]);
        }
        // AABB.transformed(by:): centre and half extent through the matrix and its absolute value
        const double cO[3] = {(pl[0] + ph[0]) * 0.5, (pl[1] + ph[1]) * 0.5, (pl[2] + ph[2]) * 0.5};
        const double eO[3] = {(ph[0] - pl[0]) * 0.5, (ph[1] - pl[1]) * 0.5, (ph[2] - pl[2]) * 0.5};
        for (int r = 0; r < 3; r++) {
            const double cN = ((M[r] * cO[0] + M[3 + r] * cO[1]) + M[6 + r] * cO[2]) + M[9 + r];
            const double eN = (fabs(M[r]) * eO[0] + fabs(M[3 + r]) * eO[1]) + fabs(M[6 + r]) * eO[2];
            lo[r] = fmin(lo[r], cN - eN);
            hi[r] = fmax(hi[r], cN + eN);
        }
    }
    for (int a = 0; a < 3; a++)
        for (int off = 16; off > 0; off >>= 1) {
            lo[a] = fmin(lo[a], __shfl_down_sync(0xffffffffu, lo[a], off));
            hi[a] = fmax(hi[a], __shfl_down_sync(0xffffffffu, hi[a], off));
        }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) for (int a = 0; a < 3; a++) { red[a][warp] = lo[a]; red[3 + a][warp] = hi[a]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        double r = red[threadIdx.x][0];
        for (int w = 1; w < RT_BOUNDS_THREADS / 32; w++) r = threadIdx.x < 3 ? fmin(r, red[threadIdx.x][w]) : fmax(r, red[threadIdx.x][w]);
        unsigned long long* dst = boundsEnc + (size_t)ck.inst * 6 + threadIdx.x;
        if (threadIdx.x < 3) atomicMin(dst, encOrdered(r));
        else atomicMax(dst, encOrdered(r));
    }
}

cudaError_t launchInstanceBounds(const DeviceScene& S, const BoundsChunk* chunks, int nChunks, const int32_t* derive,
                                 unsigned long long* boundsEnc, cudaStream_t stream) {
    if (nChunks > 0) RT_LAUNCH(instanceBoundsKernel, nChunks, RT_BOUNDS_THREADS, stream, S, chunks, derive, boundsEnc);
    return cudaGetLastError();
}

// ---- animation: TLAS refit ---------------------------------------------------------------------------------------
// One block walks the levels of the tree from the deepest to the root (order / levelStart from rt_scene_create); a
// node takes each child's box from the instance bounds (leaf) or from the union of that child's two boxes (inner
// node, refitted one level earlier) and re-encodes its record for the traversal kernels: FP32 relative to the scene
// centre, padded by P and rounded outward (the encoding of rt_scene_create).
#define RT_REFIT_THREADS 256
struct RefitFrame { double c[3]; double pad; };
__global__ void __launch_bounds__(RT_REFIT_THREADS) tlasRefitKernel(TNode* __restrict__ tn, GNode* __restrict__ out,
                                                                    const double* __restrict__ instBounds, const int* __restrict__ order,
                                                                    const int* __restrict__ levelStart, int nLevels, RefitFrame F) {
    for (int lv = 0; lv < nLevels; lv++) {
        for (int k = levelStart[lv] + (int)threadIdx.x; k < levelStart[lv + 1]; k += RT_REFIT_THREADS) {
            const int ni = order[k];
            TNode t = tn[ni];
            const int32_t ch[2] = {t.child0, t.child1};
            double* lo[2] = {t.lo0, t.lo1};
            double* hi[2] = {t.hi0, t.hi1};
            for (int c = 0; c < 2; c++) {
                if (ch[c] == kEmptyRef) continue;
                if (ch[c] < 0) {
                    const double* b = instBounds + (size_t)(~ch[c]) * 6;
                    for (int a = 0; a < 3; a++) { lo[c][a] = b[a]; hi[c][a] = b[3 + a]; }
                } else {
                    const TNode& s = tn[ch[c]];
                    for (int a = 0; a < 3; a++) { lo[c][a] = fmin(s.lo0[a], s.lo1[a]); hi[c][a] = fmax(s.hi0[a], s.hi1[a]); }
                }
            }
            tn[ni] = t;
            GNode g;
            float l0[3], h0[3], l1[3], h1[3];
            for (int a = 0; a < 3; a++) {
                l0[a] = __double2float_rd(t.lo0[a] - F.c[a] - F.pad); h0[a] = __double2float_ru(t.hi0[a] - F.c[a] + F.pad);
                l1[a] = __double2float_rd(t.lo1[a] - F.c[a] - F.pad); h1[a] = __double2float_ru(t.hi1[a] - F.c[a] + F.pad);
            }
            g.q0[0] = l0[0]; g.q0[1] = l0[1]; g.q0[2] = l0[2]; g.q0[3] = h0[0];
            g.q1[0] = h0[1]; g.q1[1] = h0[2]; g.q1[2] = l1[0]; g.q1[3] = l1[1];
            g.q2[0] = l1[2]; g.q2[1] = h1[0]; g.q2[2] = h1[1]; g.q2[3] = h1[2];
            g.child0 = t.child0; g.child1 = t.child1; g.pad0 = g.pad1 = 0;
            out[ni] = g;
        }
        __syncthreads();
    }
}

cudaError_t launchTlasRefit(TNode* tlasD, GNode* tlasOut, const double* instBounds, const int* order, const int* levelStart,
                            int nLevels, const double* centre, double pad, cudaStream_t stream) {
    RefitFrame F;
    F.c[0] = centre[0]; F.c[1] = centre[1]; F.c[2] = centre[2]; F.pad = pad;
    if (nLevels > 0) RT_LAUNCH(tlasRefitKernel, 1, RT_REFIT_THREADS, stream, tlasD, tlasOut, instBounds, order, levelStart, nLevels, F);
    return cudaGetLastError();
}

size_t scheduleWorkBytes(int nTiles) {
    return (size_t)((nTiles + RT_SCHED_THREADS - 1) / RT_SCHED_THREADS) * RT_SCHED_CLASSES * sizeof(unsigned);
}

cudaError_t launchSchedule(unsigned* tileCost, int* tileOrder, unsigned* work, int nTiles, cudaStream_t stream) {
    const int blocks = (nTiles + RT_SCHED_THREADS - 1) / RT_SCHED_THREADS;
    RT_LAUNCH(scheduleCount, blocks, RT_SCHED_THREADS, stream, tileCost, work, nTiles);
    RT_LAUNCH(scheduleScatter, blocks, RT_SCHED_THREADS, stream, tileCost, tileOrder, work, nTiles);
    return cudaGetLastError();
}

// ---- scene assembly on the device: a BLAS the device builder left in device memory goes into its slice of the scene-wide
// node array with its references made absolute (gpu_layout.h), exactly what rt_scene_create's host loop does to a
// host-built BLAS
__global__ void __launch_bounds__(256) rebaseNodesKernel(GNode* __restrict__ dst, const GNode* __restrict__ src, int n, int nodeBase,
                                                         int triBase, int f64) {
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n) return;
    GNode g = src[i];
    int32_t* ref[2] = {&g.child0, &g.child1};
    for (int c = 0; c < 2; c++) {
        const int32_t r0 = *ref[c];
        if (r0 == kEmptyRef) continue;
        if (r0 >= 0) { *ref[c] = r0 + nodeBase; continue; }
        const uint32_t r = ~(uint32_t)r0;
        *ref[c] = ~(int32_t)((((r >> 3) + (uint32_t)triBase) << 4) | (f64 ? 8u : 0u) | (r & 7u));
    }
    dst[i] = g;
}

cudaError_t launchRebaseNodes(GNode* dst, const GNode* src, int n, int nodeBase, int triBase, int f64, cudaStream_t stream) {
    if (n > 0) RT_LAUNCH(rebaseNodesKernel, (n + 255) / 256, 256, stream, dst, src, n, nodeBase, triBase, f64);
    return cudaGetLastError();
}

int renderBlockSize(int variant) { return variant == KV_FAST ? blockSizeVariantFast() : blockSizeVariantGeneral(); }
int renderStackCapacity() { return RT_STACK; }
int beamStackCapacity() { return RT_BSTACK; }

}  // namespace rtg
