// bvh_build_gpu.cu — device-side builder of the GPU BLAS layout (gpu_layout.h): SURVEY.md §8(f) row 1.
//
// Same output contract as the host builder (bvh_build.h: BlasBuildOutput) from a different construction:
//   1. per-triangle FP64 bounds + the BLAS bounds (block reduction)              triPrepKernel / boundsKernel
//   2. 63-bit Morton code of each triangle's box centre, radix sort (cub)        mortonKernel + DeviceRadixSort
//   3. leaves = pairs of consecutive sorted triangles (count <= 2)
//   4. binary radix tree over the leaf keys, one thread per inner node           radixTreeKernel
//      (Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees", HPG 2012;
//       duplicate keys are told apart by their position)
//   5. boxes bottom-up: the second child to arrive at a node computes it         fitKernel
//   6. GNode records: both children's boxes relative to the BLAS centre, padded by M * 2^-19 and rounded
//      outward exactly like the host builder; triangles written in sorted (= leaf) order        emitKernels
// The reference builds its trees on the CPU (Accelearion/BVH.swift:78-250, single-threaded, every mesh twice,
// RTContext.swift:372-382,430-435); closest-hit results do not depend on the tree, only equal-t ties may differ.
// Arrays come back to the host because rt_scene_create concatenates the BLASes of a scene there.
#include <cub/device/device_radix_sort.cuh>

#include <cmath>
#include <cstring>

#include "bvh_build.h"

namespace rtg {
namespace {

#define GCK(call)                                          \
    do {                                                   \
        cudaError_t e_ = (call);                           \
        if (e_ != cudaSuccess) { cleanup(); return e_; }   \
    } while (0)

constexpr int kThreads = 256;

struct Bounds6 { double v[6]; };   // lo xyz, hi xyz

__device__ __forceinline__ double shflMin(double a, int off) { double b = __shfl_down_sync(0xffffffffu, a, off); return b < a ? b : a; }
__device__ __forceinline__ double shflMax(double a, int off) { double b = __shfl_down_sync(0xffffffffu, a, off); return b > a ? b : a; }

// bounds of every triangle of the BLAS (FP64) and per-block partial bounds of all of them; flag[0] |= 1 when
// some coordinate is not exactly representable in FP32 (-> FP64 triangle records)
__global__ void __launch_bounds__(kThreads) triPrepKernel(const double* __restrict__ verts, const long long* __restrict__ idx,
                                                          int n, Bounds6* __restrict__ triBox, Bounds6* __restrict__ partial,
                                                          int* __restrict__ flag) {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool wide = false;
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const double* v = verts + idx[i] * 9;
        Bounds6 b;
        for (int a = 0; a < 3; a++) {
            const double p0 = v[a], p1 = v[3 + a], p2 = v[6 + a];
            b.v[a] = fmin(p0, fmin(p1, p2));
            b.v[3 + a] = fmax(p0, fmax(p1, p2));
            wide |= (double)(float)p0 != p0 || (double)(float)p1 != p1 || (double)(float)p2 != p2;
            lo[a] = fmin(lo[a], b.v[a]);
            hi[a] = fmax(hi[a], b.v[3 + a]);
        }
        triBox[i] = b;
    }
    __shared__ double sm[kThreads / 32][6];
    for (int a = 0; a < 3; a++)
        for (int off = 16; off > 0; off >>= 1) { lo[a] = shflMin(lo[a], off); hi[a] = shflMax(hi[a], off); }
    if ((threadIdx.x & 31) == 0)
        for (int a = 0; a < 3; a++) { sm[threadIdx.x >> 5][a] = lo[a]; sm[threadIdx.x >> 5][3 + a] = hi[a]; }
    if (__syncthreads_or(wide) && threadIdx.x == 0) atomicOr(flag, 1);
    if (threadIdx.x == 0) {
        Bounds6 r;
        for (int a = 0; a < 6; a++) r.v[a] = sm[0][a];
        for (int w = 1; w < kThreads / 32; w++)
            for (int a = 0; a < 3; a++) { r.v[a] = fmin(r.v[a], sm[w][a]); r.v[3 + a] = fmax(r.v[3 + a], sm[w][3 + a]); }
        partial[blockIdx.x] = r;
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) {   // 21 bits -> every third bit
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(kThreads) mortonKernel(const Bounds6* __restrict__ triBox, int n, Bounds6 root,
                                                         unsigned long long* __restrict__ keys, unsigned* __restrict__ vals) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    unsigned long long code = 0;
    for (int a = 0; a < 3; a++) {
        const double ext = root.v[3 + a] - root.v[a];
        const double c = 0.5 * (triBox[i].v[a] + triBox[i].v[3 + a]);
        double q = ext > 0 ? (c - root.v[a]) / ext * 2097152.0 : 0.0;
        q = fmin(fmax(q, 0.0), 2097151.0);
        code |= spread21((unsigned long long)q) << (2 - a);
    }
    keys[i] = code;
    vals[i] = (unsigned)i;
}

// common-prefix length of leaves i and j (keys of their first triangle; position breaks ties); -1 outside
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int L, int i, int j) {
    if (j < 0 || j >= L) return -1;
    const unsigned long long a = keys[2 * i], b = keys[2 * j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz(i ^ j);
}

// children: >= 0 inner node, < 0 leaf ~l
__global__ void __launch_bounds__(kThreads) radixTreeKernel(const unsigned long long* __restrict__ keys, int L,
                                                            int2* __restrict__ child, int* __restrict__ parentInner,
                                                            int* __restrict__ parentLeaf) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= L - 1) return;
    const int d = delta(keys, L, i, i + 1) - delta(keys, L, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, L, i, i - d);
    int lmax = 2;
    while (delta(keys, L, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, L, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, L, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, L, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int g = i + s * d + (d < 0 ? d : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    int2 c;
    if (lo == g) { c.x = ~g; parentLeaf[g] = i; } else { c.x = g; parentInner[g] = i; }
    if (hi == g + 1) { c.y = ~(g + 1); parentLeaf[g + 1] = i; } else { c.y = g + 1; parentInner[g + 1] = i; }
    child[i] = c;
    if (i == 0) parentInner[0] = -1;
}

__device__ __forceinline__ void merge(Bounds6& a, const Bounds6& b) {
    for (int k = 0; k < 3; k++) { a.v[k] = fmin(a.v[k], b.v[k]); a.v[3 + k] = fmax(a.v[3 + k], b.v[3 + k]); }
}

// leaf boxes, then inner boxes bottom-up
__global__ void __launch_bounds__(kThreads) fitKernel(const Bounds6* __restrict__ triBox, const unsigned* __restrict__ order, int n,
                                                      int L, const int2* __restrict__ child, const int* __restrict__ parentInner,
                                                      const int* __restrict__ parentLeaf, Bounds6* __restrict__ leafBox,
                                                      Bounds6* __restrict__ innerBox, int* __restrict__ arrived) {
    const int l = blockIdx.x * kThreads + threadIdx.x;
    if (l >= L) return;
    Bounds6 b = triBox[order[2 * l]];
    if (2 * l + 1 < n) merge(b, triBox[order[2 * l + 1]]);
    leafBox[l] = b;
    __threadfence();
    int node = parentLeaf[l];
    while (node >= 0) {
        if (atomicAdd(&arrived[node], 1) == 0) return;        // the sibling subtree is not finished: it will go on
        __threadfence();
        const int2 c = child[node];
        Bounds6 u = c.x < 0 ? leafBox[~c.x] : innerBox[c.x];
        merge(u, c.y < 0 ? leafBox[~c.y] : innerBox[c.y]);
        innerBox[node] = u;
        __threadfence();
        node = parentInner[node];
    }
}

// longest leaf-to-root path, in nodes (the traversal stacks must hold it)
__global__ void __launch_bounds__(kThreads) depthKernel(int L, const int* __restrict__ parentInner, const int* __restrict__ parentLeaf,
                                                        int* __restrict__ maxDepth) {
    const int l = blockIdx.x * kThreads + threadIdx.x;
    if (l >= L) return;
    int depth = 1;
    for (int node = parentLeaf[l]; node >= 0; node = parentInner[node]) depth++;
    for (int off = 16; off > 0; off >>= 1) depth = max(depth, __shfl_down_sync(0xffffffffu, depth, off));
    if ((threadIdx.x & 31) == 0) atomicMax(maxDepth, depth);
}

__global__ void __launch_bounds__(kThreads) emitNodesKernel(int L, int n, const int2* __restrict__ child, const Bounds6* __restrict__ leafBox,
                                                            const Bounds6* __restrict__ innerBox, double cx, double cy, double cz,
                                                            double pad, GNode* __restrict__ nodes) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= L - 1) return;
    const int2 c = child[i];
    const double centre[3] = {cx, cy, cz};
    float lo[2][3], hi[2][3];
    int32_t ref[2];
    for (int k = 0; k < 2; k++) {
        const int ck = k ? c.y : c.x;
        const Bounds6 b = ck < 0 ? leafBox[~ck] : innerBox[ck];
        for (int a = 0; a < 3; a++) {
            lo[k][a] = __double2float_rd(b.v[a] - centre[a] - pad);       // rounded outward, as bvh_build.cpp putBox
            hi[k][a] = __double2float_ru(b.v[3 + a] - centre[a] + pad);
        }
        if (ck < 0) {
            const int l = ~ck, count = (2 * l + 1 < n) ? 2 : 1;
            ref[k] = ~(int32_t)(((unsigned)(2 * l) << 3) | (unsigned)(count - 1));
        } else {
            ref[k] = ck;
        }
    }
    GNode g;
    g.q0[0] = lo[0][0]; g.q0[1] = lo[0][1]; g.q0[2] = lo[0][2]; g.q0[3] = hi[0][0];
    g.q1[0] = hi[0][1]; g.q1[1] = hi[0][2]; g.q1[2] = lo[1][0]; g.q1[3] = lo[1][1];
    g.q2[0] = lo[1][2]; g.q2[1] = hi[1][0]; g.q2[2] = hi[1][1]; g.q2[3] = hi[1][2];
    g.child0 = ref[0]; g.child1 = ref[1]; g.pad0 = 0; g.pad1 = 0;
    nodes[i] = g;
}

__global__ void __launch_bounds__(kThreads) emitTrisKernel(const double* __restrict__ verts, const long long* __restrict__ idx,
                                                           const unsigned* __restrict__ order, int n, GTri* __restrict__ tris,
                                                           GTriD* __restrict__ trisD) {
    const int p = blockIdx.x * kThreads + threadIdx.x;
    if (p >= n) return;
    const long long tri = idx[order[p]];
    const double* v = verts + tri * 9;
    if (trisD) {
        GTriD t;
        for (int k = 0; k < 3; k++) { t.v0[k] = v[k]; t.v1[k] = v[3 + k]; t.v2[k] = v[6 + k]; }
        t.triIndex = tri;
        trisD[p] = t;
    } else {
        GTri t;
        for (int k = 0; k < 3; k++) { t.v0[k] = (float)v[k]; t.v1[k] = (float)v[3 + k]; t.v2[k] = (float)v[6 + k]; }
        t.triIndex = (int32_t)tri; t.pad1 = 0; t.pad2 = 0;
        tris[p] = t;
    }
}

}  // namespace

// Returns cudaSuccess and fills `out` like buildBlas; out.maxDepth tells the caller whether the tree fits its stacks.
// `dVerts`: the scene's triangle vertices (9 doubles each) already on the device.
cudaError_t buildBlasDevice(const double* dVerts, const BlasBuildInput& in, BlasBuildOutput& out) {
    out = BlasBuildOutput();
    const int n = (int)in.count;
    long long* dIdx = nullptr;
    Bounds6 *dTriBox = nullptr, *dPartial = nullptr, *dLeafBox = nullptr, *dInnerBox = nullptr;
    unsigned long long *dKeys = nullptr, *dKeys2 = nullptr;
    unsigned *dVals = nullptr, *dVals2 = nullptr;
    int2* dChild = nullptr;
    int *dParentInner = nullptr, *dParentLeaf = nullptr, *dArrived = nullptr, *dFlags = nullptr;
    void* dTemp = nullptr;
    GNode* dNodes = nullptr;
    GTri* dTris = nullptr;
    GTriD* dTrisD = nullptr;
    auto cleanup = [&]() {
        cudaFree(dIdx); cudaFree(dTriBox); cudaFree(dPartial); cudaFree(dLeafBox); cudaFree(dInnerBox); cudaFree(dKeys);
        cudaFree(dKeys2); cudaFree(dVals); cudaFree(dVals2); cudaFree(dChild); cudaFree(dParentInner); cudaFree(dParentLeaf);
        cudaFree(dArrived); cudaFree(dFlags); cudaFree(dTemp); cudaFree(dNodes); cudaFree(dTris); cudaFree(dTrisD);
    };
    if (n <= 0) return cudaSuccess;
    const int L = (n + 1) / 2;
    const int blocksN = (n + kThreads - 1) / kThreads, blocksL = (L + kThreads - 1) / kThreads;
    const int redBlocks = blocksN < 1024 ? blocksN : 1024;

    GCK(cudaMalloc(&dIdx, (size_t)n * 8));
    GCK(cudaMemcpy(dIdx, in.triIndices, (size_t)n * 8, cudaMemcpyHostToDevice));
    GCK(cudaMalloc(&dTriBox, (size_t)n * sizeof(Bounds6)));
    GCK(cudaMalloc(&dPartial, (size_t)redBlocks * sizeof(Bounds6)));
    GCK(cudaMalloc(&dFlags, 2 * sizeof(int)));
    GCK(cudaMemset(dFlags, 0, 2 * sizeof(int)));
    triPrepKernel<<<redBlocks, kThreads>>>(dVerts, dIdx, n, dTriBox, dPartial, dFlags);
    GCK(cudaGetLastError());
    std::vector<Bounds6> partial((size_t)redBlocks);
    int flags[2] = {0, 0};
    GCK(cudaMemcpy(partial.data(), dPartial, (size_t)redBlocks * sizeof(Bounds6), cudaMemcpyDeviceToHost));
    GCK(cudaMemcpy(flags, dFlags, sizeof flags, cudaMemcpyDeviceToHost));
    Bounds6 root = partial[0];
    for (int b = 1; b < redBlocks; b++)
        for (int a = 0; a < 3; a++) { root.v[a] = std::fmin(root.v[a], partial[b].v[a]); root.v[3 + a] = std::fmax(root.v[3 + a], partial[b].v[3 + a]); }
    out.f64 = flags[0] != 0;
    double M = 0;
    for (int a = 0; a < 3; a++) {
        out.rootLo[a] = root.v[a]; out.rootHi[a] = root.v[3 + a]; out.centre[a] = 0.5 * (root.v[a] + root.v[3 + a]);
        M = std::fmax(M, std::fmax(std::fabs(root.v[a] - out.centre[a]), std::fabs(root.v[3 + a] - out.centre[a])));
    }
    out.pad = std::fmax(M * std::ldexp(1.0, -19), 1e-30);
    out.maxLeaf = n >= 2 ? 2 : 1;

    // Morton order
    GCK(cudaMalloc(&dKeys, (size_t)n * 8)); GCK(cudaMalloc(&dKeys2, (size_t)n * 8));
    GCK(cudaMalloc(&dVals, (size_t)n * 4)); GCK(cudaMalloc(&dVals2, (size_t)n * 4));
    mortonKernel<<<blocksN, kThreads>>>(dTriBox, n, root, dKeys, dVals);
    GCK(cudaGetLastError());
    size_t tempBytes = 0;
    GCK(cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, dKeys, dKeys2, dVals, dVals2, n, 0, 63));
    GCK(cudaMalloc(&dTemp, tempBytes));
    GCK(cub::DeviceRadixSort::SortPairs(dTemp, tempBytes, dKeys, dKeys2, dVals, dVals2, n, 0, 63));

    // triangles in leaf order
    if (out.f64) GCK(cudaMalloc(&dTrisD, (size_t)n * sizeof(GTriD))); else GCK(cudaMalloc(&dTris, (size_t)n * sizeof(GTri)));
    emitTrisKernel<<<blocksN, kThreads>>>(dVerts, dIdx, dVals2, n, dTris, dTrisD);
    GCK(cudaGetLastError());
    if (out.f64) { out.trisD.resize((size_t)n); GCK(cudaMemcpy(out.trisD.data(), dTrisD, (size_t)n * sizeof(GTriD), cudaMemcpyDeviceToHost)); }
    else { out.tris.resize((size_t)n); GCK(cudaMemcpy(out.tris.data(), dTris, (size_t)n * sizeof(GTri), cudaMemcpyDeviceToHost)); }

    if (L == 1) {   // one leaf holds everything
        out.rootRef = ~(int32_t)((0u << 3) | (unsigned)(n - 1));
        out.maxDepth = 1;
        cleanup();
        return cudaSuccess;
    }

    // hierarchy + boxes + records
    GCK(cudaMalloc(&dChild, (size_t)(L - 1) * sizeof(int2)));
    GCK(cudaMalloc(&dParentInner, (size_t)(L - 1) * 4)); GCK(cudaMalloc(&dParentLeaf, (size_t)L * 4));
    GCK(cudaMalloc(&dArrived, (size_t)(L - 1) * 4));
    GCK(cudaMemset(dArrived, 0, (size_t)(L - 1) * 4));
    GCK(cudaMalloc(&dLeafBox, (size_t)L * sizeof(Bounds6))); GCK(cudaMalloc(&dInnerBox, (size_t)(L - 1) * sizeof(Bounds6)));
    radixTreeKernel<<<(L - 1 + kThreads - 1) / kThreads, kThreads>>>(dKeys2, L, dChild, dParentInner, dParentLeaf);
    GCK(cudaGetLastError());
    fitKernel<<<blocksL, kThreads>>>(dTriBox, dVals2, n, L, dChild, dParentInner, dParentLeaf, dLeafBox, dInnerBox, dArrived);
    GCK(cudaGetLastError());
    depthKernel<<<blocksL, kThreads>>>(L, dParentInner, dParentLeaf, dFlags + 1);
    GCK(cudaGetLastError());
    GCK(cudaMalloc(&dNodes, (size_t)(L - 1) * sizeof(GNode)));
    emitNodesKernel<<<(L - 1 + kThreads - 1) / kThreads, kThreads>>>(L, n, dChild, dLeafBox, dInnerBox, out.centre[0], out.centre[1],
                                                                    out.centre[2], out.pad, dNodes);
    GCK(cudaGetLastError());
    out.nodes.resize((size_t)(L - 1));
    GCK(cudaMemcpy(out.nodes.data(), dNodes, (size_t)(L - 1) * sizeof(GNode), cudaMemcpyDeviceToHost));
    GCK(cudaMemcpy(flags, dFlags, sizeof flags, cudaMemcpyDeviceToHost));
    out.maxDepth = flags[1];
    out.rootRef = 0;
    cleanup();
    return cudaSuccess;
}

}  // namespace rtg
