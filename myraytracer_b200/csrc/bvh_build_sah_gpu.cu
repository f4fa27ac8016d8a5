// bvh_build_sah_gpu.cu — binned-SAH construction of the GPU BLAS layout ON THE DEVICE.
// COMPILE WITH -fmad=false (the split costs are compared in FP64 exactly as bvh_build.cpp compares them).
//
// The reference builds every mesh BVH on one CPU thread, top-down, 12-bin SAH (Accelearion/BVH.swift:78-250, called
// from RTContext.swift:372-382,430-435).  bvh_build.cpp is this repository's host builder (16 bins, leaf-cost
// termination, <= kMaxLeaf triangles per leaf); bvh_build_gpu.cu builds a Morton-order radix tree, which is quick to
// make but walks 66 % slower.  This file makes the HOST builder's decisions, one tree level per round of kernels:
//   prep        per-triangle FP64 box (+ motion growth), BLAS bounds, "needs FP64 records" flag
//   per level   1. centroid bounds of every open node            sahCentroids   (ordered-u64 atomicMin/Max)
//               2. 3 axes x 16 bins: box and count per bin        sahBin
//               3. one thread per open node walks the 3 x 15 candidate planes in the host builder's order and
//                  with its arithmetic, then applies its leaf rule                                  sahSplit
//               4. exclusive scan of the "splits" flags -> slots of the children in the node pool   (scan kernels)
//               5. children records (range, box = union of their bins)                             sahChildren
//               6. triangles move to their child's range, warp-aggregated cursors                  sahPartition
//   final       inner nodes numbered as the host builder numbers them (subtree counts up the levels, visiting ranks down
//               the levels, one scan), leaves sorted by triangle index, GNode and GTri/GTriD records written exactly
//               like the host builder's (same centre, padding and outward rounding)
// min, max and integer counts do not depend on the order the atomics land in, so every level decides what
// bvh_build.cpp's recursion decides: same splits, same leaves, same boxes, same node numbers: the GNode array is the host builder's
// byte for byte.  Differences: triangles inside a leaf are in index order, and a node whose centroids all coincide is
// halved with its own box handed to both halves (conservative).  Closest-hit results do not depend on
// any of that; only equal-t ties inside one leaf may resolve differently.
// A block whose 256 positions all belong to one node accumulates in shared memory and flushes once (the top levels);
// mixed blocks go to global memory directly (deep levels, where the addresses are spread out anyway).
// The arrays come back to the host, where rt_scene_create concatenates a scene's BLASes, as with the other builders.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

#include "bvh_build.h"
#include "rt_kernels.h"

namespace rtg {
cudaError_t launchExclusiveScan(unsigned* a, long long n, unsigned* blockSums, cudaStream_t stream);   // ply_ingest.cu

namespace {

typedef unsigned long long u64;
constexpr int kT = 256;
constexpr int kBins = 16;          // bvh_build.cpp
constexpr int kSlots = 3 * kBins;  // (axis, bin)
constexpr double kCostTri = 1.2, kCostNode = 1.0;
constexpr int kLeaf = -2, kHalve = -1;   // sAxis values besides 0..2

// order-preserving map double -> u64, so that atomicMin/atomicMax on the image are min/max on the doubles
__device__ __forceinline__ u64 encD(double d) {
    const u64 b = (u64)__double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double decD(u64 u) {
    const u64 b = (u >> 63) ? (u & 0x7fffffffffffffffULL) : ~u;
    return __longlong_as_double((long long)b);
}
inline double decHost(u64 u) {
    const u64 b = (u >> 63) ? (u & 0x7fffffffffffffffULL) : ~u;
    double d;
    std::memcpy(&d, &b, 8);
    return d;
}

struct SahState {
    int n;
    const double* triBox;        // n * 6: lo xyz, hi xyz
    int* order[2];               // position -> triangle (local index), ping-pong
    int* nodeOf[2];              // position -> open node, -1 once its leaf is final
    // node pool (2n): a level's nodes are contiguous, the children of level d follow it in parent order
    int* nFirst;
    int* nCount;
    int* nLeft;                  // left child (right = left + 1), -1 = leaf
    double* nBox;                // pool * 6
    // per level, indexed by j = node - levelBase
    u64* cbLo; u64* cbHi;        // A * 3: centroid bounds
    u64* binLo; u64* binHi;      // A * kSlots * 3
    unsigned* binCnt;            // A * kSlots
    unsigned* flag;              // A + 1: 1 = splits; scanned in place into the pair's slot, flag[A] = number of splits
    int* sAxis; int* sBin; int* sLc;
    double* childBox;            // A * 12
    unsigned* cursor;            // A * 2
};

__device__ __forceinline__ double areaOf(const double* lo, const double* hi) {   // Box::area, bvh_build.cpp
    const double ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
    if (ex < 0 || ey < 0 || ez < 0) return 0;
    return 2.0 * (ex * ey + ey * ez + ez * ex);
}
__device__ __forceinline__ int binOf(double c, double cmin, double scale) {
    const int k = (int)((c - cmin) * scale);
    return k > kBins - 1 ? kBins - 1 : (k < 0 ? 0 : k);
}

// true when every thread of the block passed the same node >= 0 (two barriers; all threads must call)
__device__ __forceinline__ bool blockUniform(int node) {
    __shared__ int sNode, sMixed;
    if (threadIdx.x == 0) { sNode = node; sMixed = 0; }
    __syncthreads();
    if (node != sNode) sMixed = 1;
    __syncthreads();
    const bool u = sMixed == 0 && sNode >= 0;
    __syncthreads();                 // sNode / sMixed may be rewritten by the next call
    return u;
}

__global__ void __launch_bounds__(kT) sahPrep(const double* __restrict__ verts, const long long* __restrict__ idx, int n, double mx,
                                               double my, double mz, int blur, double* __restrict__ triBox, u64* __restrict__ rootLo,
                                               u64* __restrict__ rootHi, int* __restrict__ wideFlag, int* __restrict__ order,
                                               int* __restrict__ nodeOf) {
    __shared__ u64 sLo[3], sHi[3];
    __shared__ int sWide;
    if (threadIdx.x < 3) { sLo[threadIdx.x] = ~0ULL; sHi[threadIdx.x] = 0ULL; }
    if (threadIdx.x == 0) sWide = 0;
    __syncthreads();
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i < n) {
        const double* v = verts + idx[i] * 9;
        const double m[3] = {mx, my, mz};
        bool wide = false;
        for (int a = 0; a < 3; a++) {
            const double p0 = v[a], p1 = v[3 + a], p2 = v[6 + a];
            double lo = fmin(p0, fmin(p1, p2)), hi = fmax(p0, fmax(p1, p2));
            if (blur) {                                            // the box also holds the moved triangle (RTContext.swift:283-289)
                lo = fmin(lo, fmin(p0 + m[a], fmin(p1 + m[a], p2 + m[a])));
                hi = fmax(hi, fmax(p0 + m[a], fmax(p1 + m[a], p2 + m[a])));
            }
            triBox[(long long)i * 6 + a] = lo;
            triBox[(long long)i * 6 + 3 + a] = hi;
            atomicMin(&sLo[a], encD(lo));
            atomicMax(&sHi[a], encD(hi));
            wide = wide || (double)(float)p0 != p0 || (double)(float)p1 != p1 || (double)(float)p2 != p2;
        }
        if (wide) sWide = 1;
        order[i] = i;
        nodeOf[i] = 0;
    }
    __syncthreads();
    if (threadIdx.x < 3) { atomicMin(rootLo + threadIdx.x, sLo[threadIdx.x]); atomicMax(rootHi + threadIdx.x, sHi[threadIdx.x]); }
    if (threadIdx.x == 0 && sWide) *wideFlag = 1;
}

__global__ void __launch_bounds__(kT) sahCentroids(SahState S, int cur, int levelBase) {
    __shared__ u64 sLo[3], sHi[3];
    const int p = blockIdx.x * kT + threadIdx.x;
    const int node = p < S.n ? S.nodeOf[cur][p] : -1;
    const bool uniform = blockUniform(node);
    if (uniform) {
        if (threadIdx.x < 3) { sLo[threadIdx.x] = ~0ULL; sHi[threadIdx.x] = 0ULL; }
        __syncthreads();
    }
    if (node >= 0 && S.nCount[node] > 1) {
        const double* b = S.triBox + (long long)S.order[cur][p] * 6;
        u64* lo = uniform ? sLo : S.cbLo + (long long)(node - levelBase) * 3;
        u64* hi = uniform ? sHi : S.cbHi + (long long)(node - levelBase) * 3;
        for (int a = 0; a < 3; a++) {
            const u64 c = encD(0.5 * (b[a] + b[3 + a]));
            atomicMin(lo + a, c);
            atomicMax(hi + a, c);
        }
    }
    if (uniform) {
        __syncthreads();
        if (threadIdx.x < 3) {
            atomicMin(S.cbLo + (long long)(node - levelBase) * 3 + threadIdx.x, sLo[threadIdx.x]);
            atomicMax(S.cbHi + (long long)(node - levelBase) * 3 + threadIdx.x, sHi[threadIdx.x]);
        }
    }
}

__global__ void __launch_bounds__(kT) sahBin(SahState S, int cur, int levelBase) {
    __shared__ u64 sLo[kSlots * 3], sHi[kSlots * 3];
    __shared__ unsigned sCnt[kSlots];
    const int p = blockIdx.x * kT + threadIdx.x;
    const int node = p < S.n ? S.nodeOf[cur][p] : -1;
    const bool uniform = blockUniform(node);
    if (uniform) {
        if (threadIdx.x < kSlots * 3) { sLo[threadIdx.x] = ~0ULL; sHi[threadIdx.x] = 0ULL; }
        if (threadIdx.x < kSlots) sCnt[threadIdx.x] = 0u;
        __syncthreads();
    }
    const long long j = node - levelBase;
    if (node >= 0 && S.nCount[node] > 1) {
        const double* b = S.triBox + (long long)S.order[cur][p] * 6;
        u64 eLo[3], eHi[3];
        for (int a = 0; a < 3; a++) { eLo[a] = encD(b[a]); eHi[a] = encD(b[3 + a]); }
        for (int a = 0; a < 3; a++) {
            const double cmin = decD(S.cbLo[j * 3 + a]), cmax = decD(S.cbHi[j * 3 + a]);
            if (!(cmax > cmin)) continue;
            const int slot = a * kBins + binOf(0.5 * (b[a] + b[3 + a]), cmin, kBins / (cmax - cmin));
            u64* lo = uniform ? sLo + slot * 3 : S.binLo + (j * kSlots + slot) * 3;
            u64* hi = uniform ? sHi + slot * 3 : S.binHi + (j * kSlots + slot) * 3;
            for (int c = 0; c < 3; c++) { atomicMin(lo + c, eLo[c]); atomicMax(hi + c, eHi[c]); }
            atomicAdd(uniform ? sCnt + slot : S.binCnt + j * kSlots + slot, 1u);
        }
    }
    if (uniform) {
        __syncthreads();
        if (threadIdx.x < kSlots * 3 && sCnt[threadIdx.x / 3] > 0u) {
            atomicMin(S.binLo + j * kSlots * 3 + threadIdx.x, sLo[threadIdx.x]);
            atomicMax(S.binHi + j * kSlots * 3 + threadIdx.x, sHi[threadIdx.x]);
        }
        if (threadIdx.x < kSlots && sCnt[threadIdx.x] > 0u) atomicAdd(S.binCnt + j * kSlots + threadIdx.x, sCnt[threadIdx.x]);
    }
}

__device__ __forceinline__ void growBox(double* lo, double* hi, const double* blo, const double* bhi) {
    for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], blo[a]); hi[a] = fmax(hi[a], bhi[a]); }
}

// Builder::build of bvh_build.cpp for one node, from the bins: candidate planes in axis, then bin order, the first
// strictly cheaper one wins; leaf when it is no more expensive than the best split.
__global__ void __launch_bounds__(kT) sahSplit(SahState S, int levelBase, int A) {
    const int j = blockIdx.x * kT + threadIdx.x;
    if (j >= A) return;
    const int node = levelBase + j;
    const int count = S.nCount[node];
    S.flag[j] = 0u;
    S.sAxis[j] = kLeaf;
    if (count <= 1) return;
    const double inf = INFINITY;
    double bestCost = inf;
    int bestAxis = -1, bestBin = -1;
    double blo[kBins][3], bhi[kBins][3];
    unsigned bc[kBins];
    for (int a = 0; a < 3; a++) {
        const double cmin = decD(S.cbLo[(long long)j * 3 + a]), cmax = decD(S.cbHi[(long long)j * 3 + a]);
        if (!(cmax > cmin)) continue;
        for (int k = 0; k < kBins; k++) {
            const long long s = (long long)j * kSlots + a * kBins + k;
            bc[k] = S.binCnt[s];
            for (int c = 0; c < 3; c++) {
                blo[k][c] = bc[k] ? decD(S.binLo[s * 3 + c]) : inf;
                bhi[k][c] = bc[k] ? decD(S.binHi[s * 3 + c]) : -inf;
            }
        }
        double rightArea[kBins];
        unsigned rightCnt[kBins];
        double lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
        unsigned rc = 0;
        for (int k = kBins - 1; k >= 1; k--) { growBox(lo, hi, blo[k], bhi[k]); rc += bc[k]; rightArea[k] = areaOf(lo, hi); rightCnt[k] = rc; }
        double llo[3] = {inf, inf, inf}, lhi[3] = {-inf, -inf, -inf};
        unsigned lc = 0;
        for (int k = 0; k < kBins - 1; k++) {
            growBox(llo, lhi, blo[k], bhi[k]);
            lc += bc[k];
            if (lc == 0 || rightCnt[k + 1] == 0) continue;
            const double cost = areaOf(llo, lhi) * (double)lc + rightArea[k + 1] * (double)rightCnt[k + 1];
            if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = k; }
        }
    }
    const double* nb = S.nBox + (long long)node * 6;
    double* cb = S.childBox + (long long)j * 12;
    if (bestAxis >= 0) {
        const double parentArea = areaOf(nb, nb + 3);
        const double leafCost = kCostTri * (double)count * parentArea;
        const double splitCost = kCostNode * parentArea + kCostTri * bestCost;
        if (count <= kMaxLeaf && leafCost <= splitCost) return;
        double lo[2][3] = {{inf, inf, inf}, {inf, inf, inf}}, hi[2][3] = {{-inf, -inf, -inf}, {-inf, -inf, -inf}};
        unsigned lc = 0;
        for (int k = 0; k < kBins; k++) {
            const long long s = (long long)j * kSlots + bestAxis * kBins + k;
            const unsigned c = S.binCnt[s];
            if (!c) continue;
            const int side = k > bestBin;
            if (!side) lc += c;
            for (int a = 0; a < 3; a++) {
                lo[side][a] = fmin(lo[side][a], decD(S.binLo[s * 3 + a]));
                hi[side][a] = fmax(hi[side][a], decD(S.binHi[s * 3 + a]));
            }
        }
        for (int side = 0; side < 2; side++)
            for (int a = 0; a < 3; a++) { cb[side * 6 + a] = lo[side][a]; cb[side * 6 + 3 + a] = hi[side][a]; }
        S.sAxis[j] = bestAxis; S.sBin[j] = bestBin; S.sLc[j] = (int)lc;
    } else {
        if (count <= kMaxLeaf) return;
        for (int side = 0; side < 2; side++)          // coincident centroids: halve the list, both halves keep this box
            for (int a = 0; a < 6; a++) cb[side * 6 + a] = nb[a];
        S.sAxis[j] = kHalve; S.sBin[j] = 0; S.sLc[j] = count / 2;
    }
    S.flag[j] = 1u;
}

__global__ void __launch_bounds__(kT) sahChildren(SahState S, int levelBase, int A) {
    const int j = blockIdx.x * kT + threadIdx.x;
    if (j >= A || S.sAxis[j] == kLeaf) return;
    const int node = levelBase + j;
    const int L = levelBase + A + 2 * (int)S.flag[j];
    const int first = S.nFirst[node], count = S.nCount[node], lc = S.sLc[j];
    S.nLeft[node] = L;
    S.nFirst[L] = first; S.nCount[L] = lc; S.nLeft[L] = -1;
    S.nFirst[L + 1] = first + lc; S.nCount[L + 1] = count - lc; S.nLeft[L + 1] = -1;
    for (int a = 0; a < 12; a++) S.nBox[(long long)L * 6 + a] = S.childBox[(long long)j * 12 + a];
}

__global__ void __launch_bounds__(kT) sahPartition(SahState S, int cur, int levelBase) {
    const int p = blockIdx.x * kT + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int node = p < S.n ? S.nodeOf[cur][p] : -1;
    const int tri = p < S.n ? S.order[cur][p] : 0;
    int key = -1, dest = p, child = -1, side = 0, j = 0;
    if (node >= 0) {
        j = node - levelBase;
        const int axis = S.sAxis[j];
        if (axis == kHalve) {
            child = S.nLeft[node] + (p - S.nFirst[node] >= S.sLc[j] ? 1 : 0);
        } else if (axis >= 0) {
            const double* b = S.triBox + (long long)tri * 6;
            const double cmin = decD(S.cbLo[(long long)j * 3 + axis]), cmax = decD(S.cbHi[(long long)j * 3 + axis]);
            side = binOf(0.5 * (b[axis] + b[3 + axis]), cmin, kBins / (cmax - cmin)) > S.sBin[j] ? 1 : 0;
            key = j * 2 + side;
            child = S.nLeft[node] + side;
        }
    }
    // one cursor bump per (node, side) and warp
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs((int)peers) - 1;
    unsigned base = 0;
    if (key >= 0 && lane == leader) base = atomicAdd(S.cursor + key, (unsigned)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (key >= 0) dest = S.nFirst[node] + (side ? S.sLc[j] : 0) + (int)base + __popc(peers & ((1u << lane) - 1u));
    if (p < S.n) { S.order[cur ^ 1][dest] = tri; S.nodeOf[cur ^ 1][dest] = child; }
}

// ---- final passes ----------------------------------------------------------------------------------------------------
// Inner nodes get the host builder's numbers (bvh_build.cpp flattens depth-first, left subtree first, and hands a node's
// inner children consecutive slots when the node is visited), so the GNode array is the host's, record for record:
//   count[x]  inner nodes in the subtree of x                                   bottom-up, one launch per level
//   pre[x]    x's rank among the inner nodes in that visiting order             top-down, one launch per level
//   k[pre[x]] how many of x's children are inner; exclusive scan = slots handed out before x is visited
__global__ void __launch_bounds__(kT) sahSubtreeCounts(const int* __restrict__ nLeft, int base, int A, int* __restrict__ count) {
    const int j = blockIdx.x * kT + threadIdx.x;
    if (j >= A) return;
    const int L = nLeft[base + j];
    count[base + j] = L >= 0 ? 1 + count[L] + count[L + 1] : 0;
}
__global__ void __launch_bounds__(kT) sahPreorder(const int* __restrict__ nLeft, int base, int A, const int* __restrict__ count,
                                                   int* __restrict__ pre, unsigned* __restrict__ k) {
    const int j = blockIdx.x * kT + threadIdx.x;
    if (j >= A) return;
    const int L = nLeft[base + j];
    if (L < 0) return;
    const int p = pre[base + j];
    const bool li = nLeft[L] >= 0, ri = nLeft[L + 1] >= 0;
    if (li) pre[L] = p + 1;
    if (ri) pre[L + 1] = p + 1 + count[L];
    k[p] = (li ? 1u : 0u) + (ri ? 1u : 0u);
}
__global__ void __launch_bounds__(kT) sahAssignSlots(const int* __restrict__ nLeft, int pool, const int* __restrict__ pre,
                                                      const unsigned* __restrict__ kScan, unsigned* __restrict__ gIndex) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i >= pool) return;
    const int L = nLeft[i];
    if (L < 0) return;
    if (i == 0) gIndex[0] = 0u;
    const unsigned base = 1u + kScan[pre[i]];
    const bool li = nLeft[L] >= 0;
    if (li) gIndex[L] = base;
    if (nLeft[L + 1] >= 0) gIndex[L + 1] = base + (li ? 1u : 0u);
}

// leaves: triangles in index order (the order the atomics left them in is not reproducible), largest leaf
__global__ void __launch_bounds__(kT) sahSortLeaves(SahState S, int cur, int pool, int* __restrict__ maxLeaf) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i >= pool || S.nLeft[i] >= 0) return;
    int* e = S.order[cur] + S.nFirst[i];
    const int count = S.nCount[i];
    for (int a = 1; a < count; a++) {
        const int key = e[a];
        int b = a - 1;
        while (b >= 0 && e[b] > key) { e[b + 1] = e[b]; b--; }
        e[b + 1] = key;
    }
    atomicMax(maxLeaf, count);
}

__global__ void __launch_bounds__(kT) sahEmitNodes(SahState S, int pool, const unsigned* __restrict__ gIndex, double cx, double cy,
                                                    double cz, double pad, GNode* __restrict__ nodes) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i >= pool || S.nLeft[i] < 0) return;
    const double centre[3] = {cx, cy, cz};
    float lo[2][3], hi[2][3];
    int32_t ref[2];
    for (int k = 0; k < 2; k++) {
        const int c = S.nLeft[i] + k;
        const double* b = S.nBox + (long long)c * 6;
        for (int a = 0; a < 3; a++) {
            lo[k][a] = __double2float_rd(b[a] - centre[a] - pad);          // rounded outward, as bvh_build.cpp putBox
            hi[k][a] = __double2float_ru(b[3 + a] - centre[a] + pad);
        }
        ref[k] = S.nLeft[c] >= 0 ? (int32_t)gIndex[c] : ~(int32_t)(((unsigned)S.nFirst[c] << 3) | (unsigned)(S.nCount[c] - 1));
    }
    GNode g;
    g.q0[0] = lo[0][0]; g.q0[1] = lo[0][1]; g.q0[2] = lo[0][2]; g.q0[3] = hi[0][0];
    g.q1[0] = hi[0][1]; g.q1[1] = hi[0][2]; g.q1[2] = lo[1][0]; g.q1[3] = lo[1][1];
    g.q2[0] = lo[1][2]; g.q2[1] = hi[1][0]; g.q2[2] = hi[1][1]; g.q2[3] = hi[1][2];
    g.child0 = ref[0]; g.child1 = ref[1]; g.pad0 = 0; g.pad1 = 0;
    nodes[gIndex[i]] = g;
}

__global__ void __launch_bounds__(kT) sahEmitTris(const double* __restrict__ verts, const long long* __restrict__ idx,
                                                   const int* __restrict__ order, int n, GTri* __restrict__ tris, GTriD* __restrict__ trisD) {
    const int p = blockIdx.x * kT + threadIdx.x;
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

// RTCUDA_TRACE=1 (diagnostics, like rt_api.cu's): where a build's host time goes — allocator calls against the rest
bool sahTraceOn() { static const bool on = std::getenv("RTCUDA_TRACE") != nullptr; return on; }
double sahNowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
struct SahTrace { double allocMs = 0, freeMs = 0; int allocs = 0, frees = 0; size_t bytes = 0; } gSahTrace;

struct Buf {
    void* p = nullptr;
    size_t bytes = 0;
    ~Buf() { release(); }
    void release() {
        if (!p) return;
        const double t = sahTraceOn() ? sahNowMs() : 0;
        cudaFree(p);
        if (sahTraceOn()) { gSahTrace.freeMs += sahNowMs() - t; gSahTrace.frees++; }
        p = nullptr;
    }
    cudaError_t alloc(size_t b) {
        bytes = b;
        const double t = sahTraceOn() ? sahNowMs() : 0;
        const cudaError_t e = cudaMalloc(&p, b ? b : 16);
        if (sahTraceOn()) { gSahTrace.allocMs += sahNowMs() - t; gSahTrace.allocs++; gSahTrace.bytes += b; }
        return e;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// One allocation carved into 256-byte aligned pieces.  A build used to make ~220 cudaMalloc / ~200 cudaFree calls
// (every per-level array grown on its own), which was three quarters of its host time (profiles/r2al_build_trace.txt);
// it now makes four of each (two arenas and the two output arrays) for a mesh of up to 2^20 triangles.
struct Arena {
    Buf mem;
    size_t used = 0;
    static size_t pad(size_t b) { return (b + 255) & ~(size_t)255; }
    template <class T> T* take(size_t count) {
        T* r = reinterpret_cast<T*>(static_cast<char*>(mem.p) + used);
        used += pad(count * sizeof(T));
        return r;
    }
};

// bytes of the per-level arrays for a level of `a` nodes (the take<> calls of the level loop, in its order)
inline size_t levelBytes(size_t a) {
    return 2 * Arena::pad(a * 24) + 2 * Arena::pad(a * kSlots * 24) + Arena::pad(a * kSlots * 4) + Arena::pad((a + 1) * 4) +
           3 * Arena::pad(a * 4) + Arena::pad(a * 96) + Arena::pad(a * 8);
}

inline int blocksFor(long long n) { return (int)((n + kT - 1) / kT); }

}  // namespace

// Same contract as buildBlas / buildBlasDevice.  `dVerts`: the scene's triangle vertices (9 doubles each) on the current
// device.  depthLimit: give up (out.maxDepth = 1 << 30) when the tree grows past it — the caller's traversal stack.
// *launches (optional) receives the number of kernels run.
cudaError_t buildBlasDeviceSah(const double* dVerts, const BlasBuildInput& in, BlasBuildOutput& out, int depthLimit, int64_t* launches) {
#define SCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return e_; } while (0)
    out = BlasBuildOutput();
    const int n = (int)in.count;
    if (launches) *launches = 0;
    if (n <= 0) return cudaSuccess;
    const int pool = 2 * n;
    const bool blur = in.motionBlur[0] != 0 || in.motionBlur[1] != 0 || in.motionBlur[2] != 0;
    int64_t nl = 0;
    const double tBuild = sahTraceOn() ? sahNowMs() : 0;
    if (sahTraceOn()) gSahTrace = SahTrace();

    // everything whose size follows from n, in one allocation (the tree has at most pool = 2n nodes, at most n inner)
    Arena fixed;
    const size_t nSums = ((size_t)(pool + 1) / 2048 + 2);
    SCK(fixed.mem.alloc(Arena::pad((size_t)n * 8) + Arena::pad((size_t)n * 48) + 4 * Arena::pad((size_t)n * 4) +
                        6 * Arena::pad((size_t)pool * 4) + Arena::pad((size_t)pool * 48) + Arena::pad(64) + Arena::pad(nSums * 4) +
                        Arena::pad((size_t)(n + 1) * 4)));
    long long* dIdx = fixed.take<long long>((size_t)n);
    double* dTriBox = fixed.take<double>((size_t)n * 6);
    int* dOrder0 = fixed.take<int>((size_t)n);
    int* dOrder1 = fixed.take<int>((size_t)n);
    int* dNodeOf0 = fixed.take<int>((size_t)n);
    int* dNodeOf1 = fixed.take<int>((size_t)n);
    int* dFirst = fixed.take<int>((size_t)pool);
    int* dCount = fixed.take<int>((size_t)pool);
    int* dLeft = fixed.take<int>((size_t)pool);
    unsigned* dGIndex = fixed.take<unsigned>((size_t)pool);
    int* dSubCount = fixed.take<int>((size_t)pool);
    int* dPre = fixed.take<int>((size_t)pool);
    double* dBox = fixed.take<double>((size_t)pool * 6);
    unsigned char* dSmall = fixed.take<unsigned char>(64);      // u64 rootLo[3], rootHi[3]; int wide, maxLeaf
    unsigned* dSums = fixed.take<unsigned>(nSums);
    unsigned* dKArr = fixed.take<unsigned>((size_t)n + 1);
    SCK(cudaMemcpy(dIdx, in.triIndices, (size_t)n * 8, cudaMemcpyHostToDevice));
    {
        unsigned char init[64];
        std::memset(init, 0, sizeof init);
        std::memset(init, 0xff, 24);
        SCK(cudaMemcpy(dSmall, init, sizeof init, cudaMemcpyHostToDevice));
    }
    u64* dRootLo = reinterpret_cast<u64*>(dSmall);
    u64* dRootHi = dRootLo + 3;
    int* dWide = reinterpret_cast<int*>(dRootLo + 6);
    int* dMaxLeaf = dWide + 1;

    RT_LAUNCH(sahPrep, blocksFor(n), kT, 0, dVerts, dIdx, n, in.motionBlur[0], in.motionBlur[1], in.motionBlur[2],
              blur ? 1 : 0, dTriBox, dRootLo, dRootHi, dWide, dOrder0, dNodeOf0);
    nl++;
    SCK(cudaGetLastError());
    unsigned char back[64];
    SCK(cudaMemcpy(back, dSmall, sizeof back, cudaMemcpyDeviceToHost));
    u64 enc[6];
    std::memcpy(enc, back, 48);
    int wide = 0;
    std::memcpy(&wide, back + 48, 4);
    out.f64 = wide != 0;
    double rootBox[6];
    double M = 0;
    for (int a = 0; a < 3; a++) {
        out.rootLo[a] = rootBox[a] = decHost(enc[a]);
        out.rootHi[a] = rootBox[3 + a] = decHost(enc[3 + a]);
        out.centre[a] = 0.5 * (out.rootLo[a] + out.rootHi[a]);
        M = std::fmax(M, std::fmax(std::fabs(out.rootLo[a] - out.centre[a]), std::fabs(out.rootHi[a] - out.centre[a])));
    }
    out.pad = std::fmax(M * std::ldexp(1.0, -19), 1e-30);
    {
        const int first = 0, left = -1;
        SCK(cudaMemcpy(dFirst, &first, 4, cudaMemcpyHostToDevice));
        SCK(cudaMemcpy(dCount, &n, 4, cudaMemcpyHostToDevice));
        SCK(cudaMemcpy(dLeft, &left, 4, cudaMemcpyHostToDevice));
        SCK(cudaMemcpy(dBox, rootBox, 48, cudaMemcpyHostToDevice));
    }

    SahState S{};
    S.n = n;
    S.triBox = dTriBox;
    S.order[0] = dOrder0; S.order[1] = dOrder1;
    S.nodeOf[0] = dNodeOf0; S.nodeOf[1] = dNodeOf1;
    S.nFirst = dFirst; S.nCount = dCount; S.nLeft = dLeft; S.nBox = dBox;

    // The per-level arrays (centroid boxes, bins, split choices) of a level of A nodes, carved from one allocation.  A
    // level has at most n nodes (each holds a triangle) and the widest level of a SAH tree with 1-2 triangle leaves is
    // close to n / 2, so the allocation starts at that bound — up to 2^20 nodes (2.7 GB: the bins are 2.5 KB a node) — and
    // is re-made 16 times larger if a level of a bigger mesh outgrows it.  Driver memory calls are what stalls a build
    // (one cudaFree of 1.2 s among 4 ms of them, profiles/r2an_build_trace.txt), so there are as few as possible.
    Arena level;
    size_t levelCap = 0;
    const size_t levelCap0 = std::min((size_t)n, (size_t)1 << 20);
    int levelBase = 0, A = 1, depth = 1, cur = 0, used = 1;
    std::vector<std::pair<int, int>> levels;          // (first pool node, nodes) of every level
    for (;;) {
        levels.emplace_back(levelBase, A);
        const size_t a = (size_t)A;
        if (a > levelCap) {
            levelCap = std::min(std::max(a * 16, levelCap0), std::max((size_t)n, a));
            level.mem.release();                        // kernels of earlier levels still running: cudaFree waits for them
            SCK(level.mem.alloc(levelBytes(levelCap)));
        }
        level.used = 0;
        S.cbLo = level.take<u64>(a * 3); S.cbHi = level.take<u64>(a * 3);
        S.binLo = level.take<u64>(a * kSlots * 3); S.binHi = level.take<u64>(a * kSlots * 3);
        S.binCnt = level.take<unsigned>(a * kSlots);
        S.flag = level.take<unsigned>(a + 1);
        S.sAxis = level.take<int>(a); S.sBin = level.take<int>(a); S.sLc = level.take<int>(a);
        S.childBox = level.take<double>(a * 12);
        S.cursor = level.take<unsigned>(a * 2);
        SCK(cudaMemsetAsync(S.cbLo, 0xff, a * 24, 0)); SCK(cudaMemsetAsync(S.cbHi, 0, a * 24, 0));
        SCK(cudaMemsetAsync(S.binLo, 0xff, a * kSlots * 24, 0)); SCK(cudaMemsetAsync(S.binHi, 0, a * kSlots * 24, 0));
        SCK(cudaMemsetAsync(S.binCnt, 0, a * kSlots * 4, 0));
        SCK(cudaMemsetAsync(S.flag, 0, (a + 1) * 4, 0)); SCK(cudaMemsetAsync(S.cursor, 0, a * 8, 0));
        RT_LAUNCH(sahCentroids, blocksFor(n), kT, 0, S, cur, levelBase);
        RT_LAUNCH(sahBin, blocksFor(n), kT, 0, S, cur, levelBase);
        RT_LAUNCH(sahSplit, blocksFor(A), kT, 0, S, levelBase, A);
        SCK(launchExclusiveScan(S.flag, (long long)A + 1, dSums, 0));
        nl += 6;
        SCK(cudaGetLastError());
        unsigned splits = 0;
        SCK(cudaMemcpy(&splits, S.flag + A, 4, cudaMemcpyDeviceToHost));
        if (splits == 0) break;
        if (depth + 1 > depthLimit || used + 2 * (long long)splits > pool) { out.maxDepth = 1 << 30; return cudaSuccess; }
        RT_LAUNCH(sahChildren, blocksFor(A), kT, 0, S, levelBase, A);
        RT_LAUNCH(sahPartition, blocksFor(n), kT, 0, S, cur, levelBase);
        nl += 2;
        SCK(cudaGetLastError());
        levelBase += A;
        A = 2 * (int)splits;
        used += A;
        depth++;
        cur ^= 1;
    }
    out.maxDepth = depth;
    if (sahTraceOn()) {
        cudaDeviceSynchronize();
        std::fprintf(stderr, "[rtcuda trace] sah: %d levels                  %8.2f ms (cudaMalloc x%d %.2f ms, %zu MB; cudaFree x%d %.2f ms)\n", depth,
                     sahNowMs() - tBuild, gSahTrace.allocs, gSahTrace.allocMs, gSahTrace.bytes >> 20, gSahTrace.frees, gSahTrace.freeMs);
    }

    // inner nodes -> GNode slots in the host builder's order (above)
    Buf dNodes, dTris;
    for (size_t l = levels.size(); l-- > 0;) {
        RT_LAUNCH(sahSubtreeCounts, blocksFor(levels[l].second), kT, 0, S.nLeft, levels[l].first, levels[l].second, dSubCount);
        nl++;
    }
    int inner = 0;
    SCK(cudaMemcpy(&inner, dSubCount, 4, cudaMemcpyDeviceToHost));
    if (inner > n) return cudaErrorUnknown;             // cannot happen: a binary tree over n triangles has fewer than n inner nodes
    if (inner > 0) {
        SCK(cudaMemsetAsync(dKArr, 0, (size_t)(inner + 1) * 4, 0));
        SCK(cudaMemsetAsync(dPre, 0, 4, 0));
        for (size_t l = 0; l < levels.size(); l++) {
            RT_LAUNCH(sahPreorder, blocksFor(levels[l].second), kT, 0, S.nLeft, levels[l].first, levels[l].second, dSubCount, dPre, dKArr);
            nl++;
        }
        SCK(launchExclusiveScan(dKArr, (long long)inner + 1, dSums, 0));
        RT_LAUNCH(sahAssignSlots, blocksFor(used), kT, 0, S.nLeft, used, dPre, dKArr, dGIndex);
        nl += 4;
    }
    RT_LAUNCH(sahSortLeaves, blocksFor(used), kT, 0, S, cur, used, dMaxLeaf);
    nl++;
    SCK(cudaGetLastError());
    SCK(cudaMemcpy(&out.maxLeaf, dMaxLeaf, 4, cudaMemcpyDeviceToHost));

    SCK(dTris.alloc((size_t)n * (out.f64 ? sizeof(GTriD) : sizeof(GTri))));
    RT_LAUNCH(sahEmitTris, blocksFor(n), kT, 0, dVerts, dIdx, S.order[cur], n, out.f64 ? nullptr : dTris.as<GTri>(),
              out.f64 ? dTris.as<GTriD>() : nullptr);
    nl++;
    SCK(cudaGetLastError());
    if (in.keepOnDevice) { out.dTris = dTris.p; out.nTrisDevice = n; dTris.p = nullptr; }       // the caller's from here on
    else if (out.f64) { out.trisD.resize((size_t)n); SCK(cudaMemcpy(out.trisD.data(), dTris.p, (size_t)n * sizeof(GTriD), cudaMemcpyDeviceToHost)); }
    else { out.tris.resize((size_t)n); SCK(cudaMemcpy(out.tris.data(), dTris.p, (size_t)n * sizeof(GTri), cudaMemcpyDeviceToHost)); }
    if (inner == 0) {                                    // the whole BLAS is one leaf
        out.rootRef = ~(int32_t)((0u << 3) | (unsigned)(n - 1));
    } else {
        SCK(dNodes.alloc((size_t)inner * sizeof(GNode)));
        RT_LAUNCH(sahEmitNodes, blocksFor(used), kT, 0, S, used, dGIndex, out.centre[0], out.centre[1], out.centre[2], out.pad,
                  dNodes.as<GNode>());
        nl++;
        SCK(cudaGetLastError());
        if (in.keepOnDevice) {
            SCK(cudaDeviceSynchronize());
            out.dNodes = dNodes.p; out.nNodesDevice = inner; dNodes.p = nullptr;
        } else {
            out.nodes.resize((size_t)inner);
            SCK(cudaMemcpy(out.nodes.data(), dNodes.p, (size_t)inner * sizeof(GNode), cudaMemcpyDeviceToHost));
        }
        out.rootRef = 0;
    }
    if (launches) *launches = nl;
    if (sahTraceOn()) {
        cudaDeviceSynchronize();
        std::fprintf(stderr, "[rtcuda trace] sah: emitted                    %8.2f ms (cudaMalloc x%d %.2f ms; cudaFree x%d %.2f ms)\n", sahNowMs() - tBuild,
                     gSahTrace.allocs, gSahTrace.allocMs, gSahTrace.frees, gSahTrace.freeMs);
    }
    return cudaSuccess;
#undef SCK
}

}  // namespace rtg
