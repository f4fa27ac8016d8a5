// bvh_build.h — host-side builder of the GPU BLAS layout (gpu_layout.h).
//
// Takes over the role of the reference's BVHBuilder for meshes
// (Accelearion/BVH.swift:78-250, called from RTContext.swift:430-435) but is a
// different tree: binned SAH with a leaf-cost termination and up to 4
// triangles per leaf, children boxes stored in the parent, FP32 boxes relative
// to the BLAS centre, padded + rounded outward (conservative), triangles
// re-ordered into leaf order.  Closest-hit results do not depend on the tree
// (SURVEY.md §7.3-4): only equal-t ties may resolve differently.
#pragma once
#include <cstdint>
#include <vector>

#include "gpu_layout.h"

namespace rtg {

struct BlasBuildInput {
    const double* triVertices;      // global array, 9 doubles per triangle
    const int64_t* triIndices;      // triangles of this BLAS (global indices), in prim order
    int64_t count;
    double motionBlur[3];           // box growth for Triangle.motionBlur (RTContext.swift:283-289)
    int threads;                    // 0 = hardware concurrency
    bool keepOnDevice = false;      // buildBlasDeviceSah: leave the node / triangle arrays in device memory (dNodes, dTris)
};

struct BlasBuildOutput {
    std::vector<GNode> nodes;       // inner nodes, node 0 = root when rootRef == 0
    std::vector<GTri> tris;         // leaf order (when !f64)
    std::vector<GTriD> trisD;       // leaf order (when f64)
    bool f64 = false;
    int32_t rootRef = kEmptyRef;
    double rootLo[3], rootHi[3];    // exact FP64 bounds (incl. motion blur growth)
    double centre[3];
    double pad = 0;                 // absolute padding applied to every FP32 box
    int maxDepth = 0;
    int maxLeaf = 0;
    // keepOnDevice: the arrays stay where the builder made them (the caller owns them: cudaFree); nodes / tris / trisD are
    // then empty.  References inside are BLAS-relative like the host vectors'.
    void* dNodes = nullptr;         // GNode[nNodesDevice]
    void* dTris = nullptr;          // GTri[nTrisDevice] or GTriD[nTrisDevice] (f64)
    int64_t nNodesDevice = 0, nTrisDevice = 0;
};

void buildBlas(const BlasBuildInput& in, BlasBuildOutput& out);

#if defined(__CUDACC__) || defined(RT_WITH_CUDA_RUNTIME)
}  // namespace rtg
#include <cuda_runtime.h>
namespace rtg {
// Device-side construction of the same output (bvh_build_gpu.cu): Morton order + binary radix tree, leaves of <= 2
// triangles.  dVerts: the scene's triangle vertices (9 doubles each) on the current device.  The caller checks
// out.maxDepth against its traversal stack (a radix tree over clustered keys can be deep) and may fall back to buildBlas.
cudaError_t buildBlasDevice(const double* dVerts, const BlasBuildInput& in, BlasBuildOutput& out);
// Device-side binned SAH (bvh_build_sah_gpu.cu): buildBlas's own splits and leaves, one tree level per round of
// kernels.  Grows the boxes by in.motionBlur like buildBlas.  out.maxDepth = 1 << 30 when the tree would grow past
// depthLimit levels.  *launches (optional): kernels run.
cudaError_t buildBlasDeviceSah(const double* dVerts, const BlasBuildInput& in, BlasBuildOutput& out, int depthLimit, int64_t* launches);
#endif

// Wide BVH of the beam walk (rt_beam.cuh, gpu_layout.h: WNode): the binary tree of any of the builders collapsed into
// nodes of up to kWide children — starting from a node's two children, the inner child with the largest box is
// replaced by its own two children until kWide are reached or only leaves are left.  Boxes and leaf references are
// taken over unchanged, so the leaves (and the triangle order) are the binary tree's.  References stay relative to
// the BLAS like the input's.  rootRef = kEmptyRef (no wide tree) when the binary root is a leaf.
struct WideBuildOutput {
    std::vector<WNode> nodes;       // node 0 = root
    int32_t rootRef = kEmptyRef;
    int depth = 0;                  // levels of wide nodes
};
void collapseWide(const std::vector<GNode>& nodes, int32_t rootRef, WideBuildOutput& out);

// TLAS over instance world bounds (FP64, one instance per leaf).
struct TlasBuildOutput {
    std::vector<TNode> nodes;
    int32_t rootRef = kEmptyRef;
    int maxDepth = 0;
};
void buildTlas(const double* worldBounds /* n*6 */, int64_t n, TlasBuildOutput& out);

// The reference's OWN TLAS over the instances' world bounds, for scenes with instanceMotion only (rt_trace.cuh:
// traceSceneRef explains why its topology matters there): buildTLAS (Models/RTContext.swift:459-474) =
// BVHBuilder(maxLeaf: 2, split: .sah, binCount: 12) (Accelearion/BVH.swift:78-250), restated operation by operation
// so that nodes, boxes and the primitiveIdx permutation come out identical.
struct RefTlasOutput {
    std::vector<RefNode> nodes;         // node 0 = root; children of an inner node are leftFirst, leftFirst + 1
    std::vector<int32_t> primIdx;       // instance indices in leaf order
};
void buildReferenceTlas(const double* worldBounds /* n*6 */, int64_t n, RefTlasOutput& out);

}  // namespace rtg
