// gpu_layout.h — device-side data layout of the render module (shared by the
// host builder and the CUDA kernels).
//
// Replaces the reference's runtime layout: 96-B FP64 AoS
// BVHNode + primIdx -> 160-B PrimitiveInfo -> >=320-B Triangle
// (Accelearion/BVH.swift:16-64, Models/RTContext.swift:34-92) with
//   * GNode      64 B : one inner BLAS node holding BOTH children's FP32 boxes
//                       (4 x 16-B vector loads, one 64-B aligned segment);
//                       boxes are relative to the BLAS centre, padded and rounded
//                       outward so the FP32 slab test is conservative;
//   * GTri       48 B : float-exact vertices in BVH leaf order (no primIdx hop),
//                       original triangle index in v0.w;  GTriD 80 B for meshes
//                       whose vertices are not float-exact;
//   * TNode     112 B : TLAS inner node, FP64 boxes = Instance.worldBounds unions
//                       (the TLAS is tiny; FP64 keeps the reference's culling exact);
//   * GInstance       : FP64 world<->local matrices, exact local root box, kind.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RT_ALIGN(n) __align__(n)
#else
#define RT_ALIGN(n) alignas(n)
#endif

namespace rtg {

// Child reference: >= 0 inner node index; < 0 leaf: ~ref = (first << 3) | (count - 1), count in 1..8;
// kEmptyRef = no child.  The builder emits both relative to one BLAS (node 0 = its first node, triangle 0 = its
// first leaf-ordered triangle); on the device they are ABSOLUTE in DeviceScene::nodes and ::tris / ::trisD
// (rt_api.cu rebases them when the BLASes are concatenated), with bit 3 of a leaf's ~ref telling which
// triangle array: ~ref = (first << 4) | (f64 << 3) | (count - 1), count in 1..kMaxLeaf.  The traversal then
// needs nothing from the instance record inside its loop.
static const int32_t kEmptyRef = 0x7fffffff;
static const int kMaxLeaf = 4;
static const int kBigBlasTriangles = 16384;

struct RT_ALIGN(16) GNode {
    // q0 = lo0.x lo0.y lo0.z hi0.x | q1 = hi0.y hi0.z lo1.x lo1.y | q2 = lo1.z hi1.x hi1.y hi1.z
    float q0[4], q1[4], q2[4];
    int32_t child0, child1, pad0, pad1;
};

// Wide node of the beam walk (rt_beam.cuh): kWide children, one lane of the warp per child.  A child is one 32-byte
// record — two 16-byte vector loads per lane, the warp's loads covering the node's 512 contiguous bytes: the GNode box
// of the collapsed BVH2 (same centre, same padding, same outward rounding) and `ref`, a wide node index (>= 0), a leaf
// reference (< 0, as in GNode) or kEmptyRef with an inverted box.  The children are sorted by their centre along
// the axis on which the node's children spread the most (child[0].aux), empty slots last: a beam travelling up that
// axis meets them in slot order, one travelling down it in reverse order.  No counterpart in the reference (its
// BVHNode is the 96-byte binary node of Accelearion/BVH.swift:16-30).
#ifndef RT_WIDE
#define RT_WIDE 16
#endif
static const int kWide = RT_WIDE;
struct RT_ALIGN(16) WChild {
    float lo[3], hi0;        // lo.x lo.y lo.z hi.x
    float hi1, hi2;          // hi.y hi.z
    int32_t ref;
    int32_t aux;             // child 0: the node's sort axis (0..2)
};
struct RT_ALIGN(128) WNode {
    WChild c[RT_WIDE];
};

struct RT_ALIGN(16) GTri {   // float-exact vertices (PLY meshes: Helpers/PLYReader.swift:85-101)
    float v0[3]; int32_t triIndex;
    float v1[3]; int32_t pad1;
    float v2[3]; int32_t pad2;
};

struct RT_ALIGN(16) GTriD {  // FP64 vertices
    double v0[3], v1[3], v2[3];
    int64_t triIndex;
};

struct RT_ALIGN(16) TNode {             // TLAS builder output (FP64); rt_api.cu encodes it into GNode records
    double lo0[3], hi0[3], lo1[3], hi1[3];
    int32_t child0, child1, pad0, pad1;  // leaf: ~ref = instance index (one instance per leaf)
};

enum InstKind : int32_t { INST_MESH = 0, INST_SPHERE = 2, INST_PLANE = 3, INST_TRI = 1 };

struct RT_ALIGN(16) GInstance {
    double w2l[12];        // rows 0..2 of worldToLocal, column-major 3x4: w2l[c*3+r], c = 0..3
    double l2w[12];        // same for localToWorld
    double nmat[9];        // normalMatrix, column-major
    double rootLo[3], rootHi[3];  // exact local-space bounds (= the reference's BLAS root node box)
    double centre[3];      // BLAS box centre the FP32 node boxes are relative to
    double motion[3];      // instanceMotion
    double prim[8];        // analytic payload: sphere c.xyz,r | plane c.xyz,n.xyz | unused for meshes
    int64_t matId;         // hit.mat value (materialOverride ?? primitive materialID)
    int64_t primIndex;     // analytic: primitiveIndex; mesh: unused
    int32_t kind;          // InstKind
    int32_t identity;      // 1: worldToLocal is exactly the identity (skip the transform)
    int32_t flipNormal;    // det(localToWorld 3x3) < 0  (RTContext.swift:691-698)
    int32_t smooth;        // mesh shading mode
    int32_t nodeBase;      // first GNode of this BLAS in the global node array
    int32_t triBase;       // first GTri/GTriD slot of this BLAS
    int32_t triF64;        // 1: GTriD array, 0: GTri array
    int32_t rootRef;       // child reference of the BLAS root (inner node 0 or a leaf)
    int32_t hasMotion;     // instanceMotion != 0 or mesh motion blur != 0
    int32_t bigBlas;       // mesh with more than kBigBlasTriangles triangles: incoherent packets walk it per lane (rt_packet.cuh)
    int32_t wideRoot;      // root of the mesh's wide BVH in DeviceScene::wnodes (beam walk, rt_beam.cuh); kEmptyRef: none
    int32_t pad;
    double meshMotion[3];  // Triangle.motionBlur of this mesh
    double beamSlack;      // absolute widening of a beam in this BLAS's space (rt_beam.cuh): 2^-15 of the largest |coordinate - centre|
};

struct GMaterial {
    double ambient[3], diffuse[3], specular[3], mirror[3], absorption[3];
    double phong, ior, absorptionIndex, roughness;
    int32_t type, hasAbsorption;
};

// One node of the reference-topology TLAS (scenes with instanceMotion only; rt_trace.cuh: traceSceneRef): the
// reference's BVHNode (Accelearion/BVH.swift:16-30) — exact FP64 box, leftFirst = left child (right = left + 1) or
// first entry of refTlasPrim, count > 0 for a leaf.
struct RT_ALIGN(16) RefNode {
    double lo[3], hi[3];
    int32_t leftFirst, count, pad0, pad1;
};

struct GPointLight { double position[3], intensity[3]; };
struct GAreaLight { double position[3], normal[3], size, radiance[3]; };   // Object+Extension.swift:145-186

// Everything the kernels read, passed by value as a kernel parameter.
struct DeviceScene {
    const GNode* nodes;
    const GTri* tris;
    const GTriD* trisD;
    const double* triNormals;      // original triangle index * 9
    const double* triVertsD;       // original triangle index * 9 (FP64 v0,v1,v2; hit point / flat normal)
    const GNode* tlas;             // TLAS in the BLAS node format: conservative FP32 boxes relative to tlasCentre;
                                   // child < 0: ~ref = instance index (one instance per leaf)
    const GInstance* instances;
    const GMaterial* materials;
    const GPointLight* pointLights;
    const GAreaLight* areaLights;
    const double* jitter;          // area lights: the 10x10 stratified table, xs[100] then ys[100] (:92-93,646-659)
    const RefNode* refTlas;        // scenes with instanceMotion: the reference's own TLAS topology (else null)
    const int32_t* refTlasPrim;    // its primitiveIdx array (instance indices in leaf order)
    const WNode* wnodes;           // wide BVHs of the meshes (collapsed from `nodes`), all BLASes concatenated
    int32_t refTlasNodes;          // 0: not used
    int32_t wideMode;              // beam walk (RTCUDA_WIDE): 1 every coherent packet, 2 closest-hit packets only, 3 any-hit packets only
    int32_t tlasRootRef;
    int32_t nInstances, nMaterials, nPointLights;
    int32_t maxDepth;
    int32_t nAreaLights;
    double ambient[3], background[3];
    double eps, shadowEps;
    double tlasLo[3], tlasHi[3], tlasCentre[3];   // exact world bounds of all instances and their centre
};

struct CameraParams {
    double eye[3], u[3], v[3], w[3];
    double q00[3], m[3];          // q00 = m + u*l + v*t ; m = eye - w*nd
    double du, dv;
    double aperture, focus;
    int32_t width, height, samples, n;  // n = floor(sqrt(samples))
    int32_t dof;
    int32_t shardIndex, shardCount;
    int32_t shardTiles;           // 0: 8-row chunks dealt round-robin, 1: 8x4 tiles dealt along diagonals (rt_render_params.shardMode)
};

#if defined(__CUDACC__)
#define RT_HD __host__ __device__
#else
#define RT_HD
#endif

// The tiles (8x4 pixels) one shard renders, numbered 0..count-1 in the order they are drawn without a schedule.
//   rows mode   chunk c (8 rows = two tile rows) belongs to shard c % shardCount — the reference's decomposition
//               (Object+Extension.swift:75-82) dealt round-robin
//   tiles mode  tile (tx, ty) belongs to shard (tx + ty) % shardCount: every shard samples the whole image at tile
//               granularity, so the shards' costs differ by a fraction of a percent whatever the image shows; the
//               tile grid is padded to a multiple of shardCount columns (tiles outside the image trace nothing)
struct ShardTiles {
    int tilesX, perRow, count, shardIndex, shardCount, byTile;
    RT_HD ShardTiles(int width, int height, int index, int n, int tilesMode) {
        tilesX = (width + 7) >> 3;
        const int nChunks = (height + 7) >> 3;
        shardCount = n < 1 ? 1 : n;
        shardIndex = index;
        byTile = tilesMode && shardCount > 1;
        if (byTile) {
            perRow = (tilesX + shardCount - 1) / shardCount;
            count = perRow * nChunks * 2;
        } else {
            perRow = tilesX;
            count = (nChunks - shardIndex + shardCount - 1) / shardCount * tilesX * 2;
        }
    }
    // tile -> pixel origin (x0, y0) of its 8x4 block
    RT_HD void origin(int tile, int& x0, int& y0) const {
        if (byTile) {
            const int ty = tile / perRow, j = tile - ty * perRow;
            const int first = (shardIndex + shardCount - ty % shardCount) % shardCount;
            x0 = (first + j * shardCount) * 8;
            y0 = ty * 4;
        } else {
            const int k = tile / (tilesX * 2), rem = tile - k * (tilesX * 2);
            const int ty = rem / tilesX, tx = rem - ty * tilesX;
            x0 = tx * 8;
            y0 = (k * shardCount + shardIndex) * 8 + ty * 4;
        }
    }
};

// Work units of the fast kernel: `passes` consecutive passes of 32 (pixel, sample) items of a tile, `perTile` units per
// 8x4 tile.  ns = samples per pixel.  ns divides 32: a pass holds whole pixels, so one pass is a unit — or, for
// scenes with recursion, up to RT_BATCH_PASSES passes, whose child rays are then traced together; 32 divides ns: the
// ns / 32 passes of one pixel; otherwise the whole tile (the ordered per-pixel sum is carried across its ns passes).
#ifndef RT_BATCH_PASSES
#define RT_BATCH_PASSES 1    // measured on C3 (B200): 1 -> 6.0 ms, 2 -> 9.8 ms, 4 -> 17.5 ms: a unit's rounds are serial, long units are the tail
#endif
struct FastUnits {
    int passes, perTile, perPixel;
    RT_HD FastUnits(int ns, bool recurse) {
        perPixel = 0;
        if (ns < 32 && 32 % ns == 0) { passes = recurse ? (ns < RT_BATCH_PASSES ? ns : RT_BATCH_PASSES) : 1; perTile = ns / passes; }
        else if (ns % 32 == 0) { passes = ns >> 5; perTile = 32; perPixel = 1; }
        else { passes = ns; perTile = 1; }
    }
};

}  // namespace rtg
