// oracle/bvh.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Restates the reference's acceleration-structure builder:
//   AABB                Accelearion/AABB.swift:12-92
//   BVHNode             Accelearion/BVH.swift:16-30
//   PrimitiveInfo       Accelearion/BVH.swift:33-64
//   BVHBuilder          Accelearion/BVH.swift:67-250  (binned SAH, 12 bins,
//                       maxLeaf 2, no leaf-cost termination)
//   refit               Accelearion/BVH.swift:254-292
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

#include "vecmath.hpp"

namespace orc {

// AABB.swift:12-29 (default = empty box), :31-35 area, :62-70 grow, :71-92 transformed
struct AABB {
    Vec3 minP{kInf, kInf, kInf};
    Vec3 maxP{-kInf, -kInf, -kInf};
    AABB() {}
    AABB(Vec3 a, Vec3 b) : minP(a), maxP(b) {}
    double area() const {
        Vec3 e = maxP - minP;
        return 2.0 * ((e.x * e.y + e.y * e.z) + e.z * e.x);
    }
    void grow(const AABB& b) { minP = vmin(minP, b.minP); maxP = vmax(maxP, b.maxP); }
    // AABB.swift:71-92: centre/half-extent transformed by R and |R|.
    AABB transformed(const Mat4& M) const {
        Vec3 cOld = 0.5 * (minP + maxP);
        Vec3 eOld = 0.5 * (maxP - minP);
        Mat3 r = upper3(M);
        Vec3 t{M.m[12], M.m[13], M.m[14]};
        Vec3 cNew = mul(r, cOld) + t;
        Mat3 ar;
        ar.c[0] = vabs(r.c[0]); ar.c[1] = vabs(r.c[1]); ar.c[2] = vabs(r.c[2]);
        Vec3 eNew{(ar.c[0].x * eOld.x + ar.c[1].x * eOld.y) + ar.c[2].x * eOld.z,
                  (ar.c[0].y * eOld.x + ar.c[1].y * eOld.y) + ar.c[2].y * eOld.z,
                  (ar.c[0].z * eOld.x + ar.c[1].z * eOld.y) + ar.c[2].z * eOld.z};
        return AABB(cNew - eNew, cNew + eNew);
    }
};

// BVH.swift:16-30
struct BVHNode {
    Vec3 aabbMin{0, 0, 0};
    Vec3 aabbMax{0, 0, 0};
    uint64_t leftFirst = 0;
    uint64_t primitiveCount = 0;
    bool isLeaf() const { return primitiveCount > 0; }
};

// BVH.swift:33-40
enum GeometryType : uint8_t { GEO_MESH = 0, GEO_TRIANGLE = 1, GEO_SPHERE = 2, GEO_PLANE = 3 };
enum ShadingMode : uint8_t { SHADE_FLAT = 0, SHADE_SMOOTH = 1 };

// BVH.swift:42-64
struct PrimitiveInfo {
    GeometryType type = GEO_MESH;
    int64_t primitiveIndex = 0;
    AABB bounds;
    Vec3 centroid;
    int64_t materialID = 0;
    ShadingMode shadingMode = SHADE_SMOOTH;
};

// BVH.swift:67-250.  Only split == .sah is used by the render path
// (RTContext.swift:432,472 and the default at BVH.swift:80).
struct BVHBuilder {
    std::vector<PrimitiveInfo> primitives;
    std::vector<uint64_t> primitiveIdx;
    std::vector<BVHNode> bvhNode;
    int maxLeaf = 2;
    uint64_t rootNodeIdx = 0;
    uint64_t nodesUsed = 0;
    int binCount = 12;
    int maxDepthSeen = 0;  // oracle extra: recursion depth reached (no effect on the tree)

    BVHBuilder() {}
    // BVH.swift:78-104
    explicit BVHBuilder(std::vector<PrimitiveInfo> prims, int maxLeaf_ = 2, int binCount_ = 12)
        : primitives(std::move(prims)), maxLeaf(maxLeaf_), binCount(binCount_ < 2 ? 2 : binCount_) {
        size_t n = primitives.size();
        primitiveIdx.resize(n);
        for (size_t i = 0; i < n; i++) primitiveIdx[i] = i;
        size_t nodeCount = n * 2 >= 2 ? n * 2 - 1 : 1;  // max(1, 2n-1)
        bvhNode.assign(nodeCount, BVHNode());
        bvhNode[0].leftFirst = 0;
        bvhNode[0].primitiveCount = n;
        nodesUsed = 0;
        rootNodeIdx = 0;
        updateNodeBounds(0);
        subdivide(0, 1);
    }

    // BVH.swift:108-124
    void updateNodeBounds(uint64_t nodeIdx) {
        BVHNode node = bvhNode[nodeIdx];
        node.aabbMin = Vec3(kInf);
        node.aabbMax = Vec3(-kInf);
        size_t first = node.leftFirst, count = node.primitiveCount;
        for (size_t i = 0; i < count; i++) {
            const PrimitiveInfo& prim = primitives[primitiveIdx[first + i]];
            node.aabbMin = vmin(node.aabbMin, prim.bounds.minP);
            node.aabbMax = vmax(node.aabbMax, prim.bounds.maxP);
        }
        bvhNode[nodeIdx] = node;
    }

    // BVH.swift:128-188
    void subdivide(uint64_t nodeIdx, int depth) {
        if (depth > maxDepthSeen) maxDepthSeen = depth;
        BVHNode node = bvhNode[nodeIdx];
        int64_t primCount = (int64_t)node.primitiveCount;
        if (primCount <= maxLeaf && nodeIdx != 0) return;

        int bestAxis = 0;
        double bestPos = 0;
        findBestSplitPlane(node, bestAxis, bestPos);

        int64_t i = (int64_t)node.leftFirst;
        int64_t j = i + primCount - 1;
        while (i <= j) {
            uint64_t primID = primitiveIdx[i];
            if (primitives[primID].centroid[bestAxis] < bestPos) {
                i += 1;
            } else {
                std::swap(primitiveIdx[i], primitiveIdx[j]);
                j -= 1;
            }
        }
        int64_t leftCount = i - (int64_t)node.leftFirst;
        if (leftCount == 0 || leftCount == primCount) return;

        nodesUsed += 1;
        uint64_t leftChild = nodesUsed;
        nodesUsed += 1;
        uint64_t rightChild = nodesUsed;

        bvhNode[leftChild].leftFirst = node.leftFirst;
        bvhNode[leftChild].primitiveCount = (uint64_t)leftCount;
        bvhNode[rightChild].leftFirst = (uint64_t)i;
        bvhNode[rightChild].primitiveCount = (uint64_t)(primCount - leftCount);

        node.leftFirst = leftChild;
        node.primitiveCount = 0;
        bvhNode[nodeIdx] = node;

        updateNodeBounds(leftChild);
        updateNodeBounds(rightChild);
        subdivide(leftChild, depth + 1);
        subdivide(rightChild, depth + 1);
    }

    // BVH.swift:192-250
    double findBestSplitPlane(const BVHNode& node, int& axis, double& splitPos) const {
        double bestCost = kInf;
        const int N = binCount;
        std::vector<AABB> bin(N);
        std::vector<int64_t> cnt(N);
        std::vector<double> leftArea(N - 1), rightArea(N - 1);
        std::vector<int64_t> leftCnt(N - 1), rightCnt(N - 1);
        for (int a = 0; a < 3; a++) {
            double boundsMin = kInf, boundsMax = -kInf;
            size_t count = node.primitiveCount, first = node.leftFirst;
            for (size_t i = 0; i < count; i++) {
                const PrimitiveInfo& prim = primitives[primitiveIdx[first + i]];
                boundsMin = dmin(boundsMin, prim.centroid[a]);
                boundsMax = dmax(boundsMax, prim.centroid[a]);
            }
            if (boundsMax <= boundsMin) continue;

            for (int k = 0; k < N; k++) { bin[k] = AABB(); cnt[k] = 0; }
            double scale = double(N) / (boundsMax - boundsMin);
            for (size_t i = 0; i < count; i++) {
                const PrimitiveInfo& prim = primitives[primitiveIdx[first + i]];
                int64_t idx = (int64_t)((prim.centroid[a] - boundsMin) * scale);  // Int(): toward zero
                if (idx > N - 1) idx = N - 1;
                cnt[idx] += 1;
                bin[idx].grow(prim.bounds);
            }

            AABB L, R;
            int64_t sL = 0, sR = 0;
            for (int i = 0; i < N - 1; i++) {
                sL += cnt[i];
                leftCnt[i] = sL;
                L.grow(bin[i]);
                leftArea[i] = L.area();

                sR += cnt[N - 1 - i];
                rightCnt[N - 2 - i] = sR;
                R.grow(bin[N - 1 - i]);
                rightArea[N - 2 - i] = R.area();
            }

            double step = (boundsMax - boundsMin) / double(N);
            for (int i = 0; i < N - 1; i++) {
                if (leftCnt[i] == 0 || rightCnt[i] == 0) continue;
                double cost = double(leftCnt[i]) * leftArea[i] + double(rightCnt[i]) * rightArea[i];
                if (cost < bestCost) {
                    bestCost = cost;
                    axis = a;
                    splitPos = boundsMin + step * double(i + 1);
                }
            }
        }
        return bestCost;
    }

    // BVH.swift:254-292
    void refit(const std::vector<PrimitiveInfo>& updated) {
        primitives = updated;
        if (bvhNode.empty()) return;
        refitNode(0);
    }
    void refitNode(uint64_t idx) {
        BVHNode node = bvhNode[idx];
        if (node.isLeaf()) {
            Vec3 mn(kInf), mx(-kInf);
            for (size_t i = 0; i < node.primitiveCount; i++) {
                const PrimitiveInfo& prim = primitives[primitiveIdx[node.leftFirst + i]];
                mn = vmin(mn, prim.bounds.minP);
                mx = vmax(mx, prim.bounds.maxP);
            }
            node.aabbMin = mn; node.aabbMax = mx;
            bvhNode[idx] = node;
        } else {
            uint64_t l = node.leftFirst, r = l + 1;
            refitNode(l);
            refitNode(r);
            node.aabbMin = vmin(bvhNode[l].aabbMin, bvhNode[r].aabbMin);
            node.aabbMax = vmax(bvhNode[l].aabbMax, bvhNode[r].aabbMax);
            bvhNode[idx] = node;
        }
    }
};

}  // namespace orc
