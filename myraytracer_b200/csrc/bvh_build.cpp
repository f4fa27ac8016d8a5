// bvh_build.cpp — see bvh_build.h.
#include "bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <future>
#include <limits>
#include <thread>

namespace rtg {
namespace {

constexpr int kBins = 16;
constexpr double kCostTri = 1.2;   // FP64 Moeller-Trumbore vs one FP32 two-box node test
constexpr double kCostNode = 1.0;
constexpr double kInfD = std::numeric_limits<double>::infinity();

struct Box {
    double lo[3] = {kInfD, kInfD, kInfD};
    double hi[3] = {-kInfD, -kInfD, -kInfD};
    void grow(const Box& b) {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); }
    }
    void grow(const double* p) {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); }
    }
    double area() const {
        double ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        if (ex < 0 || ey < 0 || ez < 0) return 0;
        return 2.0 * (ex * ey + ey * ez + ez * ex);
    }
};

struct Ref {
    Box b;
    double c[3];
    int64_t tri;
};

struct BNode {
    Box b;
    int32_t left = -1, right = -1;  // BNode indices, -1 = leaf
    int64_t first = 0, count = 0;   // leaf range in refs
};

struct Builder {
    std::vector<Ref>& refs;
    std::vector<BNode> pool;
    std::atomic<int64_t> used{0};
    std::atomic<int> maxDepth{0};
    std::atomic<int> spawned{0};
    int maxThreads;

    Builder(std::vector<Ref>& r, int threads) : refs(r), pool(std::max<size_t>(1, 2 * r.size())), maxThreads(threads) {}

    int32_t alloc() { return (int32_t)used.fetch_add(1); }

    int32_t build(int64_t first, int64_t count, int depth) {
        int32_t id = alloc();
        BNode& n = pool[id];
        Box cb;
        for (int64_t i = first; i < first + count; i++) { n.b.grow(refs[i].b); cb.grow(refs[i].c); }
        int d = maxDepth.load();
        while (depth > d && !maxDepth.compare_exchange_weak(d, depth)) {}

        auto makeLeaf = [&]() { n.first = first; n.count = count; return id; };
        if (count == 1) return makeLeaf();

        const double parentArea = n.b.area();
        double bestCost = kInfD;
        int bestAxis = -1, bestBin = -1;
        for (int a = 0; a < 3; a++) {
            const double cmin = cb.lo[a], cmax = cb.hi[a];
            if (!(cmax > cmin)) continue;
            Box bb[kBins];
            int64_t bc[kBins] = {0};
            const double scale = kBins / (cmax - cmin);
            for (int64_t i = first; i < first + count; i++) {
                int k = std::min(kBins - 1, std::max(0, (int)((refs[i].c[a] - cmin) * scale)));
                bc[k]++;
                bb[k].grow(refs[i].b);
            }
            double rightArea[kBins];
            int64_t rightCnt[kBins];
            Box r;
            int64_t rc = 0;
            for (int k = kBins - 1; k >= 1; k--) { r.grow(bb[k]); rc += bc[k]; rightArea[k] = r.area(); rightCnt[k] = rc; }
            Box l;
            int64_t lc = 0;
            for (int k = 0; k < kBins - 1; k++) {
                l.grow(bb[k]); lc += bc[k];
                if (lc == 0 || rightCnt[k + 1] == 0) continue;
                double cost = l.area() * (double)lc + rightArea[k + 1] * (double)rightCnt[k + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = k; }
            }
        }

        int64_t mid = -1;
        if (bestAxis >= 0) {
            const double leafCost = kCostTri * (double)count * parentArea;
            const double splitCost = kCostNode * parentArea + kCostTri * bestCost;
            if (count <= kMaxLeaf && leafCost <= splitCost) return makeLeaf();
            const double cmin = cb.lo[bestAxis], scale = kBins / (cb.hi[bestAxis] - cmin);
            auto it = std::partition(refs.begin() + first, refs.begin() + first + count, [&](const Ref& r) {
                int k = std::min(kBins - 1, std::max(0, (int)((r.c[bestAxis] - cmin) * scale)));
                return k <= bestBin;
            });
            mid = it - refs.begin();
            if (mid == first || mid == first + count) mid = -1;
        }
        if (mid < 0) {
            if (count <= kMaxLeaf) return makeLeaf();
            mid = first + count / 2;  // coincident centroids: split the list in half
        }

        const int64_t lcount = mid - first, rcount = first + count - mid;
        int32_t L, R;
        if (count > 65536 && spawned.load() < maxThreads) {
            spawned.fetch_add(1);
            auto fut = std::async(std::launch::async, [&, first, lcount, depth]() { return build(first, lcount, depth + 1); });
            R = build(mid, rcount, depth + 1);
            L = fut.get();
            spawned.fetch_sub(1);
        } else {
            L = build(first, lcount, depth + 1);
            R = build(mid, rcount, depth + 1);
        }
        BNode& n2 = pool[id];
        n2.left = L; n2.right = R;
        return id;
    }
};

inline float roundDown(double x) {
    float f = (float)x;
    if ((double)f > x) f = std::nextafterf(f, -std::numeric_limits<float>::infinity());
    return f;
}
inline float roundUp(double x) {
    float f = (float)x;
    if ((double)f < x) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
    return f;
}

}  // namespace

void buildBlas(const BlasBuildInput& in, BlasBuildOutput& out) {
    out = BlasBuildOutput();
    const int64_t n = in.count;
    std::vector<Ref> refs((size_t)n);
    bool f64 = false;
    const bool blur = in.motionBlur[0] != 0 || in.motionBlur[1] != 0 || in.motionBlur[2] != 0;
    Box root;
    for (int64_t i = 0; i < n; i++) {
        const double* v = in.triVertices + in.triIndices[i] * 9;
        Ref& r = refs[i];
        r.tri = in.triIndices[i];
        for (int k = 0; k < 3; k++) {
            r.b.grow(v + 3 * k);
            if (blur) { double p[3] = {v[3 * k] + in.motionBlur[0], v[3 * k + 1] + in.motionBlur[1], v[3 * k + 2] + in.motionBlur[2]}; r.b.grow(p); }
        }
        for (int a = 0; a < 3; a++) r.c[a] = 0.5 * (r.b.lo[a] + r.b.hi[a]);
        for (int k = 0; k < 9; k++)
            if ((double)(float)v[k] != v[k]) f64 = true;
        root.grow(r.b);
    }
    out.f64 = f64;
    for (int a = 0; a < 3; a++) { out.rootLo[a] = root.lo[a]; out.rootHi[a] = root.hi[a]; out.centre[a] = 0.5 * (root.lo[a] + root.hi[a]); }
    if (n == 0) return;

    int threads = in.threads > 0 ? in.threads : (int)std::thread::hardware_concurrency();
    Builder B(refs, std::max(1, threads) - 1);
    int32_t rootId = B.build(0, n, 1);
    out.maxDepth = B.maxDepth.load();

    // conservative FP32 boxes: relative to centre, padded by M * 2^-19 (see DESIGN.md "FP32 slab test")
    double M = 0;
    for (int a = 0; a < 3; a++) M = std::max(M, std::max(std::fabs(root.lo[a] - out.centre[a]), std::fabs(root.hi[a] - out.centre[a])));
    const double P = std::max(M * std::ldexp(1.0, -19), 1e-30);
    out.pad = P;

    // leaf-order triangles
    if (f64) out.trisD.resize((size_t)n); else out.tris.resize((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        const double* v = in.triVertices + refs[i].tri * 9;
        if (f64) {
            GTriD& t = out.trisD[i];
            std::memcpy(t.v0, v, 24); std::memcpy(t.v1, v + 3, 24); std::memcpy(t.v2, v + 6, 24);
            t.triIndex = refs[i].tri;
        } else {
            GTri& t = out.tris[i];
            for (int k = 0; k < 3; k++) { t.v0[k] = (float)v[k]; t.v1[k] = (float)v[3 + k]; t.v2[k] = (float)v[6 + k]; }
            t.triIndex = (int32_t)refs[i].tri; t.pad1 = 0; t.pad2 = 0;
        }
    }

    auto leafRef = [&](const BNode& b) -> int32_t {
        if (b.count > out.maxLeaf) out.maxLeaf = (int)b.count;
        return ~(int32_t)((b.first << 3) | (b.count - 1));
    };
    const BNode& rootN = B.pool[rootId];
    if (rootN.left < 0) {  // whole BLAS is one leaf
        out.rootRef = leafRef(rootN);
        return;
    }
    // DFS flatten; inner BNode -> GNode index assigned in pre-order
    struct Item { int32_t bnode; int32_t gnode; };
    std::vector<Item> stack;
    out.nodes.reserve((size_t)B.used.load() / 2 + 1);
    out.nodes.emplace_back();
    stack.push_back({rootId, 0});
    out.rootRef = 0;
    auto putBox = [&](const Box& b, float* lo, float* hi) {
        for (int a = 0; a < 3; a++) { lo[a] = roundDown(b.lo[a] - out.centre[a] - P); hi[a] = roundUp(b.hi[a] - out.centre[a] + P); }
    };
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        const BNode& bn = B.pool[it.bnode];
        const BNode& l = B.pool[bn.left];
        const BNode& r = B.pool[bn.right];
        float lo0[3], hi0[3], lo1[3], hi1[3];
        putBox(l.b, lo0, hi0);
        putBox(r.b, lo1, hi1);
        int32_t c0, c1;
        if (l.left < 0) c0 = leafRef(l); else { c0 = (int32_t)out.nodes.size(); out.nodes.emplace_back(); }
        if (r.left < 0) c1 = leafRef(r); else { c1 = (int32_t)out.nodes.size(); out.nodes.emplace_back(); }
        GNode& g = out.nodes[it.gnode];
        g.q0[0] = lo0[0]; g.q0[1] = lo0[1]; g.q0[2] = lo0[2]; g.q0[3] = hi0[0];
        g.q1[0] = hi0[1]; g.q1[1] = hi0[2]; g.q1[2] = lo1[0]; g.q1[3] = lo1[1];
        g.q2[0] = lo1[2]; g.q2[1] = hi1[0]; g.q2[2] = hi1[1]; g.q2[3] = hi1[2];
        g.child0 = c0; g.child1 = c1; g.pad0 = 0; g.pad1 = 0;
        if (r.left >= 0) stack.push_back({bn.right, c1});
        if (l.left >= 0) stack.push_back({bn.left, c0});
    }
}

void collapseWide(const std::vector<GNode>& nodes, int32_t rootRef, WideBuildOutput& out) {
    out = WideBuildOutput();
    if (rootRef < 0 || rootRef == kEmptyRef || nodes.empty()) return;
    struct Child { float lo[3], hi[3]; int32_t ref; double area; };
    auto children = [&](int32_t n, Child& a, Child& b) {
        const GNode& g = nodes[(size_t)n];
        a.lo[0] = g.q0[0]; a.lo[1] = g.q0[1]; a.lo[2] = g.q0[2]; a.hi[0] = g.q0[3]; a.hi[1] = g.q1[0]; a.hi[2] = g.q1[1];
        b.lo[0] = g.q1[2]; b.lo[1] = g.q1[3]; b.lo[2] = g.q2[0]; b.hi[0] = g.q2[1]; b.hi[1] = g.q2[2]; b.hi[2] = g.q2[3];
        a.ref = g.child0; b.ref = g.child1;
        for (Child* c : {&a, &b}) {
            const double ex = (double)c->hi[0] - c->lo[0], ey = (double)c->hi[1] - c->lo[1], ez = (double)c->hi[2] - c->lo[2];
            c->area = ex * ey + ey * ez + ez * ex;
        }
    };
    struct Item { int32_t bnode, wnode; int depth; };
    std::vector<Item> todo;
    out.nodes.emplace_back();
    out.rootRef = 0;
    todo.push_back({rootRef, 0, 1});
    std::vector<Child> list;
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        if (it.depth > out.depth) out.depth = it.depth;
        list.assign(2, Child());
        children(it.bnode, list[0], list[1]);
        for (size_t k = 0; k < list.size();)              // a binary node with one child only (kEmptyRef): drop the slot
            if (list[k].ref == kEmptyRef) list.erase(list.begin() + (long)k); else k++;
        while ((int)list.size() < kWide) {
            int best = -1;
            for (int k = 0; k < (int)list.size(); k++)
                if (list[k].ref >= 0 && (best < 0 || list[k].area > list[best].area)) best = k;
            if (best < 0) break;
            Child a, b;
            children(list[best].ref, a, b);
            list[best] = a;
            list.push_back(b);
            for (size_t k = 0; k < list.size();)
                if (list[k].ref == kEmptyRef) list.erase(list.begin() + (long)k); else k++;
        }
        // front-to-back order: sort by the box centre along the axis on which the centres spread the most
        const int n = (int)list.size();
        int axis = 0;
        {
            double best = -1;
            for (int a = 0; a < 3; a++) {
                double cmin = kInfD, cmax = -kInfD;
                for (int k = 0; k < n; k++) { const double c = (double)list[k].lo[a] + list[k].hi[a]; cmin = std::min(cmin, c); cmax = std::max(cmax, c); }
                if (cmax - cmin > best) { best = cmax - cmin; axis = a; }
            }
            std::stable_sort(list.begin(), list.end(), [&](const Child& x, const Child& y) {
                return (double)x.lo[axis] + x.hi[axis] < (double)y.lo[axis] + y.hi[axis];
            });
        }
        WNode w;
        std::memset(&w, 0, sizeof w);
        for (int k = 0; k < kWide; k++) {
            WChild& c = w.c[k];
            if (k < n) {
                c.lo[0] = list[k].lo[0]; c.lo[1] = list[k].lo[1]; c.lo[2] = list[k].lo[2];
                c.hi0 = list[k].hi[0]; c.hi1 = list[k].hi[1]; c.hi2 = list[k].hi[2];
                if (list[k].ref >= 0) {
                    c.ref = (int32_t)out.nodes.size();
                    out.nodes.emplace_back();
                    todo.push_back({list[k].ref, c.ref, it.depth + 1});
                } else {
                    c.ref = list[k].ref;
                }
            } else {        // empty slot: an inverted box fails the beam test
                c.lo[0] = c.lo[1] = c.lo[2] = 1e30f; c.hi0 = c.hi1 = c.hi2 = -1e30f;
                c.ref = kEmptyRef;
            }
        }
        w.c[0].aux = axis;
        out.nodes[(size_t)it.wnode] = w;
    }
}

void buildTlas(const double* wb, int64_t n, TlasBuildOutput& out) {
    out = TlasBuildOutput();
    if (n <= 0) return;
    if (n == 1) { out.rootRef = ~0; return; }
    std::vector<int32_t> idx((size_t)n);
    for (int64_t i = 0; i < n; i++) idx[i] = (int32_t)i;
    auto boxOf = [&](int64_t first, int64_t count, double* lo, double* hi) {
        for (int a = 0; a < 3; a++) { lo[a] = kInfD; hi[a] = -kInfD; }
        for (int64_t i = first; i < first + count; i++)
            for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], wb[idx[i] * 6 + a]); hi[a] = std::max(hi[a], wb[idx[i] * 6 + 3 + a]); }
    };
    // recursive median split on the widest centroid axis; returns child reference
    struct Rec {
        std::vector<int32_t>& idx; const double* wb; TlasBuildOutput& out; decltype(boxOf)& boxFn;
        int32_t go(int64_t first, int64_t count, int depth) {
            if (depth > out.maxDepth) out.maxDepth = depth;
            if (count == 1) return ~idx[first];
            double cl[3] = {kInfD, kInfD, kInfD}, ch[3] = {-kInfD, -kInfD, -kInfD};
            for (int64_t i = first; i < first + count; i++)
                for (int a = 0; a < 3; a++) {
                    double c = 0.5 * (wb[idx[i] * 6 + a] + wb[idx[i] * 6 + 3 + a]);
                    if (!(c == c)) c = 0;
                    cl[a] = std::min(cl[a], c); ch[a] = std::max(ch[a], c);
                }
            int axis = 0;
            for (int a = 1; a < 3; a++) if (ch[a] - cl[a] > ch[axis] - cl[axis]) axis = a;
            int64_t mid = first + count / 2;
            std::nth_element(idx.begin() + first, idx.begin() + mid, idx.begin() + first + count, [&](int32_t x, int32_t y) {
                return wb[x * 6 + axis] + wb[x * 6 + 3 + axis] < wb[y * 6 + axis] + wb[y * 6 + 3 + axis];
            });
            int32_t me = (int32_t)out.nodes.size();
            out.nodes.emplace_back();
            int32_t c0 = go(first, mid - first, depth + 1);
            int32_t c1 = go(mid, first + count - mid, depth + 1);
            TNode& t = out.nodes[me];
            boxFn(first, mid - first, t.lo0, t.hi0);
            boxFn(mid, first + count - mid, t.lo1, t.hi1);
            t.child0 = c0; t.child1 = c1; t.pad0 = t.pad1 = 0;
            return me;
        }
    } rec{idx, wb, out, boxOf};
    out.rootRef = rec.go(0, n, 1);
}

// ---- the reference's TLAS, rebuilt exactly (scenes with instanceMotion; see bvh_build.h) ---------------------------
namespace {
struct RefBox {                          // AABB.swift:12-35,62-70
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    void grow(const double* l, const double* h) {
        for (int a = 0; a < 3; a++) { lo[a] = l[a] < lo[a] ? l[a] : lo[a]; hi[a] = h[a] > hi[a] ? h[a] : hi[a]; }
    }
    double area() const {
        const double ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return 2.0 * ((ex * ey + ey * ez) + ez * ex);
    }
};
struct RefBuilder {
    const double* wb; int64_t n;
    std::vector<double> centroid;        // (minP + maxP) * 0.5 per instance (RTContext.swift:464)
    std::vector<RefNode>& nodes; std::vector<int32_t>& idx;
    int32_t used = 0;
    static const int N = 12, maxLeaf = 2;

    void updateBounds(int32_t ni) {                                        // BVH.swift:107-125
        RefNode& nd = nodes[ni];
        RefBox b;
        for (int i = 0; i < nd.count; i++) { const int64_t p = idx[nd.leftFirst + i]; b.grow(wb + p * 6, wb + p * 6 + 3); }
        for (int a = 0; a < 3; a++) { nd.lo[a] = b.lo[a]; nd.hi[a] = b.hi[a]; }
    }
    double bestSplit(const RefNode& nd, int& axis, double& pos) const {    // BVH.swift:189-250
        double bestCost = INFINITY;
        for (int a = 0; a < 3; a++) {
            double bmin = INFINITY, bmax = -INFINITY;
            for (int i = 0; i < nd.count; i++) {
                const double c = centroid[(size_t)idx[nd.leftFirst + i] * 3 + a];
                bmin = c < bmin ? c : bmin; bmax = c > bmax ? c : bmax;
            }
            if (bmax <= bmin) continue;
            RefBox bin[N]; int cnt[N] = {};
            const double scale = (double)N / (bmax - bmin);
            for (int i = 0; i < nd.count; i++) {
                const int64_t p = idx[nd.leftFirst + i];
                int k = (int)((centroid[(size_t)p * 3 + a] - bmin) * scale);
                if (k > N - 1) k = N - 1;
                cnt[k]++;
                bin[k].grow(wb + p * 6, wb + p * 6 + 3);
            }
            double leftArea[N - 1], rightArea[N - 1]; int leftCnt[N - 1], rightCnt[N - 1];
            RefBox L, R; int sL = 0, sR = 0;
            for (int i = 0; i < N - 1; i++) {
                sL += cnt[i]; leftCnt[i] = sL; L.grow(bin[i].lo, bin[i].hi); leftArea[i] = L.area();
                sR += cnt[N - 1 - i]; rightCnt[N - 2 - i] = sR; R.grow(bin[N - 1 - i].lo, bin[N - 1 - i].hi); rightArea[N - 2 - i] = R.area();
            }
            const double step = (bmax - bmin) / (double)N;
            for (int i = 0; i < N - 1; i++) {
                if (leftCnt[i] == 0 || rightCnt[i] == 0) continue;
                const double cost = (double)leftCnt[i] * leftArea[i] + (double)rightCnt[i] * rightArea[i];
                if (cost < bestCost) { bestCost = cost; axis = a; pos = bmin + step * (double)(i + 1); }
            }
        }
        return bestCost;
    }
    void subdivide(int32_t ni) {                                           // BVH.swift:128-186
        const RefNode nd = nodes[ni];
        const int count = nd.count;
        if (count <= maxLeaf && ni != 0) return;
        int axis = 0; double pos = 0;
        bestSplit(nd, axis, pos);
        int i = nd.leftFirst, j = i + count - 1;
        while (i <= j) {
            if (centroid[(size_t)idx[i] * 3 + axis] < pos) i++;
            else { std::swap(idx[i], idx[j]); j--; }
        }
        const int leftCount = i - nd.leftFirst;
        if (leftCount == 0 || leftCount == count) return;
        const int32_t l = ++used, r = ++used;
        nodes[l].leftFirst = nd.leftFirst; nodes[l].count = leftCount;
        nodes[r].leftFirst = i; nodes[r].count = count - leftCount;
        nodes[ni].leftFirst = l; nodes[ni].count = 0;
        updateBounds(l); updateBounds(r);
        subdivide(l); subdivide(r);
    }
};
}  // namespace

void buildReferenceTlas(const double* worldBounds, int64_t n, RefTlasOutput& out) {
    out.nodes.assign((size_t)std::max<int64_t>(1, 2 * n - 1), RefNode{});      // BVH.swift:90-91
    out.primIdx.resize((size_t)n);
    for (int64_t i = 0; i < n; i++) out.primIdx[i] = (int32_t)i;
    RefBuilder b{worldBounds, n, {}, out.nodes, out.primIdx};
    b.centroid.resize((size_t)n * 3);
    for (int64_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) b.centroid[i * 3 + a] = (worldBounds[i * 6 + a] + worldBounds[i * 6 + 3 + a]) * 0.5;
    out.nodes[0].leftFirst = 0; out.nodes[0].count = (int32_t)n;
    b.updateBounds(0);
    b.subdivide(0);
    out.nodes.resize((size_t)b.used + 1);
}

}  // namespace rtg
