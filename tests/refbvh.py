"""Independent restatement (pure Python) of the reference's BVHBuilder — Accelearion/BVH.swift:78-250 with
AABB.area (Accelearion/AABB.swift:29-33): pre-sized node array, the root always tries to split, binned SAH over the
three axes (first strictly cheaper plane wins), in-place two-pointer partition by `centroid < splitPos` (which falls
back to axis 0, position 0 when no plane was found: the initial values of bestAxis / bestPos), children allocated as
nodesUsed + 1, + 2, depth-first recursion left then right.  Written from the Swift source, NOT from oracle/bvh.hpp or
csrc/bvh_build.cpp: the CPU tests hold both against this file.  Test infrastructure only."""
from __future__ import annotations

import math

INF = math.inf


class Builder:
    def __init__(self, bounds, centroids, max_leaf=2, bins=12):
        """bounds: n x 6 (min xyz, max xyz); centroids: n x 3 (python floats or numpy rows)."""
        self.b = [tuple(float(x) for x in r) for r in bounds]
        self.c = [tuple(float(x) for x in r) for r in centroids]
        n = len(self.b)
        self.prim_idx = list(range(n))
        self.max_leaf, self.bins = max_leaf, max(2, bins)
        count = max(1, n * 2 - 1)
        self.lo = [[INF] * 3 for _ in range(count)]
        self.hi = [[-INF] * 3 for _ in range(count)]
        self.left_first = [0] * count
        self.count = [0] * count
        self.count[0] = n
        self.used = 0
        self.update_bounds(0)
        self.subdivide(0)

    def update_bounds(self, k):                                       # :107-124
        lo, hi = [INF] * 3, [-INF] * 3
        for i in range(self.left_first[k], self.left_first[k] + self.count[k]):
            b = self.b[self.prim_idx[i]]
            for a in range(3):
                lo[a] = min(lo[a], b[a])
                hi[a] = max(hi[a], b[3 + a])
        self.lo[k], self.hi[k] = lo, hi

    @staticmethod
    def _area(lo, hi):                                                # AABB.swift:29-33
        ex, ey, ez = hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]
        return 2.0 * (ex * ey + ey * ez + ez * ex)

    def find_best_plane(self, k):                                     # :191-249
        best_cost, axis, pos = INF, 0, 0.0
        first, count, N = self.left_first[k], self.count[k], self.bins
        for a in range(3):
            bmin, bmax = INF, -INF
            for i in range(first, first + count):
                c = self.c[self.prim_idx[i]][a]
                bmin, bmax = min(bmin, c), max(bmax, c)
            if bmax <= bmin:
                continue
            blo = [[INF] * 3 for _ in range(N)]
            bhi = [[-INF] * 3 for _ in range(N)]
            cnt = [0] * N
            scale = float(N) / (bmax - bmin)
            for i in range(first, first + count):
                p = self.prim_idx[i]
                idx = min(N - 1, int((self.c[p][a] - bmin) * scale))
                cnt[idx] += 1
                for x in range(3):
                    blo[idx][x] = min(blo[idx][x], self.b[p][x])
                    bhi[idx][x] = max(bhi[idx][x], self.b[p][3 + x])
            left_area, right_area = [0.0] * (N - 1), [0.0] * (N - 1)
            left_cnt, right_cnt = [0] * (N - 1), [0] * (N - 1)
            Llo, Lhi, Rlo, Rhi = [INF] * 3, [-INF] * 3, [INF] * 3, [-INF] * 3
            sl = sr = 0
            for i in range(N - 1):
                sl += cnt[i]
                left_cnt[i] = sl
                for x in range(3):
                    Llo[x], Lhi[x] = min(Llo[x], blo[i][x]), max(Lhi[x], bhi[i][x])
                left_area[i] = self._area(Llo, Lhi) if sl else 0.0    # an empty side is skipped below; avoid inf * 0
                sr += cnt[N - 1 - i]
                right_cnt[N - 2 - i] = sr
                for x in range(3):
                    Rlo[x], Rhi[x] = min(Rlo[x], blo[N - 1 - i][x]), max(Rhi[x], bhi[N - 1 - i][x])
                right_area[N - 2 - i] = self._area(Rlo, Rhi) if sr else 0.0
            step = (bmax - bmin) / float(N)
            for i in range(N - 1):
                if left_cnt[i] == 0 or right_cnt[i] == 0:
                    continue
                cost = float(left_cnt[i]) * left_area[i] + float(right_cnt[i]) * right_area[i]
                if cost < best_cost:
                    best_cost, axis, pos = cost, a, bmin + step * float(i + 1)
        return axis, pos

    def subdivide(self, k):                                           # :127-188
        n = self.count[k]
        if n <= self.max_leaf and k != 0:
            return
        axis, pos = self.find_best_plane(k)
        first = self.left_first[k]
        i, j = first, first + n - 1
        while i <= j:
            if self.c[self.prim_idx[i]][axis] < pos:
                i += 1
            else:
                self.prim_idx[i], self.prim_idx[j] = self.prim_idx[j], self.prim_idx[i]
                j -= 1
        left = i - first
        if left == 0 or left == n:
            return
        self.used += 1
        lc = self.used
        self.used += 1
        rc = self.used
        assert rc < len(self.count), "BVH node overflow"
        self.left_first[lc], self.count[lc] = first, left
        self.left_first[rc], self.count[rc] = i, n - left
        self.left_first[k], self.count[k] = lc, 0
        self.update_bounds(lc)
        self.update_bounds(rc)
        self.subdivide(lc)
        self.subdivide(rc)


def hit_aabb(lo, hi, o, inv, eps):
    """hitAABB of the traversal loops (Models/RTContext.swift:561-569): entry distance, or inf for a miss."""
    t1 = [(lo[a] - o[a]) * inv[a] for a in range(3)]
    t2 = [(hi[a] - o[a]) * inv[a] for a in range(3)]
    tmin = max(max(min(t1[0], t2[0]), min(t1[1], t2[1])), min(t1[2], t2[2]))
    tmax = min(max(t1[0], t2[0]), min(max(t1[1], t2[1]), max(t1[2], t2[2])))
    return tmin if tmax >= max(tmin, eps) else INF


def closest_in_blas(B: Builder, o, d, eps, t_min, test):
    """intersectBLAS (:545-618): explicit stack, node re-tested when popped, near child pushed last, leaves walk
    primitiveIdx in order.  test(prim, best_t) -> t or None applies the primitive's own rules (t > max(eps, tMin),
    t < best_t: the FIRST primitive visited keeps an equal-t tie).  Returns (prim or -1, t, node visits)."""
    inv = [1.0 / x for x in d]
    best, best_t, visits = -1, INF, 0
    stack = [0]
    while stack:
        visits += 1
        k = stack.pop()
        if hit_aabb(B.lo[k], B.hi[k], o, inv, eps) == INF:
            continue
        if B.count[k] > 0:
            for i in range(B.left_first[k], B.left_first[k] + B.count[k]):
                p = B.prim_idx[i]
                t = test(p, best_t)
                if t is not None:
                    best, best_t = p, t
            continue
        L, R = B.left_first[k], B.left_first[k] + 1
        d1, d2 = hit_aabb(B.lo[L], B.hi[L], o, inv, eps), hit_aabb(B.lo[R], B.hi[R], o, inv, eps)
        if d1 > d2:
            d1, d2, L, R = d2, d1, R, L
        if d2 != INF:
            stack.append(R)
        if d1 != INF:
            stack.append(L)
    return best, best_t, visits
