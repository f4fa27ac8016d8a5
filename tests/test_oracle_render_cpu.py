"""CPU tests of the oracle's scene-level behaviour: BVH builder invariants, PLY ingest through the reference's
own CPly sources, a brute-force (no BVH) differential test of the primary hits, golden-render drift."""
import ctypes as C
import math
import os
import struct

import numpy as np
import pytest

import oracle_lib
import refmath
from golden_cases import GOLDEN_CASES
from myraytracer_b200 import scenegen

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------------------------- BVHBuilder (BVH.swift:78-250)
@pytest.mark.parametrize("n", [1, 2, 3, 17, 500])
def test_bvh_builder_invariants(n):
    L = oracle_lib.lib()
    rng = np.random.default_rng(n)
    c = rng.uniform(-5, 5, size=(n, 3))
    h = rng.uniform(0.01, 0.3, size=(n, 3))
    bounds = np.concatenate([c - h, c + h], axis=1).copy()
    nb = np.zeros((2 * n, 6)); meta = np.zeros((2 * n, 2), np.int64); prim = np.zeros(n, np.int64)
    used = L.orc_bvh_build(n, _p(bounds), _p(c.copy()), _p(nb), _p(meta), _p(prim))
    assert 1 <= used <= 2 * n - 1 + (1 if n > 1 else 0)          # node 1 is skipped when the root splits (:90-91,169)
    assert sorted(prim.tolist()) == list(range(n))               # primIdx is a permutation
    seen = np.zeros(n, int)

    def walk(k, depth):
        lf, cnt = meta[k]
        if cnt > 0:                                              # leaf iff primitiveCount > 0 (BVH.swift:16-30)
            for q in prim[lf:lf + cnt]:
                seen[q] += 1
                assert (bounds[q, :3] >= nb[k, :3] - 1e-12).all() and (bounds[q, 3:] <= nb[k, 3:] + 1e-12).all()
            return cnt, depth
        a, da = walk(lf, depth + 1)                              # children adjacent: leftFirst, leftFirst + 1
        b, db = walk(lf + 1, depth + 1)
        for ch in (lf, lf + 1):
            assert (nb[ch, :3] >= nb[k, :3] - 1e-12).all() and (nb[ch, 3:] <= nb[k, 3:] + 1e-12).all()
        return a + b, max(da, db)

    total, depth = walk(0, 1)
    assert total == n and (seen == 1).all() and depth <= 64      # fixed 64-entry traversal stacks (RTContext.swift:550)
    leaves = meta[:used][meta[:used, 1] > 0]
    if n >= 17:
        assert (leaves[:, 1] <= 2).mean() > 0.9                  # maxLeaf 2 unless the split degenerates (:128-188)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 17, 200, 1500])
def test_bvh_builder_matches_independent_restatement(n):
    """oracle/bvh.hpp against tests/refbvh.py (pure Python, written from BVH.swift:78-250): node boxes, leftFirst /
    primitiveCount of every node and the primitiveIdx permutation identical — the topology the reference's traversal
    order, its tie-breaking and (with instance motion) its image depend on.  n = 5: coincident centroids, no split."""
    import refbvh
    L = oracle_lib.lib()
    rng = np.random.default_rng(n)
    c = rng.uniform(-5, 5, size=(n, 3))
    h = rng.uniform(0.01, 0.3, size=(n, 3))
    if n == 5:
        c[:] = c[0]
    bounds = np.concatenate([c - h, c + h], axis=1).copy()
    nb = np.zeros((2 * n, 6)); meta = np.zeros((2 * n, 2), np.int64); prim = np.zeros(n, np.int64)
    used = L.orc_bvh_build(n, _p(bounds), _p(c.copy()), _p(nb), _p(meta), _p(prim))
    B = refbvh.Builder(bounds, c)
    k = B.used + 1
    assert used == k
    assert prim.tolist() == B.prim_idx
    assert np.array_equal(meta[:k, 0], B.left_first[:k]) and np.array_equal(meta[:k, 1], B.count[:k])
    assert np.array_equal(nb[:k, :3], np.array(B.lo[:k])) and np.array_equal(nb[:k, 3:], np.array(B.hi[:k]))


# ---------------------------------------------------------------------------------- PLY ingest
def _write_quad_ply(path):
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\n"
                b"element face 2\nproperty list uchar int vertex_indices\nend_header\n")
        for v in [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0.5, 2, 0)]:
            f.write(struct.pack("<fff", *v))
        f.write(struct.pack("<Biiii", 4, 0, 1, 2, 3))
        f.write(struct.pack("<Biii", 3, 3, 2, 4))


def _ply_load(path, use_ref):
    L = oracle_lib.lib()
    nv, ni, nn = C.c_int64(), C.c_int64(), C.c_int64()
    cply = oracle_lib.CPLY_SO.encode() if use_ref else None
    rc = L.orc_ply_load(path.encode(), 1 if use_ref else 0, cply, C.byref(nv), C.byref(ni), C.byref(nn), None, None, None)
    assert rc == 0, rc
    pos = np.zeros((nv.value, 3)); idx = np.zeros(ni.value, np.int64); nrm = np.zeros((nn.value, 3))
    L.orc_ply_load(path.encode(), 1 if use_ref else 0, cply, C.byref(nv), C.byref(ni), C.byref(nn), _p(pos), _p(idx), _p(nrm))
    return pos, idx


def test_ply_quad_triangulation_order(tmp_path):
    """miniply splits a quad 0,1,3 / 2,3,1 (miniply.cpp:1976-1985); the Swift loader takes that path when a face
    is not a triangle (Helpers/PLYReader.swift:186-198)."""
    path = str(tmp_path / "quad.ply")
    _write_quad_ply(path)
    have_ref = os.path.exists(oracle_lib.CPLY_SO)
    if have_ref:
        pos, idx = _ply_load(path, True)                         # the reference's own CPly, compiled from its sources
        assert idx.tolist() == [0, 1, 3, 2, 3, 1, 3, 2, 4]
        assert pos.shape == (5, 3) and pos[4].tolist() == [0.5, 2.0, 0.0]
    # the built-in reader (used only when oracle/_ref is absent) reads all-triangle files and refuses others
    L = oracle_lib.lib()
    nv, ni, nn = C.c_int64(), C.c_int64(), C.c_int64()
    assert L.orc_ply_load(path.encode(), 0, None, C.byref(nv), C.byref(ni), C.byref(nn), None, None, None) == -1


def test_ply_positions_are_float32_exact(tmp_path):
    pos, tris = scenegen.terrain_mesh(12, 9)
    path = str(tmp_path / "t.ply")
    scenegen.write_ply(path, pos, tris)
    for use_ref in ([True, False] if os.path.exists(oracle_lib.CPLY_SO) else [False]):
        got, idx = _ply_load(path, use_ref)
        assert np.array_equal(got, pos.astype(np.float32).astype(np.float64))     # PLYReader.swift:85-101: Float -> Double
        assert np.array_equal(idx.reshape(-1, 3), tris)


# ---------------------------------------------------------------------------------- brute-force differential
def _random_scene(tmp, seed, ntri=150, samples=4, W=40, H=30, aperture=0.0, focus=0.0, lookat=True):
    rng = np.random.default_rng(seed)
    base = rng.uniform(-3, 3, size=(ntri, 1, 3))
    tri = base + rng.uniform(-0.7, 0.7, size=(ntri, 3, 3))
    verts = np.concatenate([tri.reshape(-1, 3), [[0.5, 0.2, 1.0]]])
    faces = (np.arange(ntri * 3, dtype=np.int32) + 1).reshape(-1, 3)
    sc = scenegen.Scene(maxRecursionDepth=0, backgroundColor=(1, 2, 3), ambient=(10, 10, 10))
    sc.vertexData = verts
    sc.materials = [scenegen.Material(id=1)]
    sc.objects = [scenegen.SceneObject(scenegen.OBJ_MESH, id=1, material=1, smooth=False, faces=faces),
                  scenegen.SceneObject(scenegen.OBJ_SPHERE, id=2, material=1, centerIdx=ntri * 3 + 1, radius=0.8)]
    sc.pointLights = [scenegen.PointLight((0, 8, 8), (500, 500, 500))]
    cam = dict(width=W, height=H, numSamples=samples, position=(0.3, 0.4, 9.0), gazePoint=(0, 0, 0), gaze=(0, 0, -1),
               up=(0, 1, 0), nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=50.0, isLookAt=True, apertureSize=aperture,
               focusDistance=focus)
    if not lookat:                       # "simple" camera: gaze vector + explicit near plane (:404-427), up not orthogonal
        cam.update(isLookAt=False, fovy=None, gaze=(-0.05, -0.04, -1.0), up=(0.1, 1.0, -0.2), nearPlane=(-0.55, 0.5, -0.4, 0.42),
                   nearDistance=1.2)
    sc.cameras = [scenegen.Camera(id=1, width=W, height=H, numSamples=samples, position=cam["position"],
                                  gazePoint=cam["gazePoint"], gaze=cam["gaze"], up=cam["up"], nearPlane=cam["nearPlane"],
                                  nearDistance=cam["nearDistance"], fovy=cam["fovy"], isLookAt=cam["isLookAt"],
                                  apertureSize=aperture, focusDistance=focus)]
    path = os.path.join(tmp, f"rand{seed}_{aperture}.rtscene")
    scenegen.write_rtscene(sc, path)
    return path, cam, tri, verts[-1], 0.8


@pytest.mark.parametrize("seed,aperture,focus,lookat", [(11, 0.0, 0.0, True), (12, 0.0, 0.0, True), (13, 0.45, 8.5, True),
                                                        (14, 0.0, 0.0, False), (15, 0.3, 9.0, False)])
def test_primary_hits_match_brute_force(tmp_path, seed, aperture, focus, lookat):
    """Oracle (reference BVH + reference traversal order) vs an independent no-BVH search with independently
    written PCG32 / camera code: same primitive and bit-identical t for every primary ray — also through the thin
    lens (Object+Extension.swift:325-338: two more draws per sample, origin on the lens, image-plane tMin from there)."""
    path, cam, tri, sc_c, sc_r = _random_scene(str(tmp_path), seed, aperture=aperture, focus=focus, lookat=lookat)
    r = oracle_lib.OracleScene(path).render(want_ids=True)
    ids = r["ids"]
    eps = 1e-6
    mism = 0
    for j in range(cam["height"]):
        for i in range(cam["width"]):
            for s, (o, d, tmin) in enumerate(refmath.primary_rays(cam, i, j)):
                k, t = refmath.brute_force_triangles(o, d, tri, eps, tmin)
                ts = refmath.brute_force_sphere(o, d, sc_c, sc_r, eps, tmin)
                h = ids[j, i, s]
                if ts < t:
                    want = (2, 0, ts)
                elif k >= 0:
                    want = (1, k, t)
                else:
                    want = (0, -1, math.inf)
                if (int(h["kind"]), int(h["prim"])) != want[:2] or float(h["t"]) != want[2]:
                    mism += 1
    assert mism == 0


# ---------------------------------------------------------------------------------- golden drift
@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_oracle_matches_golden(name, scene_dir):
    fn, kw = GOLDEN_CASES[name]
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    r = oracle_lib.OracleScene(fn(scene_dir, **kw)).render(want_ids=True)
    assert np.array_equal(r["rgba"], g["rgba"])
    assert np.array_equal(r["ids"]["prim"], g["prim"]) and np.array_equal(r["ids"]["kind"], g["kind"])
    assert np.array_equal(r["ids"]["t"], g["t"])
    assert [r["counters"][k] for k in ("raysPrimary", "raysShadow", "raysSecondary")] == g["rays"].tolist()


def test_oracle_chunk_sampling_is_consistent(scene_dir):
    """chunk_stride/offset (the bounded CPU-baseline sample of bench.py) renders exactly the rows it claims."""
    path = scenegen.config_c1(scene_dir, width=64, height=50)
    o = oracle_lib.OracleScene(path)
    full = o.render()
    part = o.render(chunk_stride=3, chunk_offset=1)
    rows = np.zeros(50, bool)
    for c in range(1, 7, 3):
        rows[c * 8:(c + 1) * 8] = True
    assert np.array_equal(part["rgba"][rows], full["rgba"][rows])
    assert part["counters"]["raysPrimary"] == int(rows.sum()) * 64


# ---------------------------------------------------------------------------------- traversal order and ties
def test_tie_breaking_follows_the_reference_traversal_order(tmp_path):
    """Hits are committed on strict t < hit.t (Models/RTContext.swift:494), so among primitives at exactly the same
    distance the first one VISITED wins, and the visiting order is the reference's: its BVH (BVH.swift:78-250), its
    near-first rule and stack discipline (:590-616), primitiveIdx order inside a leaf.  A mesh in which every triangle
    exists three times makes every hit a three-way tie; the oracle must report the triangle that a pure-Python walk
    of tests/refbvh.py's tree reports — for every primary ray."""
    import refbvh
    rng = np.random.default_rng(21)
    nt = 40
    base = rng.uniform(-3, 3, size=(nt, 1, 3)) + rng.uniform(-0.9, 0.9, size=(nt, 3, 3))
    tri = np.concatenate([base, base, base])[rng.permutation(3 * nt)]          # 120 triangles, each position 3 times
    verts = tri.reshape(-1, 3)
    faces = (np.arange(len(verts), dtype=np.int32) + 1).reshape(-1, 3)
    W, H, samples = 36, 28, 1
    sc = scenegen.Scene(maxRecursionDepth=0, backgroundColor=(1, 2, 3), ambient=(10, 10, 10))
    sc.vertexData = verts
    sc.materials = [scenegen.Material(id=1)]
    sc.objects = [scenegen.SceneObject(scenegen.OBJ_MESH, id=1, material=1, smooth=False, faces=faces)]
    sc.pointLights = [scenegen.PointLight((0, 8, 8), (500, 500, 500))]
    cam = dict(width=W, height=H, numSamples=samples, position=(0.2, 0.3, 9.0), gazePoint=(0, 0, 0), gaze=(0, 0, -1),
               up=(0, 1, 0), nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=50.0, isLookAt=True)
    sc.cameras = [scenegen.Camera(id=1, width=W, height=H, numSamples=samples, position=cam["position"],
                                  gazePoint=cam["gazePoint"], up=cam["up"], nearDistance=1.0, fovy=50.0, isLookAt=True)]
    path = os.path.join(str(tmp_path), "ties.rtscene")
    scenegen.write_rtscene(sc, path)
    ids = oracle_lib.OracleScene(path).render(want_ids=True)["ids"]

    bounds = np.concatenate([tri.min(axis=1), tri.max(axis=1)], axis=1)
    cent = ((tri[:, 0] + tri[:, 1]) + tri[:, 2]) / 3                            # tri.centroid, RTContext.swift:339
    B = refbvh.Builder(bounds, cent)
    eps = 1e-6
    hits = ties = 0
    for j in range(H):
        for i in range(W):
            (o, d, tmin), = refmath.primary_rays(cam, i, j)

            def test(p, best_t):
                h = refmath.mt_single(o, d, tri[p, 0], tri[p, 1], tri[p, 2], eps)
                return h[0] if h and h[0] > max(eps, tmin) and h[0] < best_t else None

            prim, t, _ = refbvh.closest_in_blas(B, list(o), list(d), eps, tmin, test)
            r = ids[j, i, 0]
            if prim < 0:
                assert int(r["kind"]) == 0
                continue
            hits += 1
            assert int(r["kind"]) == 1 and float(r["t"]) == t
            assert int(r["prim"]) == prim, (i, j, int(r["prim"]), prim)
            k, tb = refmath.brute_force_triangles(o, d, tri, eps, tmin)
            ties += int(tb == t and k != prim)
    assert hits > 100 and ties > 20, (hits, ties)       # the winner is often not the lowest-numbered of the three
