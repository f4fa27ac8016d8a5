"""Pins the oracle's area-light path (Object+Extension.swift:145-186: per-chunk jitter index, table look-up, light
sample, shadow ray with tMax = dist - shadowEps, raw Phong exponent) against an independent restatement in
tests/refmath.py — no BVH, no shared code with oracle/."""
import math
import os

import numpy as np

import oracle_lib
import refmath
from myraytracer_b200 import scenegen


def _scene(tmp, W=36, H=20):
    verts = np.asarray([(-8, 0, -8), (8, 0, -8), (8, 0, 8), (-8, 0, 8), (0.4, 1.1, 0.3)], np.float64)
    sc = scenegen.Scene(maxRecursionDepth=0, backgroundColor=(3, 4, 5), ambient=(12, 11, 10), shadowRayEpsilon=1e-3,
                        intersectionTestEpsilon=1e-6)
    mat = dict(ambient=(0.5, 0.6, 0.7), diffuse=(0.7, 0.6, 0.5), specular=(0.4, 0.3, 0.2), phong=7.5)
    sc.materials = [scenegen.Material(id=1, ambient=mat["ambient"], diffuse=mat["diffuse"], specular=mat["specular"],
                                      phong=mat["phong"])]
    sc.vertexData = verts
    sc.objects = [scenegen.SceneObject(scenegen.OBJ_TRIANGLE, id=1, material=1, indices=(1, 3, 2)),
                  scenegen.SceneObject(scenegen.OBJ_TRIANGLE, id=2, material=1, indices=(1, 4, 3)),
                  scenegen.SceneObject(scenegen.OBJ_SPHERE, id=3, material=1, centerIdx=5, radius=1.1)]
    lights = [dict(position=(1.0, 6.0, 2.0), normal=(0.2, -1.0, 0.1), size=2.5, radiance=(900.0, 800.0, 700.0)),
              dict(position=(-5.0, 3.0, 4.0), normal=(1.0, -0.4, -0.6), size=1.5, radiance=(300.0, 500.0, 900.0))]
    sc.areaLights = [scenegen.AreaLight(l["position"], l["normal"], l["size"], l["radiance"]) for l in lights]
    cam = dict(width=W, height=H, numSamples=1, position=(0.5, 4.0, 10.0), gazePoint=(0, 0.5, 0), gaze=(0, 0, -1), up=(0, 1, 0),
               nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=45.0, isLookAt=True)
    sc.cameras = [scenegen.Camera(id=1, width=W, height=H, numSamples=1, position=cam["position"], gazePoint=cam["gazePoint"],
                                  up=cam["up"], nearDistance=1.0, fovy=45.0, isLookAt=True)]
    path = os.path.join(tmp, "area_ref.rtscene")
    scenegen.write_rtscene(sc, path)
    tris = [(verts[0], verts[2], verts[1]), (verts[0], verts[3], verts[2])]
    return path, cam, sc, mat, lights, tris, verts[4], 1.1


def _closest(o, d, tris, c, r, eps, t_min):
    best = (math.inf, None)
    for v0, v1, v2 in tris:
        h = refmath.mt_single(o, d, v0, v1, v2, eps)
        if h and h[0] > max(eps, t_min) and h[0] < best[0]:
            p = (v0 + h[1] * (v1 - v0)) + h[2] * (v2 - v0)
            best = (h[0], ("tri", p, refmath._norm(np.cross(v1 - v0, v2 - v0))))
    ts = refmath.brute_force_sphere(o, d, c, r, eps, t_min)
    if ts < best[0]:
        p = o + d * ts
        best = (ts, ("sphere", p, refmath._norm(p - c)))
    return best


def _occluded(o, d, t_max, tris, c, r, eps):
    for v0, v1, v2 in tris:
        h = refmath.mt_single(o, d, v0, v1, v2, eps)
        if h and eps < h[0] < t_max:
            return True
    oc = o - c                                                    # sphereShadowHit, RTContext.swift:851-861
    a, b, cc = refmath._dot(d, d), 2.0 * refmath._dot(oc, d), refmath._dot(oc, oc) - r * r
    disc = b * b - 4 * a * cc
    if disc < 0:
        return False
    sd = math.sqrt(disc)
    t = (-b - sd) / (2 * a)
    if t < eps:
        t = (-b + sd) / (2 * a)
    return eps < t < t_max


def test_oracle_area_lights_match_independent_restatement(tmp_path):
    path, cam, sc, mat, lights, tris, c, r = _scene(str(tmp_path))
    ref = oracle_lib.OracleScene(path).render(want_linear=True)
    xs, ys = refmath.jitter_table()
    eps, seps = sc.intersectionTestEpsilon, sc.shadowRayEpsilon
    W, H = cam["width"], cam["height"]
    want = np.zeros((H, W, 3))
    shadowed = lit = 0
    for row0 in range(0, H, 8):                                  # one jitter index per 8-row chunk task (:75-82,288)
        k = 0
        for j in range(row0, min(row0 + 8, H)):
            for i in range(W):
                (o, d, t_min), = refmath.primary_rays(cam, i, j)
                t, hit = _closest(o, d, tris, c, r, eps, t_min)
                if hit is None:
                    want[j, i] = sc.backgroundColor
                    continue
                _, p, n_geo = hit
                N = n_geo if refmath._dot(d, n_geo) < 0 else -n_geo
                Lo = np.asarray(sc.ambient, np.float64) * np.asarray(mat["ambient"], np.float64)
                for L in lights:
                    s = refmath.area_light_sample(L, k, xs, ys, p, N)
                    k += 1
                    if s is None:
                        continue
                    _, wi, dist, dist2, ndl, ln = s
                    if _occluded(p + wi * seps, wi, dist - seps, tris, c, r, eps):
                        shadowed += 1
                        continue
                    lit += 1
                    Lo = Lo + refmath.area_light_term(L, mat, N, d, wi, dist2, ndl, ln)
                want[j, i] = Lo
    assert lit > 200 and shadowed > 20                            # the scene exercises both outcomes
    got = ref["linear"]
    assert np.allclose(got, want, rtol=1e-12, atol=1e-12), float(np.abs(got - want).max())


def test_oracle_mesh_motion_blur_matches_brute_force(tmp_path):
    """Mesh.motionBlur: the triangle test sees the ray origin moved by -motionBlur * time (RTContext.swift:480-481)
    with time the third draw of the sample (:340); the BLAS boxes grown by the vector (:283-289) never hide a hit.
    Independent PCG32 / camera / no-BVH search: same primitive, bit-identical t, for every primary ray."""
    rng = np.random.default_rng(5)
    ntri = 120
    tri = rng.uniform(-3, 3, size=(ntri, 1, 3)) + rng.uniform(-0.7, 0.7, size=(ntri, 3, 3))
    mb = (0.9, -0.5, 0.4)
    sc = scenegen.Scene(maxRecursionDepth=0, backgroundColor=(1, 2, 3), ambient=(10, 10, 10))
    sc.vertexData = tri.reshape(-1, 3)
    sc.materials = [scenegen.Material(id=1)]
    faces = (np.arange(ntri * 3, dtype=np.int32) + 1).reshape(-1, 3)
    sc.objects = [scenegen.SceneObject(scenegen.OBJ_MESH, id=1, material=1, smooth=False, faces=faces, motionBlur=mb)]
    sc.pointLights = [scenegen.PointLight((0, 8, 8), (500, 500, 500))]
    W, H = 36, 28
    cam = dict(width=W, height=H, numSamples=4, position=(0.3, 0.4, 9.0), gazePoint=(0, 0, 0), gaze=(0, 0, -1), up=(0, 1, 0),
               nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=50.0, isLookAt=True)
    sc.cameras = [scenegen.Camera(id=1, width=W, height=H, numSamples=4, position=cam["position"], gazePoint=cam["gazePoint"],
                                  up=cam["up"], nearDistance=1.0, fovy=50.0, isLookAt=True)]
    path = os.path.join(str(tmp_path), "mesh_motion.rtscene")
    scenegen.write_rtscene(sc, path)
    ids = oracle_lib.OracleScene(path).render(want_ids=True)["ids"]
    still = 0
    for j in range(H):
        for i in range(W):
            for s, (o, d, t_min, time) in enumerate(refmath.primary_rays(cam, i, j, with_time=True)):
                k, t = refmath.brute_force_triangles(o - np.asarray(mb) * time, d, tri, 1e-6, t_min)
                h = ids[j, i, s]
                assert (int(h["prim"]), float(h["t"])) == ((k, t) if k >= 0 else (-1, math.inf)), (i, j, s)
                still += refmath.brute_force_triangles(o, d, tri, 1e-6, t_min)[0] == k
    assert still < 0.97 * W * H * 4                              # the motion changes which triangle is hit
