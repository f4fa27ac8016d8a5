"""Pins the oracle's shading recursion (Extensions/Object+Extension.swift:96-283) against the independent pure-Python
restatement in tests/reftrace.py: no BVH, no code shared with oracle/.  Covers what BASELINE's C3 exercises — ambient,
point lights with shadow rays, mirror, dielectric (Fresnel, total internal reflection, Beer's law, light only on the
outside), conductor, the depth rules — plus glossy roughness drawing from the pixel's PCG32 stream inside the
recursion (reflection subtree before refraction subtree)."""
import os

import numpy as np
import pytest

import oracle_lib
import refmath
import reftrace
from myraytracer_b200 import scenegen

TYPES = {scenegen.MAT_PLAIN: "plain", scenegen.MAT_MIRROR: "mirror", scenegen.MAT_DIELECTRIC: "dielectric",
         scenegen.MAT_CONDUCTOR: "conductor"}


def _build(tmp, rough, depth, W=30, H=20, samples=4, inf_light=False):
    verts = np.asarray([(0, -1.0, 0), (-7, -1, -6), (7, -1, -6), (7, 6, -6), (-7, 6, -6),          # 1 plane point, 2-5 wall
                        (-1.9, 0.1, 0.4), (1.7, 0.0, -0.3), (0.1, 0.35, 2.1), (-0.3, 2.3, -1.8)], np.float64)
    M = scenegen.Material
    mats = [M(id=1, ambient=(0.4, 0.5, 0.6), diffuse=(0.6, 0.5, 0.4), specular=(0.3, 0.3, 0.3), phong=12.5),
            M(id=2, type=scenegen.MAT_MIRROR, ambient=(0.1, 0.1, 0.1), diffuse=(0.2, 0.2, 0.25), specular=(0.5, 0.5, 0.5), phong=40.0,
              mirror=(0.8, 0.75, 0.7), roughness=rough),
            M(id=3, type=scenegen.MAT_DIELECTRIC, ambient=(0.05, 0.05, 0.05), diffuse=(0.05, 0.06, 0.07), specular=(0.6, 0.6, 0.6),
              phong=80.0, ior=1.5, absorption=(0.3, 0.1, 0.05), roughness=rough),
            M(id=4, type=scenegen.MAT_CONDUCTOR, ambient=(0.1, 0.1, 0.1), diffuse=(0.3, 0.25, 0.1), specular=(0.7, 0.6, 0.3), phong=25.0,
              mirror=(0.9, 0.8, 0.5), ior=0.37, absorptionIndex=2.82, roughness=rough),
            M(id=5, ambient=(0.3, 0.3, 0.3), diffuse=(0.3, 0.7, 0.3), specular=(0.1, 0.1, 0.1), phong=0.5)]   # phong < 1: max(1, phong)
    O = scenegen.SceneObject
    objects = [O(scenegen.OBJ_PLANE, id=1, material=1, centerIdx=1, normal=(0.0, 2.0, 0.0)),       # normal not unit on purpose
               O(scenegen.OBJ_TRIANGLE, id=2, material=2, indices=(2, 3, 4)),
               O(scenegen.OBJ_TRIANGLE, id=3, material=2, indices=(2, 4, 5)),
               O(scenegen.OBJ_SPHERE, id=4, material=3, centerIdx=6, radius=1.1),
               O(scenegen.OBJ_SPHERE, id=5, material=4, centerIdx=7, radius=1.0),
               O(scenegen.OBJ_SPHERE, id=6, material=3, centerIdx=8, radius=0.65),
               O(scenegen.OBJ_SPHERE, id=7, material=5, centerIdx=9, radius=0.8)]
    lights = [((4.0, 7.0, 6.0), (38000.0, 36000.0, 34000.0)), ((-5.0, 4.0, 3.0), (6000.0, 7000.0, 9000.0))]
    if inf_light:                        # a shading point it reaches has a non-finite Lo: that level returns black (:277-280)
        lights[1] = (lights[1][0], (float("inf"),) * 3)
    sc = scenegen.Scene(maxRecursionDepth=depth, backgroundColor=(20, 30, 50), ambient=(25, 25, 25), shadowRayEpsilon=1e-3,
                        intersectionTestEpsilon=1e-6)
    sc.materials, sc.objects, sc.vertexData = mats, objects, verts
    sc.pointLights = [scenegen.PointLight(p, i) for p, i in lights]
    cam = dict(width=W, height=H, numSamples=samples, position=(0.4, 2.6, 8.5), gazePoint=(0.0, 0.4, 0.0), gaze=(0, 0, -1),
               up=(0, 1, 0), nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=42.0, isLookAt=True)
    sc.cameras = [scenegen.Camera(id=1, width=W, height=H, numSamples=samples, position=cam["position"],
                                  gazePoint=cam["gazePoint"], up=cam["up"], nearDistance=1.0, fovy=42.0, isLookAt=True)]
    path = os.path.join(tmp, f"trace_ref_{rough}_{depth}_{samples}_{int(inf_light)}.rtscene")
    scenegen.write_rtscene(sc, path)

    prims = []
    for ob in objects:
        if ob.kind == scenegen.OBJ_PLANE:
            prims.append(dict(kind="plane", mat=ob.material, c=verts[ob.centerIdx - 1], n=np.asarray(ob.normal, np.float64)))
        elif ob.kind == scenegen.OBJ_TRIANGLE:
            prims.append(dict(kind="tri", mat=ob.material, v=tuple(verts[k - 1] for k in ob.indices)))
        else:
            prims.append(dict(kind="sphere", mat=ob.material, c=verts[ob.centerIdx - 1], r=ob.radius))
    materials = [dict(type=TYPES[m.type], ambient=m.ambient, diffuse=m.diffuse, specular=m.specular, mirror=m.mirror,
                      absorption=m.absorption, phong=m.phong, ior=m.ior, absorptionIndex=m.absorptionIndex, roughness=m.roughness)
                 for m in mats]
    ref = reftrace.Scene(prims, materials, lights, sc.ambient, sc.backgroundColor, sc.intersectionTestEpsilon,
                         sc.shadowRayEpsilon, depth)
    return path, ref, cam


@pytest.mark.parametrize("rough,depth,samples,inf_light", [(0.0, 4, 4, False), (0.15, 3, 4, False), (0.0, 0, 4, False),
                                                           (0.0, 1, 7, False), (0.0, 2, 1, True)])
def test_oracle_trace_matches_independent_restatement(rough, depth, samples, inf_light, tmp_path):
    # 7 samples: n = floor(sqrt(7)) = 2 -> four rays, their sum divided by 7 (Object+Extension.swift:299-356)
    path, ref, cam = _build(str(tmp_path), rough, depth, samples=samples, inf_light=inf_light)
    want, counts = reftrace.render(ref, cam)
    got = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    lin = got["linear"]
    assert lin.shape == want.shape
    err = np.abs(lin - want) / np.maximum(1.0, np.abs(want))
    assert err.max() <= 1e-13, (err.max(), np.unravel_index(err.argmax(), err.shape))      # measured: identical bit for bit
    c = got["counters"]
    assert (c["raysPrimary"], c["raysShadow"], c["raysSecondary"]) == (counts["primary"], counts["shadow"], counts["secondary"])
    # RGBA8 (RayTracer.swift:186-195): clamp to [0, 255], UInt8(x) truncates towards zero, alpha 255
    rgb = np.clip(want, 0.0, 255.0).astype(np.uint8)
    assert np.array_equal(got["rgba"][..., :3], rgb) and (got["rgba"][..., 3] == 255).all()
    if inf_light:
        assert (want == 0).all(axis=-1).mean() > 0.2 and (want != 0).any(axis=-1).mean() > 0.05, "guard not exercised"
    if depth >= 3:
        assert rgb.max() == 255 and 0 < rgb.min() < 255               # the clamp and the truncation are both exercised
        assert counts["secondary"] > counts["primary"]                # the recursion really branches


def test_oracle_instances_and_smooth_normals_match_independent_restatement(tmp_path):
    """The instance step of intersectTLAS (Models/RTContext.swift:651-704) and the smooth normals of RTContext.init
    (:305-326, :497-500): a smooth mesh with a rotation + non-uniform scale, a mirrored MeshInstance of it with a
    material override (negative determinant -> flipped normal), an instance of that instance, and a sphere stretched
    into an ellipsoid — against a no-BVH restatement whose matrices come from numpy (inverse, inverse transpose,
    determinant), so agreement is to rounding, not bit for bit: t to 1e-9 relative, colours to 1e-7."""
    sg = scenegen
    g = 5
    rng = np.random.default_rng(3)
    xs, zs = np.meshgrid(np.linspace(-1.5, 1.5, g), np.linspace(-1.5, 1.5, g), indexing="ij")
    pos = np.stack([xs.ravel(), 0.35 * rng.standard_normal(g * g), zs.ravel()], axis=1)
    faces0 = []
    for a in range(g - 1):
        for b in range(g - 1):
            i = a * g + b
            faces0 += [(i, i + 1, i + g), (i + 1, i + g + 1, i + g)]
    faces0 = np.asarray(faces0, np.int32)
    verts = np.concatenate([pos, [[0.2, 0.1, 0.3]]])
    M_mesh = sg.mat_translate((-1.6, 0.4, 0.0)) @ sg.mat_rotate((0.3, 1.0, 0.2), 25.0) @ sg.mat_scale((1.2, 0.8, 1.1))
    M_a = sg.mat_translate((2.0, 0.9, -1.0)) @ sg.mat_scale((-1.0, 1.0, 1.0)) @ sg.mat_rotate((0.0, 0.0, 1.0), 30.0)
    M_b = sg.mat_translate((0.0, 2.6, -2.0)) @ sg.mat_rotate((1.0, 0.0, 0.0), -50.0) @ sg.mat_scale((0.7, 0.7, 0.7))
    M_s = sg.mat_translate((0.4, -0.3, 2.2)) @ sg.mat_scale((1.5, 0.6, 1.0))
    mats = [sg.Material(id=1, ambient=(0.3, 0.3, 0.3), diffuse=(0.7, 0.6, 0.5), specular=(0.3, 0.3, 0.3), phong=20.0),
            sg.Material(id=2, ambient=(0.2, 0.3, 0.4), diffuse=(0.3, 0.5, 0.8), specular=(0.5, 0.5, 0.5), phong=45.0),
            sg.Material(id=3, ambient=(0.4, 0.2, 0.2), diffuse=(0.8, 0.3, 0.2), specular=(0.2, 0.2, 0.2), phong=8.0)]
    O = sg.SceneObject
    objects = [O(sg.OBJ_MESH, id=10, material=1, smooth=True, faces=faces0 + 1, transform=sg.colmajor(M_mesh)),
               O(sg.OBJ_MESH_INSTANCE, id=11, material=2, baseMeshID=10, transform=sg.colmajor(M_a)),
               O(sg.OBJ_MESH_INSTANCE, id=12, material=3, baseMeshID=11, transform=sg.colmajor(M_b)),
               O(sg.OBJ_SPHERE, id=13, material=2, centerIdx=g * g + 1, radius=0.7, transform=sg.colmajor(M_s))]
    lights = [((3.0, 8.0, 7.0), (700.0, 700.0, 650.0))]
    W, H, samples = 36, 24, 4
    sc = sg.Scene(maxRecursionDepth=0, backgroundColor=(5, 8, 12), ambient=(30, 30, 30))
    sc.materials, sc.objects, sc.vertexData = mats, objects, verts
    sc.pointLights = [sg.PointLight(p, i) for p, i in lights]
    cam = dict(width=W, height=H, numSamples=samples, position=(0.3, 3.2, 9.0), gazePoint=(0.0, 0.8, 0.0), gaze=(0, 0, -1),
               up=(0, 1, 0), nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=40.0, isLookAt=True)
    sc.cameras = [sg.Camera(id=1, width=W, height=H, numSamples=samples, position=cam["position"], gazePoint=cam["gazePoint"],
                            up=cam["up"], nearDistance=1.0, fovy=40.0, isLookAt=True)]
    path = os.path.join(str(tmp_path), "instances_ref.rtscene")
    sg.write_rtscene(sc, path)

    tri = pos[faces0]                                                     # (T, 3, 3)
    vn = reftrace.smooth_vertex_normals(pos, faces0)[faces0]
    mesh = dict(kind="mesh", v=tri, normals=vn, first=0)
    prims = [dict(mesh, M=M_a, mat=2), dict(mesh, M=M_b, mat=3), dict(mesh, M=M_mesh, mat=1),
             dict(kind="sphere", c=verts[-1], r=0.7, M=M_s, mat=2)]
    materials = [dict(type="plain", ambient=m.ambient, diffuse=m.diffuse, specular=m.specular, mirror=m.mirror,
                      absorption=m.absorption, phong=m.phong, ior=m.ior, absorptionIndex=m.absorptionIndex, roughness=0.0)
                 for m in mats]
    ref = reftrace.prepare(reftrace.Scene(prims, materials, lights, sc.ambient, sc.backgroundColor, 1e-6, 1e-3, 0))

    got = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    ids, lin = got["ids"], got["linear"]
    agree = total = hits = 0
    worst = 0.0
    for j in range(H):
        for i in range(W):
            colour = np.zeros(3)
            same = True
            for s, (o, d, tmin) in enumerate(refmath.primary_rays(cam, i, j)):
                h = reftrace.closest_any(ref, o, d, tmin)
                r = ids[j, i, s]
                total += 1
                if h is None:
                    ok = int(r["kind"]) == 0
                    colour = colour + ref.background
                else:
                    hits += 1
                    kind = 2 if h[4] == "sphere" else 1
                    ok = int(r["kind"]) == kind and (kind == 2 or int(r["prim"]) == h[5]) and \
                        abs(float(r["t"]) - h[0]) <= 1e-9 * h[0]
                    colour = colour + reftrace.shade_direct(ref, d, h)
                agree += ok
                same = same and ok
            if same:
                want = colour / float(samples)
                worst = max(worst, float(np.max(np.abs(lin[j, i] - want) / np.maximum(1.0, np.abs(want)))))
    print("hits", hits, "agree", agree, "of", total, "worst colour error", worst)
    assert hits > 0.15 * total                     # the objects fill a good part of the frame
    assert agree >= 0.998 * total, (agree, total)  # silhouette samples may flip under the matrices' rounding
    assert worst <= 1e-7, worst




def _aabb_transformed(box, M):
    """AABB.transformed(by:), Accelearion/AABB.swift:71-92: centre through the matrix, half extents through |R|."""
    lo, hi = np.asarray(box[:3]), np.asarray(box[3:])
    c_old, e_old = 0.5 * (lo + hi), 0.5 * (hi - lo)
    R, t = M[:3, :3], M[:3, 3]
    c_new = ((R[:, 0] * c_old[0] + R[:, 1] * c_old[1]) + R[:, 2] * c_old[2]) + t
    A = np.abs(R)
    e_new = (A[:, 0] * e_old[0] + A[:, 1] * e_old[1]) + A[:, 2] * e_old[2]
    return np.concatenate([c_new - e_new, c_new + e_new])


def test_oracle_instance_records_match_independent_restatement(tmp_path):
    """makeInstance (Models/RTContext.swift:437-455): worldBounds = union over the BLAS's primitives of their boxes
    moved by AABB.transformed (every triangle's box, not the mesh's box), worldToLocal = inverse, normal matrix =
    inverse transpose of the upper 3x3 — for the analytic primitives, the MeshInstances (one mirrored, one an instance
    of an instance) and the base mesh of config_instances, in RTContext's order (:96-410)."""
    import ctypes as C
    sg = scenegen
    path = sg.config_instances(str(tmp_path), width=16, height=12, samples=1, frame=0.35)
    orc = oracle_lib.OracleScene(path)
    inst = orc.instances()
    L = oracle_lib.lib()

    # unit: AABB.transformed on random boxes and matrices (rotation, shear, mirror)
    rng = np.random.default_rng(8)
    for _ in range(20):
        lo = rng.uniform(-3, 3, 3)
        box = np.concatenate([lo, lo + rng.uniform(0, 2, 3)])
        M = np.eye(4)
        M[:3, :4] = rng.uniform(-2, 2, (3, 4))
        out = np.zeros(6)
        Mc = np.ascontiguousarray(M.T).ravel()                               # column-major
        L.orc_aabb_transformed(box.ctypes.data_as(C.c_void_p), Mc.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert np.allclose(out, _aabb_transformed(box, M), rtol=1e-14, atol=1e-14)

    # the scene's instances
    pos, tris = sg.icosphere_mesh(2, radius=1.0, seed=7, disp=0.15)
    pos = pos.astype(np.float32).astype(np.float64)                           # PLYReader.swift:85-101
    tri = pos[tris]
    tri_boxes = np.concatenate([tri.min(axis=1), tri.max(axis=1)], axis=1)
    l2w = inst["l2w"].reshape(-1, 4, 4).transpose(0, 2, 1)                    # stored column-major
    n = len(l2w)
    assert n == 6                                                             # plane, sphere, 3 mesh instances, base mesh
    seen_mesh = 0
    for k in range(n):
        M = l2w[k]
        assert np.allclose(inst["w2l"][k].reshape(4, 4).T, np.linalg.inv(M), rtol=1e-12, atol=1e-12)
        assert np.allclose(inst["nmat"][k].reshape(3, 3).T, np.linalg.inv(M[:3, :3]).T, rtol=1e-12, atol=1e-12)
        got = inst["bounds"][k]
        cands = {"plane": _aabb_transformed(np.array([-1e5] * 3 + [1e5] * 3), M),
                 "sphere": _aabb_transformed(np.array([0.8 - 1.0, 1.1 - 1.0, 4.0 - 1.0, 0.8 + 1.0, 1.1 + 1.0, 4.0 + 1.0]), M)}
        moved = np.asarray([_aabb_transformed(b, M) for b in tri_boxes])
        cands["mesh"] = np.concatenate([moved[:, :3].min(axis=0), moved[:, 3:].max(axis=0)])
        kind = [name for name, want in cands.items() if np.allclose(got, want, rtol=1e-12, atol=1e-9)]
        assert len(kind) == 1, (k, got, cands)
        seen_mesh += kind[0] == "mesh"
    assert seen_mesh == 4


def _mul(M, v, w):
    """simd_double4x4 * SIMD4(v, w): columns scaled and added left to right (column-major M given as 4x4 numpy)."""
    return (((M[:3, 0] * v[0] + M[:3, 1] * v[1]) + M[:3, 2] * v[2]) + M[:3, 3] * w)


def test_oracle_instance_motion_matches_independent_restatement(tmp_path):
    """MeshInstance.motionBlur (Models/RTContext.swift:651-704): the TLAS boxes do not follow the motion (makeInstance,
    :437-455), the box tests see the undisplaced ray, and only inside an instance is the origin moved by
    -instanceMotion * time — so which displaced instances a ray can reach depends on the reference's own TLAS.  A
    pure-Python walk of tests/refbvh.py's tree over the instance bounds, brute force inside each instance reached,
    against the oracle's primary hits: same primitive, t bit for bit, for every sample (4 spp: 4 ray times a pixel)."""
    import refbvh
    sg = scenegen
    W, H, samples, motion = 40, 30, 4, 1.2
    path = sg.config_instances(str(tmp_path), width=W, height=H, samples=samples, instance_motion=motion)
    orc = oracle_lib.OracleScene(path)
    inst = orc.instances()
    ids = orc.render(want_ids=True)["ids"]
    cam = dict(width=W, height=H, numSamples=samples, position=(0.5, 4.5, 12.0), gazePoint=(0.0, 1.2, 0.0), gaze=(0, 0, -1),
               up=(0, 1, 0), nearPlane=(-1, 1, -1, 1), nearDistance=1.0, fovy=45.0, isLookAt=True)
    pos, tris = sg.icosphere_mesh(2, radius=1.0, seed=7, disp=0.15)
    tri = pos.astype(np.float32).astype(np.float64)[tris]
    w2l = inst["w2l"].reshape(-1, 4, 4).transpose(0, 2, 1)
    move = {6: (0.9 * motion, 0.0, 0.3 * motion), 7: (0.0, 0.7 * motion, 0.0), 8: (-0.5 * motion, -0.4 * motion, 0.6 * motion)}
    obj = inst["meta"][:, 0]
    n = len(obj)
    bounds = inst["bounds"]
    tlas = refbvh.Builder(bounds, (bounds[:, :3] + bounds[:, 3:]) * 0.5)      # buildTLAS, :459-474
    eps = 1e-6
    box = (np.full(3, -1e5), np.full(3, 1e5))
    moved_hits = total_hits = 0
    for j in range(H):
        for i in range(W):
            for s, (o, d, tmin, time) in enumerate(refmath.primary_rays(cam, i, j, with_time=True)):
                found = {}

                def test(k, best_t):
                    ol = _mul(w2l[k], o - np.asarray(move.get(int(obj[k]), (0.0, 0.0, 0.0))) * time, 1.0)
                    dl = _mul(w2l[k], d, 0.0)
                    if inst["meta"][k, 2] > 1:
                        p, t = refmath.brute_force_triangles(ol, dl, tri, eps, tmin)
                        kind = 1
                    elif int(obj[k]) == 2:
                        p, t, kind = 0, refmath.brute_force_sphere(ol, dl, np.array([0.8, 1.1, 4.0]), 1.0, eps, tmin), 2
                    else:
                        p, t, kind = 0, np.inf, 3
                        with np.errstate(divide="ignore"):
                            inv = 1.0 / dl
                        nrm, ctr = np.array([0.0, 1.0, 0.0]), np.zeros(3)                    # intersectPlane, :526-536
                        denom = refmath._dot(nrm, dl)
                        if reftrace._hit_aabb(box[0], box[1], ol, inv, eps) and abs(denom) >= eps:
                            tp = refmath._dot(ctr - ol, nrm) / denom
                            if tp > max(eps, tmin):
                                t = tp
                    if t < best_t:
                        found[k] = (kind, p)
                        return t
                    return None

                k, t, _ = refbvh.closest_in_blas(tlas, list(o), list(d), eps, tmin, test)
                r = ids[j, i, s]
                if k < 0:
                    assert int(r["kind"]) == 0, (i, j, s)
                    continue
                kind, p = found[k]
                total_hits += 1
                moved_hits += int(obj[k]) in move
                assert int(r["kind"]) == kind and float(r["t"]) == t, (i, j, s, r, kind, t)
                if kind == 1:
                    assert int(r["prim"]) == p and int(r["inst"]) == k, (i, j, s, r, p, k)
    assert total_hits > 0.5 * W * H * samples and moved_hits > 100, (total_hits, moved_hits)
