"""Independent restatement (pure Python on top of tests/refmath.py) of the reference's `trace` recursion —
Extensions/Object+Extension.swift:96-283: ambient + point lights with shadow rays, mirror / dielectric / conductor
children in the reference's order (reflection before refraction), glossy perturbation through the pixel's PCG32 stream,
Fresnel weights, Beer's law, the depth rules and the non-finite guard — and of the pixel loop around it (:294-356).
Written from the Swift source, NOT from oracle/: no BVH (every primitive is tried, Models/RTContext.swift:479-537 for the
closest hit, :833-870 for the shadow tests), Python recursion instead of the oracle's C++.  The CPU tests hold the
oracle's linear colours against this file, so an error in the shading path has to be made twice, independently, to go
unnoticed (SURVEY.md §8c iv).  Scenes: top-level spheres, triangles (flat) and planes — the 1-primitive BLAS + instance
of RTContext.init (:96-250) — and meshes with flat or smooth normals; a primitive without "M" is the identity instance
(bit-exact path), one with "M" (4x4 localToWorld) goes through the instance step of intersectTLAS (:651-704): ray to
local space without re-normalising the direction, hit point back through localToWorld, normal through the
inverse-transpose and flipped when the determinant is negative, material override.  Test infrastructure only."""
from __future__ import annotations

import math

import numpy as np

import refmath
from refmath import _dot, _norm

INF = math.inf


def _v(x):
    return np.asarray(x, np.float64)


class Scene:
    """prims: list of dicts {kind: 'sphere'|'tri'|'plane', mat: 1-based material id, ...geometry};
    materials: list of dicts with the Material fields trace reads; lights: [(position, intensity)]."""

    def __init__(self, prims, materials, lights, ambient, background, eps, shadow_eps, max_depth):
        self.prims, self.materials, self.lights = prims, materials, lights
        self.ambient, self.background = _v(ambient), _v(background)
        self.eps, self.shadow_eps, self.max_depth = eps, shadow_eps, max_depth


def _hit_aabb(lo, hi, o, inv, eps):
    """hitAABB, RTContext.swift:561-569 (the plane's box is finite: +-1e5, :172-174)."""
    with np.errstate(invalid="ignore"):
        t1, t2 = (lo - o) * inv, (hi - o) * inv
    tmin = max(max(min(t1[0], t2[0]), min(t1[1], t2[1])), min(t1[2], t2[2]))
    tmax = min(max(t1[0], t2[0]), min(max(t1[1], t2[1]), max(t1[2], t2[2])))
    return tmax >= max(tmin, eps)


_BOX = (_v((-1e5,) * 3), _v((1e5,) * 3))


def closest(sc: Scene, o, d, t_min):
    """intersectTLAS over 1-primitive instances with identity transforms: (t, p, n, mat, kind) or None.  The instance
    step re-normalises the normal (normalMatrix * n, :688) and adds a zero motion offset to p (:701)."""
    best = None
    best_t = INF
    with np.errstate(divide="ignore"):
        inv = 1.0 / d
    for pr in sc.prims:
        if pr["kind"] == "sphere":
            t = refmath.brute_force_sphere(o, d, pr["c"], pr["r"], sc.eps, t_min)
            if t < best_t:
                p = o + d * t
                best_t, best = t, (t, p, _norm(p - pr["c"]), pr["mat"], "sphere")
        elif pr["kind"] == "tri":
            v0, v1, v2 = pr["v"]
            h = refmath.mt_single(o, d, v0, v1, v2, sc.eps)
            if h and h[0] > max(sc.eps, t_min) and h[0] < best_t:
                e1, e2 = v1 - v0, v2 - v0
                p = (v0 + h[1] * e1) + h[2] * e2
                best_t, best = h[0], (h[0], p, _norm(np.cross(e1, e2)), pr["mat"], "tri")
        else:
            if not _hit_aabb(_BOX[0], _BOX[1], o, inv, sc.eps):
                continue
            denom = _dot(pr["n"], d)                                     # intersectPlane, :526-536
            if abs(denom) < sc.eps:
                continue
            t = _dot(pr["c"] - o, pr["n"]) / denom
            if t <= max(sc.eps, t_min) or t >= best_t:
                continue
            best_t, best = t, (t, o + d * t, _norm(pr["n"]), pr["mat"], "plane")
    if best is None:
        return None
    t, p, n, mat, kind = best
    return t, p + 0.0, _norm(n), mat, kind


def _local_ray(pr, o, d):
    """world -> local, :658-668 (no instance motion here)."""
    W = pr["W"]                                                            # worldToLocal = M.inverse (:449)
    ol = (W @ np.append(o, 1.0))[:3]
    dl = (W @ np.append(d, 0.0))[:3]
    return ol, dl


def prepare(sc: Scene):
    """Per-instance matrices: worldToLocal, normal matrix (inverse transpose of the upper 3x3, :24-31), handedness."""
    for pr in sc.prims:
        if "M" in pr:
            M = np.asarray(pr["M"], np.float64)
            pr["W"] = np.linalg.inv(M)
            pr["N"] = np.linalg.inv(M[:3, :3]).T
            pr["flip"] = np.linalg.det(M[:3, :3]) < 0
    return sc


def _closest_transformed(sc: Scene, pr, o, d, t_min, best_t):
    """One instance with a transform: (t, p_world, n_world, mat, kind, prim) or None; only hits closer than best_t."""
    ol, dl = _local_ray(pr, o, d)
    found = None
    if pr["kind"] == "sphere":
        t = refmath.brute_force_sphere(ol, dl, pr["c"], pr["r"], sc.eps, t_min)
        if t < best_t:
            p = ol + dl * t
            found = (t, p, _norm(p - pr["c"]), "sphere", 0)
    elif pr["kind"] == "mesh":
        for k, (v0, v1, v2) in enumerate(pr["v"]):
            h = refmath.mt_single(ol, dl, v0, v1, v2, sc.eps)
            if h and h[0] > max(sc.eps, t_min) and h[0] < best_t:
                e1, e2 = v1 - v0, v2 - v0
                if pr.get("normals") is not None:                         # smooth: w*n0 + u*n1 + v*n2 (:497-500)
                    n0, n1, n2 = pr["normals"][k]
                    n = _norm(((1.0 - h[1] - h[2]) * n0 + h[1] * n1) + h[2] * n2)
                else:
                    n = _norm(np.cross(e1, e2))
                best_t = h[0]
                found = (h[0], (v0 + h[1] * e1) + h[2] * e2, n, "tri", pr["first"] + k)
    else:
        raise ValueError(pr["kind"])
    if found is None:
        return None
    t, p, n, kind, prim = found
    pw = (np.asarray(pr["M"], np.float64) @ np.append(p, 1.0))[:3]
    nw = _norm(pr["N"] @ n)
    if pr["flip"]:
        nw = -nw
    return t, pw, nw, pr["mat"], kind, prim


def closest_any(sc: Scene, o, d, t_min):
    """closest() over identity instances and transformed ones: (t, p, n, mat, kind, prim) or None."""
    plain = Scene([p for p in sc.prims if "M" not in p], sc.materials, sc.lights, sc.ambient, sc.background, sc.eps,
                  sc.shadow_eps, sc.max_depth)
    h = closest(plain, o, d, t_min)
    best = None if h is None else h + (-1,)
    for pr in sc.prims:
        if "M" in pr:
            c = _closest_transformed(sc, pr, o, d, t_min, INF if best is None else best[0])
            if c is not None:
                best = c
    return best


def occluded_any(sc: Scene, o, d, t_max):
    plain = Scene([p for p in sc.prims if "M" not in p], sc.materials, sc.lights, sc.ambient, sc.background, sc.eps,
                  sc.shadow_eps, sc.max_depth)
    if occluded(plain, o, d, t_max):
        return True
    for pr in sc.prims:
        if "M" not in pr:
            continue
        ol, dl = _local_ray(pr, o, d)
        if pr["kind"] == "sphere":
            one = Scene([dict(kind="sphere", c=pr["c"], r=pr["r"], mat=pr["mat"])], sc.materials, sc.lights, sc.ambient,
                        sc.background, sc.eps, sc.shadow_eps, sc.max_depth)
            if occluded(one, ol, dl, t_max):
                return True
        else:
            for v0, v1, v2 in pr["v"]:
                h = refmath.mt_single(ol, dl, v0, v1, v2, sc.eps)
                if h and max(sc.eps, 0.0) < h[0] < t_max:
                    return True
    return False


def shade_direct(sc: Scene, d, hit):
    """Depth-0 shading of a plain material (:105-143): ambient + point lights with shadow rays."""
    t, p, n_geo, mat_id = hit[:4]
    mat = sc.materials[max(0, min(len(sc.materials) - 1, mat_id - 1))]
    front = _dot(d, n_geo) < 0
    N = n_geo if front else -n_geo
    Lo = sc.ambient * _v(mat["ambient"])
    for lpos, lint in sc.lights:
        wi = _v(lpos) - p
        dist = math.sqrt(_dot(wi, wi))
        wi = _norm(wi)
        if occluded_any(sc, p + wi * sc.shadow_eps, wi, dist):
            continue
        n_dot_l = max(0.0, _dot(N, wi))
        if n_dot_l > 0:
            ld = _v(mat["diffuse"]) * n_dot_l
            h = _norm(wi + _norm(-d))
            ls = _v(mat["specular"]) * math.pow(max(0.0, _dot(N, h)), max(1.0, mat["phong"]))
            Lo = Lo + (ld + ls) * (_v(lint) / max(dist * dist, 1e-12))
    return Lo


def smooth_vertex_normals(pos, faces0):
    """RTContext.swift:305-326: unit face normals added to their three vertices in (face, corner) order, then
    normalised.  faces0: zero-based (T, 3)."""
    vn = np.zeros_like(pos)
    for f in faces0:
        v0, v1, v2 = pos[f[0]], pos[f[1]], pos[f[2]]
        fn = _norm(np.cross(v1 - v0, v2 - v0))
        for c in range(3):
            vn[f[c]] = vn[f[c]] + fn
    return np.asarray([_norm(x) for x in vn])


def occluded(sc: Scene, o, d, t_max):
    """occludedTLAS / occludedBLAS, :723-870, for a shadow ray with tMin = 0."""
    with np.errstate(divide="ignore"):
        inv = 1.0 / d
    for pr in sc.prims:
        if pr["kind"] == "sphere":                                        # sphereShadowHit, :851-861
            oc = o - pr["c"]
            a, b, c = _dot(d, d), 2.0 * _dot(oc, d), _dot(oc, oc) - pr["r"] * pr["r"]
            disc = b * b - 4 * a * c
            if disc < 0:
                continue
            sd = math.sqrt(disc)
            t = (-b - sd) / (2 * a)
            if t < max(sc.eps, 0.0):
                t = (-b + sd) / (2 * a)
            if max(sc.eps, 0.0) < t < t_max:
                return True
        elif pr["kind"] == "tri":                                         # triShadowHit, :833-849
            h = refmath.mt_single(o, d, *pr["v"], sc.eps)
            if h and max(sc.eps, 0.0) < h[0] < t_max:
                return True
        else:                                                             # planeShadowHit, :863-869
            if not _hit_aabb(_BOX[0], _BOX[1], o, inv, sc.eps):
                continue
            denom = _dot(pr["n"], d)
            if abs(denom) < sc.eps:
                continue
            t = _dot(pr["c"] - o, pr["n"]) / denom
            if max(sc.eps, 0.0) < t < t_max:
                return True
    return False


def reflect(d, n):
    return d - (2.0 * _dot(d, n)) * n                                      # :467


def fresnel_conductor_rgb(eta, k, cos_i):
    """:492-505, vector form (all three components equal)."""
    r = refmath.fresnel_conductor(eta, k, cos_i)
    return _v((r, r, r))


def _glossy(rd, rough, rng):
    """:191-197 / :259-265."""
    t, b = refmath.orthonormal_basis(rd)
    rand1 = rng.next_float() - 0.5
    rand2 = rng.next_float() - 0.5
    rd = (rd + (rough * rand1) * b) + (rough * rand2) * t
    return _norm(rd)


def trace(sc: Scene, o, d, t_min, depth, rng, counts):
    """-> (colour, first hit or None).  counts: dict of ray counters (primary is counted by the caller)."""
    if depth > sc.max_depth:
        return sc.background.copy(), None
    hit = closest(sc, o, d, t_min)
    if hit is None:
        return sc.background.copy(), None
    t, p, n_geo, mat_id, _kind = hit
    mat = sc.materials[max(0, min(len(sc.materials) - 1, mat_id - 1))]
    front = _dot(d, n_geo) < 0
    N = n_geo if front else -n_geo
    has_refraction = mat["ior"] > 0
    direct = (not has_refraction) or front
    Lo = sc.ambient * _v(mat["ambient"]) if direct else np.zeros(3)
    if direct:
        for lpos, lint in sc.lights:
            wi = _v(lpos) - p
            dist = math.sqrt(_dot(wi, wi))
            wi = _norm(wi)
            counts["shadow"] += 1
            if occluded(sc, p + wi * sc.shadow_eps, wi, dist):
                continue
            n_dot_l = max(0.0, _dot(N, wi))
            if n_dot_l > 0:
                shininess = max(1.0, mat["phong"])
                ld = _v(mat["diffuse"]) * n_dot_l
                view = _norm(-d)
                h = _norm(wi + view)
                ls = _v(mat["specular"]) * math.pow(max(0.0, _dot(N, h)), shininess)
                atten = _v(lint) / max(dist * dist, 1e-12)
                Lo = Lo + (ld + ls) * atten
    rough = mat["roughness"]
    if mat["type"] == "mirror" and depth < sc.max_depth:
        rd = _norm(reflect(d, N))
        if rough != 0:
            rd = _glossy(rd, rough, rng)
        counts["secondary"] += 1
        li, _ = trace(sc, p + N * sc.shadow_eps, rd, 0.0, depth + 1, rng, counts)
        Lo = Lo + _v(mat["mirror"]) * li
    elif mat["type"] == "dielectric" and depth < sc.max_depth:
        mN = N
        if rough != 0:
            tb, bb = refmath.orthonormal_basis(N)
            r1 = (rng.next_float() * 2.0) - 1.0
            r2 = (rng.next_float() * 2.0) - 1.0
            mN = _norm((N + (tb * r1) * rough) + (bb * r2) * rough)
        n1, n2 = (1.0, mat["ior"]) if front else (mat["ior"], 1.0)
        cos_i = -_dot(d, mN)
        R, cos_t = refmath.fresnel_dielectric(n1, n2, cos_i)
        rd = _norm(reflect(d, mN))
        counts["secondary"] += 1
        li_r, _ = trace(sc, p + rd * sc.shadow_eps, rd, 0.0, depth + 1, rng, counts)
        if cos_t is None:
            Lo = Lo + li_r
        else:
            eta = n1 / n2
            td = _norm((d * eta) + (mN * (eta * cos_i - cos_t)))
            counts["secondary"] += 1
            li_t, t_hit = trace(sc, p + td * sc.shadow_eps, td, 0.0, depth + 1, rng, counts)
            absorb = _v(mat["absorption"])
            if front and absorb.any() and t_hit is not None and t_hit[3] == mat_id:
                dd = max(t_hit[0], 0.0)
                if math.isfinite(dd) and dd > 0:                         # beerAttenuate, :470-474
                    li_t = li_t * _v([math.exp(-absorb[k] * dd) for k in range(3)])
                else:
                    li_t = li_t * li_t
            Lo = Lo + (li_r * R + li_t * (1.0 - R))
    elif mat["type"] == "conductor" and depth < sc.max_depth:
        cos_i = max(0.0, -_dot(d, N))
        rf = fresnel_conductor_rgb(mat["ior"], mat["absorptionIndex"], cos_i)
        rd = _norm(reflect(d, N))
        if rough != 0:
            rd = _glossy(rd, rough, rng)
        counts["secondary"] += 1
        li, _ = trace(sc, p + N * sc.shadow_eps, rd, 0.0, depth + 1, rng, counts)
        Lo = Lo + (rf * _v(mat["mirror"])) * li
    if not np.isfinite(Lo).all():
        return np.zeros(3), hit
    return Lo, hit


def render(sc: Scene, cam: dict):
    """Renderer.render's pixel loop (:294-356) without depth of field -> (linear (H, W, 3), counts)."""
    W, H = max(1, cam["width"]), max(1, cam["height"])
    samples = max(1, cam["numSamples"])
    n = int(math.sqrt(samples))
    out = np.zeros((H, W, 3))
    counts = dict(primary=0, shadow=0, secondary=0)
    basis = refmath.camera_basis(cam)
    for j in range(H):
        for i in range(W):
            rng = refmath.PCG32(refmath.pixel_seed(i, j))
            colour = np.zeros(3)
            for sy in range(n):
                for sx in range(n):
                    xi1, xi2 = rng.next_float(), rng.next_float()
                    ci = float(i) + (float(sx) + xi1) / float(n)
                    cj = float(j) + (float(sy) + xi2) / float(n)
                    s = (basis["q00"] - basis["v"] * (cj * basis["dv"])) + basis["u"] * (ci * basis["du"])
                    d = _norm(s - basis["e"])
                    rng.next_float()                                      # ray time (:340): drawn, unused without motion
                    denom = _dot(d, basis["w"])
                    t_img = _dot(basis["m"] - basis["e"], basis["w"]) / (denom if denom != 0 else 5e-324)
                    counts["primary"] += 1
                    c, _ = trace(sc, basis["e"].copy(), d, max(t_img, 0.0), 0, rng, counts)
                    colour = colour + c
            out[j, i] = colour / float(samples)
    return out, counts
