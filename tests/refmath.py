"""Second, independent restatement (numpy / pure Python) of the pieces of the reference that pin the oracle:
PCG32, the pixel seed, primary-ray generation and a brute-force (no BVH) closest hit.  Written from the
reference's source (cited per function), NOT from oracle/: the CPU tests check oracle/ against this file, so an
error has to be made twice, independently, to go unnoticed (SURVEY.md §8c iv).  Test infrastructure only."""
from __future__ import annotations

import math

import numpy as np

M64 = (1 << 64) - 1


class PCG32:
    """Extensions/Object+Extension.swift:556-589 (non-standard seeding :561-567)."""

    def __init__(self, seed: int):
        self.state = 0
        self.inc = ((seed << 1) | 1) & M64
        self.next()
        self.state = (self.state + 0x9E3779B97F4A7C15) & M64
        self.next()

    def next(self) -> int:
        old = self.state
        self.state = (old * 6364136223846793005 + self.inc) & M64
        xorshifted = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = (old >> 59) & 31
        return ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xFFFFFFFF

    def next_float(self) -> float:
        return float(self.next()) * 2.3283064365386963e-10


def pixel_seed(i: int, j: int) -> int:
    """:294 — Swift: `^` and `&+` share AdditionPrecedence and associate left."""
    return ((((j << 32) & M64) ^ i) + 0x9E3779B97F4A7C15) & M64


def _norm(v):
    return v * (1.0 / math.sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]))


def _dot(a, b):
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]


def camera_basis(cam: dict):
    """makeCameraBasis, :382-427, + du, dv, m, q00 (:67-70).  cam: dict of the Camera fields."""
    W, H = max(1, cam["width"]), max(1, cam["height"])
    aspect = W / H
    e = np.asarray(cam["position"], np.float64)
    nd = cam["nearDistance"]
    if cam["isLookAt"]:
        w = -_norm(np.asarray(cam["gazePoint"], np.float64) - e)
        u = _norm(np.cross(_norm(np.asarray(cam["up"], np.float64)), w))
        v = _norm(np.cross(w, u))
        t = nd * math.tan((cam["fovy"] * math.pi) / (2.0 * 180.0)) if cam.get("fovy") is not None else nd * 0.5
        b, r = -t, t * aspect
        l = -r
    else:
        l, r, b, t = cam["nearPlane"]
        w = -_norm(np.asarray(cam["gaze"], np.float64))
        u = _norm(np.cross(_norm(np.asarray(cam["up"], np.float64)), w))
        v = _norm(np.cross(w, u))
    m = e - w * nd
    q00 = (m + u * l) + v * t
    return dict(e=e, u=u, v=v, w=w, m=m, q00=q00, du=(r - l) / W, dv=(t - b) / H, W=W, H=H)


def primary_rays(cam: dict, i: int, j: int, with_time: bool = False):
    """All n*n primary rays of pixel (i, j): list of (origin, dir, tMin[, time]), :294-344, thin lens (:325-338)
    when the camera has apertureSize > 0 and focusDistance > 0."""
    B = camera_basis(cam)
    samples = max(1, cam["numSamples"])
    n = int(math.sqrt(samples))
    aperture, focus = cam.get("apertureSize", 0.0), cam.get("focusDistance", 0.0)
    rng = PCG32(pixel_seed(i, j))
    out = []
    for sy in range(n):
        for sx in range(n):
            xi1, xi2 = rng.next_float(), rng.next_float()
            ci = float(i) + (float(sx) + xi1) / float(n)
            cj = float(j) + (float(sy) + xi2) / float(n)
            s = (B["q00"] - B["v"] * (cj * B["dv"])) + B["u"] * (ci * B["du"])
            d = _norm(s - B["e"])
            eye = B["e"].copy()
            if aperture > 0 and focus > 0:
                denom = _dot(d, -B["w"])
                t_focus = focus if abs(denom) < 1e-6 else focus / denom
                p_focus = B["e"] + d * t_focus
                u_rand = rng.next_float() - 0.5
                v_rand = rng.next_float() - 0.5
                eye = B["e"] + (u_rand * B["u"] + v_rand * B["v"]) * aperture
                d = _norm(p_focus - eye)
            time = rng.next_float()  # ray time (:340)
            t_img = _dot(B["m"] - eye, B["w"]) / _dot(d, B["w"])
            out.append((eye, d, max(t_img, 0.0), time) if with_time else (eye, d, max(t_img, 0.0)))
    return out


def brute_force_triangles(o, d, tris, eps, t_min):
    """Closest Moeller-Trumbore hit over ALL triangles, RTContext.swift:479-494 — vectorised, no BVH.
    tris: (N, 3, 3).  Returns (index or -1, t)."""
    v0, e1, e2 = tris[:, 0], tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]
    pvec = np.cross(d[None, :], e2)
    det = (e1[:, 0] * pvec[:, 0] + e1[:, 1] * pvec[:, 1]) + e1[:, 2] * pvec[:, 2]
    ok = np.abs(det) >= eps
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / det
        tvec = o[None, :] - v0
        u = ((tvec[:, 0] * pvec[:, 0] + tvec[:, 1] * pvec[:, 1]) + tvec[:, 2] * pvec[:, 2]) * inv
        q = np.cross(tvec, e1)
        v = ((d[0] * q[:, 0] + d[1] * q[:, 1]) + d[2] * q[:, 2]) * inv
        t = ((e2[:, 0] * q[:, 0] + e2[:, 1] * q[:, 1]) + e2[:, 2] * q[:, 2]) * inv
    ok &= (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > max(eps, t_min))
    if not ok.any():
        return -1, math.inf
    t = np.where(ok, t, np.inf)
    k = int(np.argmin(t))
    return k, float(t[k])


def brute_force_sphere(o, d, c, r, eps, t_min):
    """intersectSphere, RTContext.swift:512-523."""
    oc = o - c
    a = _dot(d, d)
    b = 2.0 * _dot(oc, d)
    cc = _dot(oc, oc) - r * r
    disc = b * b - 4 * a * cc
    if disc < 0:
        return math.inf
    sd = math.sqrt(disc)
    t = (-b - sd) / (2 * a)
    if t < eps:
        t = (-b + sd) / (2 * a)
    return t if t > max(eps, t_min) else math.inf


def fresnel_dielectric(n1, n2, cos_i):
    """:476-490 -> (R, cosT or None)."""
    c = max(0.0, min(1.0, abs(cos_i)))
    eta = n1 / n2
    s2 = eta * eta * max(0.0, 1.0 - c * c)
    if s2 > 1.0:
        return 1.0, None
    cp = math.sqrt(max(0.0, 1.0 - s2))
    rs = (n2 * c - n1 * cp) / (n2 * c + n1 * cp)
    rp = (n1 * c - n2 * cp) / (n1 * c + n2 * cp)
    return 0.5 * (rs * rs + rp * rp), cp


def fresnel_conductor(eta, k, cos_i):
    """:492-505."""
    c = max(0.0, min(1.0, abs(cos_i)))
    c2 = c * c
    e = eta * eta + k * k
    t = 2.0 * eta * c
    rs = (e - t + c2) / (e + t + c2)
    rp = ((e * c2) - t + 1.0) / ((e * c2) + t + 1.0)
    return 0.5 * (rs + rp)


# ---------------------------------------------------------------------------------- area lights (:145-186)
def jitter_table():
    """buildStratifiedJitter, :646-659, from PCG32(seed: 0x123456789ABCDEF) (:92-93)."""
    rng = PCG32(0x123456789ABCDEF)
    xs, ys = [], []
    for gy in range(10):
        for gx in range(10):
            xs.append(float(gx) + rng.next_float())
            ys.append(float(gy) + rng.next_float())
    return xs, ys


def orthonormal_basis(n):
    """:530-552."""
    sign = 1.0 if n[2] >= 0 else -1.0
    a = -1.0 / (sign + n[2])
    b = n[0] * n[1] * a
    t = _norm(np.array([1.0 + sign * n[0] * n[0] * a, sign * b, -sign * n[0]]))
    bt = _norm(np.array([b, sign + n[1] * n[1] * a, -n[1]]))
    return t, bt


def mt_single(o, d, v0, v1, v2, eps):
    """One Moeller-Trumbore test (RTContext.swift:479-494) -> (t, u, v) or None; the caller applies the t window."""
    e1, e2 = v1 - v0, v2 - v0
    pvec = np.cross(d, e2)
    det = _dot(e1, pvec)
    if abs(det) < eps:
        return None
    inv = 1.0 / det
    tvec = o - v0
    u = _dot(tvec, pvec) * inv
    if u < 0 or u > 1:
        return None
    q = np.cross(tvec, e1)
    v = _dot(d, q) * inv
    if v < 0 or u + v > 1:
        return None
    return _dot(e2, q) * inv, u, v


def area_light_sample(light: dict, k: int, xs, ys, p, N):
    """:146-171 for jitter entry k: (nL, wi, dist, dist2, NdotL, Ln), or None when the sample does not face p."""
    nL = _norm(np.asarray(light["normal"], np.float64))
    t, b = orthonormal_basis(nL)
    size = light["size"]
    r1 = xs[k % 100] / 10.0 - 0.5
    r2 = ys[k % 100] / 10.0 - 0.5
    sample = (np.asarray(light["position"], np.float64) + t * (r1 * size)) + b * (r2 * size)
    wi = sample - p
    dist2 = _dot(wi, wi)
    dist = math.sqrt(dist2)
    wi = wi / dist
    n_dot_l = _dot(N, wi)
    if n_dot_l <= 0:
        return None
    ln = abs(_dot(nL, -wi))
    if ln <= 0:
        return None
    return nL, wi, dist, dist2, n_dot_l, ln


def area_light_term(light: dict, mat: dict, N, ray_dir, wi, dist2, n_dot_l, ln):
    """:175-185: (diffuse*NdotL + specular*pow(max(0, N.h), phong)) * radiance * (Ln / dist2) * size^2."""
    view = _norm(-ray_dir)
    h = _norm(wi + view)
    ld = np.asarray(mat["diffuse"], np.float64) * n_dot_l
    ls = np.asarray(mat["specular"], np.float64) * math.pow(max(0.0, _dot(N, h)), mat["phong"])
    return (((ld + ls) * np.asarray(light["radiance"], np.float64)) * (ln / dist2)) * (light["size"] * light["size"])
