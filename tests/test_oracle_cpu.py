"""CPU tests of the oracle (oracle/liboracle.so): known-answer tests derived by hand from the reference's
source (the reference ships no tests of its own) and differential tests against the second,
independent restatement in tests/refmath.py.  No GPU needed."""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

import oracle_lib
import refmath
from myraytracer_b200 import scenegen

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INF = float("inf")


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def v3(*x):
    return np.asarray(x, np.float64)


# ---------------------------------------------------------------------------------- PCG32 / seeds
def test_pcg32_matches_independent_python_and_golden():
    L = oracle_lib.lib()
    gold = json.load(open(os.path.join(GOLDEN, "pcg32.json")))
    for seed_s, want in gold.items():
        seed = int(seed_s)
        out = np.zeros(len(want), np.uint32)
        L.orc_pcg32(seed, len(want), _p(out))
        r = refmath.PCG32(seed)
        mine = [r.next() for _ in want]
        assert out.tolist() == mine == want


def test_pcg32_is_not_the_canonical_seeding():
    # canonical pcg32_srandom(42, 54) first output is 0xa15c02b7; the reference's recipe (:561-567) differs
    r = refmath.PCG32(42)
    assert r.next() != 0xA15C02B7


def test_pixel_seed_uses_swift_precedence():
    L = oracle_lib.lib()
    differs = 0
    for i, j in [(0, 0), (1, 0), (0, 1), (799, 799), (1919, 1079), (3, 0xFFFFFFFF)]:
        want = ((((j << 32) & refmath.M64) ^ i) + 0x9E3779B97F4A7C15) & refmath.M64
        assert L.orc_pixel_seed(i, j) == want == refmath.pixel_seed(i, j)
        c_precedence = (((j << 32) & refmath.M64) ^ ((i + 0x9E3779B97F4A7C15) & refmath.M64))
        differs += want != c_precedence       # `(j<<32) ^ (i + C)` is the wrong (C-precedence) reading
    assert differs >= 3


def test_jitter_table():
    L = oracle_lib.lib()
    xs, ys = np.zeros(100), np.zeros(100)
    L.orc_jitter_table(_p(xs), _p(ys))
    r = refmath.PCG32(0x123456789ABCDEF)          # :92, table filled gy-major, x draw then y draw (:646-659)
    k = 0
    for gy in range(10):
        for gx in range(10):
            assert xs[k] == gx + r.next_float()
            assert ys[k] == gy + r.next_float()
            k += 1


# ---------------------------------------------------------------------------------- Fresnel / ONB
def test_fresnel_dielectric_kats():
    L = oracle_lib.lib()
    out = np.zeros(4)
    L.orc_fresnel_dielectric(1.0, 1.5, 1.0, _p(out))          # normal incidence air->glass: ((1.5-1)/(1.5+1))^2
    assert abs(out[0] - 0.04) < 1e-15 and out[1] == 1.0 and out[2] == 1.0
    L.orc_fresnel_dielectric(1.5, 1.0, 0.1, _p(out))          # beyond the critical angle: TIR
    assert out[0] == 1.0 and out[1] == 0.0 and out[3] > 1.0
    L.orc_fresnel_dielectric(1.0, 1.5, -0.3, _p(out))         # |cosI| is used (:479)
    R, cp = refmath.fresnel_dielectric(1.0, 1.5, 0.3)
    assert out[0] == R and out[2] == cp
    rng = np.random.default_rng(7)
    for _ in range(200):
        n1, n2, c = rng.uniform(1, 2.5), rng.uniform(1, 2.5), rng.uniform(-1.2, 1.2)
        L.orc_fresnel_dielectric(n1, n2, c, _p(out))
        R, cp = refmath.fresnel_dielectric(n1, n2, c)
        assert out[0] == R and (out[1] == 1.0) == (cp is not None)


def test_fresnel_conductor_kats():
    L = oracle_lib.lib()
    # normal incidence: ((eta-1)^2 + k^2) / ((eta+1)^2 + k^2)
    eta, k = 0.37, 2.82
    want = ((eta - 1) ** 2 + k * k) / ((eta + 1) ** 2 + k * k)
    assert abs(L.orc_fresnel_conductor(eta, k, 1.0) - want) < 1e-15
    for c in (0.0, 0.2, 0.77, 1.0, 1.7, -0.4):
        assert L.orc_fresnel_conductor(eta, k, c) == refmath.fresnel_conductor(eta, k, c)


def test_orthonormal_basis():
    L = oracle_lib.lib()
    rng = np.random.default_rng(3)
    for _ in range(100):
        n = rng.normal(size=3)
        n /= np.linalg.norm(n)
        t, b = np.zeros(3), np.zeros(3)
        L.orc_orthonormal_basis(_p(n), _p(t), _p(b))
        assert abs(np.dot(t, n)) < 1e-12 and abs(np.dot(b, n)) < 1e-12 and abs(np.dot(t, b)) < 1e-12
        assert abs(np.linalg.norm(t) - 1) < 1e-12 and abs(np.linalg.norm(b) - 1) < 1e-12


# ---------------------------------------------------------------------------------- primitive tests
def test_slab_test_rules():
    L = oracle_lib.lib()
    lo, hi = v3(-1, -1, -1), v3(1, 1, 1)
    hit = lambda o, d, eps=1e-6: L.orc_hit_aabb(_p(lo), _p(hi), _p(v3(*o)), _p(v3(*d)), eps)
    assert hit((0, 0, -5), (0, 0, 1)) == 4.0                      # tmin returned
    assert hit((0, 0, 0), (0, 0, 1)) == -1.0                      # origin inside: NEGATIVE tmin (RTContext.swift:557-565)
    assert hit((0, 0, 5), (0, 0, 1)) == INF                       # box behind: tmax < eps
    assert hit((3, 0, -5), (0, 0, 1)) == INF                      # parallel miss (inf * 0 lanes)
    assert hit((0, 0, -5), (0, 0, 1), eps=10.0) == INF            # hit iff tmax >= max(tmin, eps): tmax = 6 < 10
    assert hit((0, 0, -5), (0, 0, 1), eps=6.0) == 4.0             # tmax == eps counts


def test_moller_trumbore_rules():
    L = oracle_lib.lib()
    tri = np.asarray([0, 0, 0, 1, 0, 0, 0, 1, 0], np.float64)
    f = lambda o, d, tmin=0.0, hitt=INF, eps=1e-6: L.orc_intersect_triangle(_p(tri), _p(v3(*o)), _p(v3(*d)), tmin, hitt, eps, None)
    assert f((0.25, 0.25, 1), (0, 0, -1)) == 1.0
    assert f((0.25, 0.25, -1), (0, 0, 1)) == 1.0                  # two-sided
    assert f((0.75, 0.75, 1), (0, 0, -1)) == INF                  # u + v > 1
    assert f((-0.1, 0.5, 1), (0, 0, -1)) == INF                   # u < 0
    assert f((0.5, 0.5, 1), (0, 0, -1)) == 1.0                    # u + v == 1 is inside
    assert f((0.25, 0.25, 1), (0, 0, -1), tmin=1.0) == INF        # t <= max(eps, tMin) rejected
    assert f((0.25, 0.25, 1), (0, 0, -1), hitt=1.0) == INF        # t >= hit.t rejected
    assert f((0.25, 0.25, 1), (0, 0, -1), eps=2.0) == INF         # |det| = 1 < eps: the epsilon is scale dependent (:485)
    assert f((0.25, 0.25, 1), (1, 0, 0)) == INF                   # parallel
    out = np.zeros(6)
    L.orc_intersect_triangle(_p(tri), _p(v3(0.25, 0.5, 3)), _p(v3(0, 0, -1)), 0.0, INF, 1e-6, _p(out))
    assert out.tolist() == [0, 0, 1, 0.25, 0.5, 0]                # flat normal normalize(e1 x e2), p = v0 + u e1 + v e2


def test_sphere_second_root_rules():
    L = oracle_lib.lib()
    c = v3(0, 0, 0)
    sph = lambda mode, o, d, tmin, tmax, eps=1e-6: L.orc_sphere(mode, _p(c), 1.0, _p(v3(*o)), _p(v3(*d)), tmin, tmax, eps)
    assert sph(0, (0, 0, -3), (0, 0, 1), 0.0, INF) == 2.0
    assert sph(0, (0, 0, 0), (0, 0, 1), 0.0, INF) == 1.0          # inside: second root
    assert sph(0, (0, 0, -3), (0, 0, 2), 0.0, INF) == 1.0         # direction not unit: a = d.d (RTContext.swift:515)
    # closest-hit switches to the far root only if t < eps (:522) -> near root 2 <= tMin 2.5 rejects the sphere
    assert sph(0, (0, 0, -3), (0, 0, 1), 2.5, INF) == INF
    # shadow test switches if t < max(eps, tMin) (:860) -> far root 4 is inside (2.5, 10): occluded
    assert sph(1, (0, 0, -3), (0, 0, 1), 2.5, 10.0) == 1.0
    assert sph(1, (0, 0, -3), (0, 0, 1), 0.0, 1.5) == 0.0         # tMax before the sphere


def test_plane_rules():
    L = oracle_lib.lib()
    c, n = v3(0, 0, 0), v3(0, 1, 0)
    pl = lambda mode, o, d, tmin, tmax, eps=1e-6: L.orc_plane(mode, _p(c), _p(n), _p(v3(*o)), _p(v3(*d)), tmin, tmax, eps)
    assert pl(0, (0, 2, 0), (0, -1, 0), 0.0, INF) == 2.0
    assert pl(0, (0, 2, 0), (1, 0, 0), 0.0, INF) == INF           # |n.d| < eps
    assert pl(0, (0, 2, 0), (0, 1, 0), 0.0, INF) == INF           # behind
    assert pl(1, (0, 2, 0), (0, -1, 0), 0.0, 1.5) == 0.0 and pl(1, (0, 2, 0), (0, -1, 0), 0.0, 2.5) == 1.0
