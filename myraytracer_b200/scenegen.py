"""Seeded synthetic scenes for the five BASELINE.json configs.

The reference ships no scene files, fixtures or PLY meshes, and
there is no network, so every input is generated here:

* ``Scene`` & friends mirror the field set of the reference's (absent)
  ParsingKit model that the render path reads; the C++ side of
  the same model is ``host/scene_model.hpp``.
* ``write_rtscene`` serialises a Scene into the ``.rtscene`` container read by
  ``pk::loadSceneBlob``.
* ``write_ply`` writes ``binary_little_endian 1.0`` PLY (float x/y/z, ``list
  uchar int vertex_indices``), the layout the reference's loader takes through
  its no-triangulation branch (Helpers/PLYReader.swift:186-198).
* ``config_c1`` … ``config_c5`` build the five configs; all randomness comes from
  splitmix64 of the stated seed.
"""
from __future__ import annotations

import math
import os
import struct
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

MAT_PLAIN, MAT_MIRROR, MAT_DIELECTRIC, MAT_CONDUCTOR = 0, 1, 2, 3
OBJ_SPHERE, OBJ_PLANE, OBJ_TRIANGLE, OBJ_MESH, OBJ_MESH_INSTANCE = 0, 1, 2, 3, 4

IDENTITY = (1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0)


# --------------------------------------------------------------------------
# model (mirror of host/scene_model.hpp)
# --------------------------------------------------------------------------
@dataclass
class Camera:
    id: int = 1
    imageName: str = "out.png"
    width: int = 800
    height: int = 800
    numSamples: int = 1
    apertureSize: float = 0.0
    focusDistance: float = 0.0
    position: Sequence[float] = (0.0, 0.0, 0.0)
    gaze: Sequence[float] = (0.0, 0.0, -1.0)
    gazePoint: Sequence[float] = (0.0, 0.0, -1.0)
    up: Sequence[float] = (0.0, 1.0, 0.0)
    nearPlane: Sequence[float] = (-1.0, 1.0, -1.0, 1.0)
    nearDistance: float = 1.0
    fovy: Optional[float] = None
    isLookAt: bool = False


@dataclass
class Material:
    id: int = 1
    type: int = MAT_PLAIN
    ambient: Sequence[float] = (1.0, 1.0, 1.0)
    diffuse: Sequence[float] = (1.0, 1.0, 1.0)
    specular: Sequence[float] = (0.0, 0.0, 0.0)
    mirror: Sequence[float] = (0.0, 0.0, 0.0)
    absorption: Sequence[float] = (0.0, 0.0, 0.0)
    phong: float = 1.0
    ior: float = 0.0
    absorptionIndex: float = 0.0
    roughness: float = 0.0


@dataclass
class PointLight:
    position: Sequence[float]
    intensity: Sequence[float]


@dataclass
class AreaLight:
    position: Sequence[float]
    normal: Sequence[float]
    size: float
    radiance: Sequence[float]


@dataclass
class SceneObject:
    kind: int
    id: int = 0
    material: Optional[int] = None  # numeric material id string; None = nil / ""
    transform: Sequence[float] = IDENTITY  # column-major local-to-world
    centerIdx: int = 0
    radius: float = 0.0
    normal: Sequence[float] = (0.0, 0.0, 0.0)
    indices: Sequence[int] = (0, 0, 0)
    smooth: bool = False
    motionBlur: Sequence[float] = (0.0, 0.0, 0.0)
    plyPath: str = ""
    faces: Optional[np.ndarray] = None  # int32, 1-based into vertexData
    baseMeshID: int = 0


@dataclass
class Scene:
    intersectionTestEpsilon: float = 1e-6
    shadowRayEpsilon: float = 1e-3
    maxRecursionDepth: int = 0
    backgroundColor: Sequence[float] = (0.0, 0.0, 0.0)
    ambient: Sequence[float] = (0.0, 0.0, 0.0)
    cameras: List[Camera] = field(default_factory=list)
    materials: List[Material] = field(default_factory=list)
    pointLights: List[PointLight] = field(default_factory=list)
    areaLights: List[AreaLight] = field(default_factory=list)
    vertexData: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float64))
    objects: List[SceneObject] = field(default_factory=list)


# --------------------------------------------------------------------------
# serialisation
# --------------------------------------------------------------------------
def _s(b: bytes) -> bytes:
    return struct.pack("<I", len(b)) + b


def _v3(v) -> bytes:
    return struct.pack("<3d", float(v[0]), float(v[1]), float(v[2]))


def write_rtscene(scene: Scene, path: str) -> None:
    """Serialise ``scene`` (layout documented in host/scene_model.hpp)."""
    out = [b"RTSCENE1"]
    out.append(struct.pack("<ddi", scene.intersectionTestEpsilon, scene.shadowRayEpsilon, scene.maxRecursionDepth))
    out.append(_v3(scene.backgroundColor) + _v3(scene.ambient))
    out.append(struct.pack("<I", len(scene.cameras)))
    for c in scene.cameras:
        out.append(struct.pack("<i", c.id) + _s(c.imageName.encode()))
        out.append(struct.pack("<iiidd", c.width, c.height, c.numSamples, c.apertureSize, c.focusDistance))
        out.append(_v3(c.position) + _v3(c.gaze) + _v3(c.gazePoint) + _v3(c.up))
        out.append(struct.pack("<4d", *[float(x) for x in c.nearPlane]))
        out.append(struct.pack("<didi", c.nearDistance, 1 if c.fovy is not None else 0,
                               float(c.fovy) if c.fovy is not None else 0.0, 1 if c.isLookAt else 0))
    out.append(struct.pack("<I", len(scene.materials)))
    for m in scene.materials:
        out.append(struct.pack("<ii", m.id, m.type))
        out.append(_v3(m.ambient) + _v3(m.diffuse) + _v3(m.specular) + _v3(m.mirror) + _v3(m.absorption))
        out.append(struct.pack("<4d", m.phong, m.ior, m.absorptionIndex, m.roughness))
    out.append(struct.pack("<I", len(scene.pointLights)))
    for l in scene.pointLights:
        out.append(_v3(l.position) + _v3(l.intensity))
    out.append(struct.pack("<I", len(scene.areaLights)))
    for a in scene.areaLights:
        out.append(_v3(a.position) + _v3(a.normal) + struct.pack("<d", a.size) + _v3(a.radiance))
    vd = np.ascontiguousarray(scene.vertexData, dtype="<f8").reshape(-1, 3)
    out.append(struct.pack("<I", vd.shape[0]) + vd.tobytes())
    out.append(struct.pack("<I", len(scene.objects)))
    for o in scene.objects:
        out.append(struct.pack("<iiii", o.kind, o.id, 1 if o.material is not None else 0,
                               o.material if o.material is not None else -1))
        out.append(struct.pack("<16d", *[float(x) for x in o.transform]))
        out.append(struct.pack("<id", o.centerIdx, o.radius) + _v3(o.normal))
        out.append(struct.pack("<3i", *[int(x) for x in o.indices]))
        out.append(struct.pack("<i", 1 if o.smooth else 0) + _v3(o.motionBlur))
        out.append(_s(o.plyPath.encode()))
        faces = np.zeros(0, "<i4") if o.faces is None else np.ascontiguousarray(o.faces, dtype="<i4").reshape(-1)
        out.append(struct.pack("<I", faces.size) + faces.tobytes())
        out.append(struct.pack("<i", o.baseMeshID))
    with open(path, "wb") as f:
        f.write(b"".join(out))


def write_ply(path: str, positions: np.ndarray, tris: np.ndarray, normals: Optional[np.ndarray] = None) -> None:
    """binary_little_endian PLY: float x,y,z [nx,ny,nz]; list uchar int vertex_indices (0-based)."""
    pos = np.ascontiguousarray(positions, dtype="<f4").reshape(-1, 3)
    tri = np.ascontiguousarray(tris, dtype="<i4").reshape(-1, 3)
    hdr = ["ply", "format binary_little_endian 1.0", f"element vertex {pos.shape[0]}",
           "property float x", "property float y", "property float z"]
    if normals is not None:
        hdr += ["property float nx", "property float ny", "property float nz"]
        pos = np.concatenate([pos, np.ascontiguousarray(normals, dtype="<f4").reshape(-1, 3)], axis=1)
    hdr += [f"element face {tri.shape[0]}", "property list uchar int vertex_indices", "end_header"]
    rec = np.zeros(tri.shape[0], dtype=[("n", "u1"), ("i", "<i4", (3,))])
    rec["n"] = 3
    rec["i"] = tri
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode())
        f.write(np.ascontiguousarray(pos).tobytes())
        f.write(rec.tobytes())


# --------------------------------------------------------------------------
# seeded randomness and transforms
# --------------------------------------------------------------------------
class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def uniform(self, lo: float = 0.0, hi: float = 1.0) -> float:
        return lo + (hi - lo) * ((self.next() >> 11) * (1.0 / (1 << 53)))


def mat_translate(t) -> np.ndarray:
    m = np.eye(4)
    m[:3, 3] = t
    return m


def mat_scale(s) -> np.ndarray:
    return np.diag([s[0], s[1], s[2], 1.0])


def mat_rotate(axis, deg: float) -> np.ndarray:
    a = np.asarray(axis, float)
    a = a / np.linalg.norm(a)
    c, s = math.cos(math.radians(deg)), math.sin(math.radians(deg))
    x, y, z = a
    r = np.array([[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s, 0],
                  [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s, 0],
                  [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c), 0],
                  [0, 0, 0, 1.0]])
    return r


def colmajor(m: np.ndarray) -> tuple:
    return tuple(float(x) for x in np.asarray(m, float).T.reshape(-1))


# --------------------------------------------------------------------------
# meshes
# --------------------------------------------------------------------------
def terrain_mesh(n: int, seed: int, half: float = 10.0, amp: float = 1.0):
    """n×n float32 vertices on [-half, half]² (x,z), y = Σ5 seeded sine/cosine octaves.

    Returns (positions float32 (n², 3), triangles int32 (2(n-1)², 3)), CCW seen from +y.
    188 → 69,938 triangles (C2); 708 → 999,698 triangles (C4).
    """
    rng = SplitMix64(seed)
    xs = np.linspace(-half, half, n)
    X, Z = np.meshgrid(xs, xs, indexing="xy")
    Y = np.zeros_like(X)
    for o in range(5):
        f = (0.25 * (1.9 ** o)) * (1.0 + 0.2 * rng.uniform())
        a = amp * (0.55 ** o)
        th = rng.uniform(0.0, 2.0 * math.pi)
        ph1, ph2 = rng.uniform(0.0, 2.0 * math.pi), rng.uniform(0.0, 2.0 * math.pi)
        U = X * math.cos(th) + Z * math.sin(th)
        V = -X * math.sin(th) + Z * math.cos(th)
        Y += a * np.sin(f * U + ph1) * np.cos(f * V + ph2)
    pos = np.stack([X, Y, Z], axis=-1).reshape(-1, 3).astype(np.float32)
    i = np.arange(n - 1)
    I, J = np.meshgrid(i, i, indexing="xy")  # I: column (x), J: row (z)
    v00 = (J * n + I).reshape(-1)
    v10 = v00 + 1
    v01 = v00 + n
    v11 = v01 + 1
    # +y facing: (v00, v01, v10) and (v10, v01, v11)
    t0 = np.stack([v00, v01, v10], axis=-1)
    t1 = np.stack([v10, v01, v11], axis=-1)
    tris = np.stack([t0, t1], axis=1).reshape(-1, 3).astype(np.int32)
    return pos, tris


def icosphere_mesh(level: int, radius: float = 1.0, seed: Optional[int] = None, disp: float = 0.0):
    """Subdivided icosahedron: 20·4^level triangles (level 4 → 5,120), outward facing."""
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [np.asarray(p, float) / np.linalg.norm(p) for p in v]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2),
             (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11),
             (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(level):
        cache = {}
        nf = []

        def mid(a, b):
            k = (a, b) if a < b else (b, a)
            if k not in cache:
                m = verts[a] + verts[b]
                verts.append(m / np.linalg.norm(m))
                cache[k] = len(verts) - 1
            return cache[k]

        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = nf
    P = np.asarray(verts, float)
    if seed is not None and disp > 0:
        rng = SplitMix64(seed)
        ph = [rng.uniform(0, 2 * math.pi) for _ in range(3)]
        P = P * (1.0 + disp * np.sin(5 * P[:, :1] + ph[0]) * np.sin(4 * P[:, 1:2] + ph[1]) * np.sin(6 * P[:, 2:3] + ph[2]))
    return (P * radius).astype(np.float32), np.asarray(faces, np.int32)


# --------------------------------------------------------------------------
# configs
# --------------------------------------------------------------------------
def _plain(id, rgb, spec=(0.3, 0.3, 0.3), phong=50.0):
    return Material(id=id, type=MAT_PLAIN, ambient=rgb, diffuse=rgb, specular=spec, phong=phong, ior=0.0)


def config_c1(out_dir: str, width: int = 800, height: int = 800, samples: int = 1) -> str:
    """C1: simple scene — ground quad (2 lone triangles), 1 lone triangle, 3 spheres, 2 point lights."""
    rng = SplitMix64(1)
    sc = Scene(maxRecursionDepth=0, backgroundColor=(0, 0, 0), ambient=(25, 25, 25),
               shadowRayEpsilon=1e-3, intersectionTestEpsilon=1e-6)
    sc.materials = [
        _plain(1, (0.55, 0.55, 0.5), (0.1, 0.1, 0.1), 10.0),
        _plain(2, (0.8, 0.25, 0.2), (0.5, 0.5, 0.5), 50.0),
        _plain(3, (0.2, 0.7, 0.3), (0.6, 0.6, 0.6), 100.0),
        _plain(4, (0.25, 0.35, 0.85), (0.4, 0.4, 0.4), 50.0),
        _plain(5, (0.85, 0.75, 0.2), (0.2, 0.2, 0.2), 10.0),
    ]
    j = lambda: rng.uniform(-0.2, 0.2)
    verts = [(-10, 0, -10), (10, 0, -10), (10, 0, 10), (-10, 0, 10),          # 1-4 ground
             (-4 + j(), 0.2, -3 + j()), (-1 + j(), 0.4, -5 + j()), (-2.5 + j(), 4 + j(), -4 + j()),  # 5-7 lone triangle
             (-2.0 + j(), 1.0, 1.0 + j()), (1.5 + j(), 1.5, -1.0 + j()), (3.5 + j(), 0.75, 2.5 + j())]  # 8-10 centres
    sc.vertexData = np.asarray(verts, np.float64)
    sc.objects = [
        SceneObject(OBJ_TRIANGLE, id=1, material=1, indices=(1, 3, 2)),
        SceneObject(OBJ_TRIANGLE, id=2, material=1, indices=(1, 4, 3)),
        SceneObject(OBJ_TRIANGLE, id=3, material=5, indices=(5, 6, 7)),
        SceneObject(OBJ_SPHERE, id=4, material=2, centerIdx=8, radius=1.0),
        SceneObject(OBJ_SPHERE, id=5, material=3, centerIdx=9, radius=1.5),
        SceneObject(OBJ_SPHERE, id=6, material=4, centerIdx=10, radius=0.75),
    ]
    sc.pointLights = [PointLight((6.0, 10.0, 6.0), (24000, 24000, 24000)),
                      PointLight((-7.0, 10.0, 3.0), (16000, 14000, 12000))]
    sc.cameras = [Camera(id=1, imageName="c1.png", width=width, height=height, numSamples=samples,
                         position=(0.5, 5.0, 13.0), gazePoint=(0.0, 1.0, 0.0), up=(0, 1, 0),
                         nearDistance=1.0, fovy=45.0, isLookAt=True)]
    path = os.path.join(out_dir, "c1.rtscene")
    write_rtscene(sc, path)
    return path


def _terrain_scene(out_dir: str, name: str, seed: int, n: int, width: int, height: int, samples: int,
                   lights: int) -> str:
    pos, tris = terrain_mesh(n, seed, amp=1.5)
    write_ply(os.path.join(out_dir, f"{name}_terrain.ply"), pos, tris)
    sc = Scene(maxRecursionDepth=0, backgroundColor=(10, 20, 40), ambient=(20, 20, 20))
    sc.materials = [_plain(1, (0.6, 0.55, 0.4), (0.25, 0.25, 0.25), 30.0)]
    sc.objects = [SceneObject(OBJ_MESH, id=1, material=1, smooth=True, plyPath=f"{name}_terrain.ply")]
    sc.pointLights = [PointLight((8.0, 9.0, 8.0), (22000, 21000, 19000))]
    if lights > 1:
        sc.pointLights.append(PointLight((-9.0, 7.0, -2.0), (9000, 10000, 12000)))
    aspect = width / height
    sc.cameras = [Camera(id=1, imageName=f"{name}.png", width=width, height=height, numSamples=samples,
                         position=(0.0, 12.0, 19.0), gaze=(0.0, -0.68, -1.0), up=(0, 1, 0),
                         nearPlane=(-0.42 * aspect, 0.42 * aspect, -0.42, 0.42), nearDistance=1.0)]
    path = os.path.join(out_dir, f"{name}.rtscene")
    write_rtscene(sc, path)
    return path


def config_c2(out_dir: str, n: int = 188, width: int = 1024, height: int = 1024, samples: int = 1) -> str:
    """C2: height-field terrain PLY, 188² vertices → 69,938 triangles, smooth, 1 point light."""
    return _terrain_scene(out_dir, "c2", 2, n, width, height, samples, lights=1)


def config_c4(out_dir: str, n: int = 708, width: int = 1920, height: int = 1080, samples: int = 16) -> str:
    """C4: terrain PLY, 708² vertices → 999,698 triangles, smooth, 2 point lights, 16 spp."""
    return _terrain_scene(out_dir, "c4", 4, n, width, height, samples, lights=2)


def config_c3(out_dir: str, width: int = 1024, height: int = 1024, samples: int = 4, depth: int = 6,
              ico_level: int = 4, roughness: float = 0.0, aperture: float = 0.0, focus: float = 0.0,
              area_lights: int = 0, inf_light: bool = False) -> str:
    """C3: closed box (10 lone triangles), mirror / dielectric / conductor spheres, dielectric PLY icosphere.
    Beyond BASELINE.json's C3: `roughness` != 0 makes the mirror, conductor and dielectric
    materials glossy (random draws inside `trace`, Object+Extension.swift:191-197,210-216,259-265); `aperture` and
    `focus` > 0 give the camera a thin lens (:325-338); `area_lights` (0..3) adds square area lights sampled
    through the per-chunk jitter index (:145-186) and drops the second point light; `inf_light` gives the last light
    (area light if there is one) an infinite intensity, so every level it reaches hits the non-finite guard (:277-280)."""
    rng = SplitMix64(3)
    pos, tris = icosphere_mesh(ico_level, radius=1.0)
    write_ply(os.path.join(out_dir, "c3_ico.ply"), pos, tris)
    sc = Scene(maxRecursionDepth=depth, backgroundColor=(5, 5, 8), ambient=(18, 18, 18))
    sc.materials = [
        _plain(1, (0.7, 0.7, 0.7), (0.0, 0.0, 0.0), 1.0),      # floor/ceiling/back
        _plain(2, (0.75, 0.2, 0.2), (0.0, 0.0, 0.0), 1.0),     # left wall
        _plain(3, (0.2, 0.3, 0.75), (0.0, 0.0, 0.0), 1.0),     # right wall
        Material(id=4, type=MAT_MIRROR, ambient=(0, 0, 0), diffuse=(0.05, 0.05, 0.05), specular=(0.3, 0.3, 0.3),
                 mirror=(0.9, 0.9, 0.9), phong=100.0, ior=0.0),
        Material(id=5, type=MAT_DIELECTRIC, ambient=(0, 0, 0), diffuse=(0.02, 0.02, 0.02), specular=(0.2, 0.2, 0.2),
                 mirror=(1, 1, 1), absorption=(0.01, 0.05, 0.05), phong=200.0, ior=1.5),
        Material(id=6, type=MAT_CONDUCTOR, ambient=(0, 0, 0), diffuse=(0.05, 0.04, 0.02), specular=(0.3, 0.3, 0.3),
                 mirror=(1.0, 0.86, 0.57), phong=80.0, ior=0.37, absorptionIndex=2.82),
        Material(id=7, type=MAT_DIELECTRIC, ambient=(0, 0, 0), diffuse=(0.02, 0.02, 0.02), specular=(0.2, 0.2, 0.2),
                 mirror=(1, 1, 1), absorption=(0.08, 0.02, 0.01), phong=200.0, ior=1.33),
    ]
    H = 5.0
    box = [(-H, 0, -H), (H, 0, -H), (H, 0, H), (-H, 0, H), (-H, 2 * H, -H), (H, 2 * H, -H), (H, 2 * H, H), (-H, 2 * H, H)]
    j = lambda: rng.uniform(-0.15, 0.15)
    centres = [(-2.6 + j(), 1.3, -1.5 + j()), (0.4 + j(), 1.1, 1.6 + j()), (0.0, 0.0, 0.0)]
    sc.vertexData = np.asarray(box + centres, np.float64)
    quad = lambda a, b, c, d, m, i: [SceneObject(OBJ_TRIANGLE, id=i, material=m, indices=(a, b, c)),
                                     SceneObject(OBJ_TRIANGLE, id=i + 1, material=m, indices=(a, c, d))]
    objs = []
    objs += quad(1, 4, 3, 2, 1, 1)     # floor (+y normal)
    objs += quad(5, 6, 7, 8, 1, 3)     # ceiling (-y)
    objs += quad(1, 2, 6, 5, 1, 5)     # back wall z=-H (+z)
    objs += quad(1, 5, 8, 4, 2, 7)     # left x=-H (+x)
    objs += quad(2, 3, 7, 6, 3, 9)     # right x=+H (-x)
    objs.append(SceneObject(OBJ_SPHERE, id=11, material=4, centerIdx=9, radius=1.3))
    objs.append(SceneObject(OBJ_SPHERE, id=12, material=5, centerIdx=10, radius=1.1))
    # conductor sphere: unit sphere at origin placed by translate·scale (exercises world<->local and normal matrix)
    Mc = mat_translate((2.7 + j(), 1.0, -1.2 + j())) @ mat_scale((1.0, 1.0, 1.0))
    objs.append(SceneObject(OBJ_SPHERE, id=13, material=6, centerIdx=11, radius=1.0, transform=colmajor(Mc)))
    Mi = mat_translate((-1.0 + j(), 3.2, 0.2 + j())) @ mat_rotate((0.3, 1.0, 0.2), 25.0) @ mat_scale((0.9, 0.7, 0.9))
    objs.append(SceneObject(OBJ_MESH, id=14, material=7, smooth=False, plyPath="c3_ico.ply", transform=colmajor(Mi)))
    sc.objects = objs
    sc.pointLights = [PointLight((0.0, 9.0, 2.0), (6000, 6000, 6000)), PointLight((-3.0, 6.0, 4.0), (2600, 2400, 2200))]
    if area_lights > 0:
        sc.pointLights = sc.pointLights[:1]
        sc.areaLights = [AreaLight((0.5, 9.9, 0.5), (0.0, -1.0, 0.0), 3.0, (1500, 1450, 1300)),
                         AreaLight((-4.9, 5.0, 1.0), (1.0, -0.3, 0.4), 2.0, (900, 1100, 1400)),
                         AreaLight((2.0, 6.0, 4.5), (-0.2, -0.5, -1.0), 1.5, (1200, 800, 600))][:area_lights]
    if inf_light:
        inf = (float("inf"),) * 3
        if area_lights > 0:
            a = sc.areaLights[-1]
            sc.areaLights[-1] = AreaLight(a.position, a.normal, a.size, inf)
        else:
            sc.pointLights[-1] = PointLight(sc.pointLights[-1].position, inf)
    for m in sc.materials:
        if m.type != MAT_PLAIN:
            m.roughness = roughness
    sc.cameras = [Camera(id=1, imageName="c3.png", width=width, height=height, numSamples=samples,
                         position=(0.0, 4.0, 14.5), gazePoint=(0.0, 3.0, 0.0), up=(0, 1, 0),
                         nearDistance=1.0, fovy=40.0, isLookAt=True, apertureSize=aperture, focusDistance=focus)]
    path = os.path.join(out_dir, "c3.rtscene")
    write_rtscene(sc, path)
    return path


def config_c5(out_dir: str, n: int = 708, width: int = 3840, height: int = 2160, samples: int = 64,
              meshes: int = 10, depth: int = 2) -> str:
    """C5: `meshes` distinct terrains (own PLY, own non-uniform L2W) + 3 reflective spheres, depth 2."""
    rng = SplitMix64(5)
    sc = Scene(maxRecursionDepth=depth, backgroundColor=(10, 20, 40), ambient=(20, 20, 20))
    sc.materials = [_plain(1, (0.6, 0.55, 0.4), (0.25, 0.25, 0.25), 30.0),
                    _plain(2, (0.35, 0.6, 0.35), (0.2, 0.2, 0.2), 20.0),
                    Material(id=3, type=MAT_MIRROR, ambient=(0, 0, 0), diffuse=(0.05, 0.05, 0.05),
                             specular=(0.4, 0.4, 0.4), mirror=(0.85, 0.85, 0.85), phong=100.0),
                    Material(id=4, type=MAT_CONDUCTOR, ambient=(0, 0, 0), diffuse=(0.05, 0.04, 0.02),
                             specular=(0.3, 0.3, 0.3), mirror=(1.0, 0.86, 0.57), phong=80.0, ior=0.37,
                             absorptionIndex=2.82)]
    objs = []
    cols = (meshes + 1) // 2
    for k in range(meshes):
        pos, tris = terrain_mesh(n, 500 + k)
        write_ply(os.path.join(out_dir, f"c5_terrain{k}.ply"), pos, tris)
        cx = (k % cols - (cols - 1) / 2.0) * 21.0
        cz = (k // cols) * -21.0
        M = (mat_translate((cx, rng.uniform(-0.5, 0.5), cz)) @ mat_rotate((0, 1, 0), rng.uniform(-20, 20))
             @ mat_rotate((1, 0, 0), rng.uniform(-4, 4)) @ mat_scale((1.05, rng.uniform(0.6, 1.6), 1.05)))
        objs.append(SceneObject(OBJ_MESH, id=k + 1, material=1 + (k % 2), smooth=True,
                                plyPath=f"c5_terrain{k}.ply", transform=colmajor(M)))
    sc.vertexData = np.asarray([(-8.0, 4.5, -4.0), (0.0, 5.0, -8.0), (9.0, 4.0, -3.0)], np.float64)
    objs.append(SceneObject(OBJ_SPHERE, id=101, material=3, centerIdx=1, radius=2.2))
    objs.append(SceneObject(OBJ_SPHERE, id=102, material=4, centerIdx=2, radius=2.6))
    objs.append(SceneObject(OBJ_SPHERE, id=103, material=3, centerIdx=3, radius=1.8))
    sc.objects = objs
    sc.pointLights = [PointLight((20.0, 30.0, 15.0), (260000, 250000, 230000)),
                      PointLight((-30.0, 25.0, -20.0), (120000, 130000, 150000))]
    aspect = width / height
    sc.cameras = [Camera(id=1, imageName="c5.png", width=width, height=height, numSamples=samples,
                         position=(0.0, 22.0, 42.0), gaze=(0.0, -0.45, -1.0), up=(0, 1, 0),
                         nearPlane=(-0.5 * aspect, 0.5 * aspect, -0.5, 0.5), nearDistance=1.0)]
    path = os.path.join(out_dir, "c5.rtscene")
    write_rtscene(sc, path)
    return path


def config_instances(out_dir: str, width: int = 256, height: int = 192, samples: int = 4, depth: int = 3,
                     mesh_motion: Sequence[float] = (0.0, 0.0, 0.0), instance_motion: float = 0.0, frame: float = 0.0,
                     name: str = "instances") -> str:
    """Not one of BASELINE.json's configs: the object kinds they leave out.  An infinite ground plane (its 1e5 bounds,
    RTContext.swift:172-174, stretch the scene box), one PLY mesh and three MeshInstances of it sharing its BLAS
    (:384-401): one with a material override, one mirrored (negative determinant -> flipped normals, :691-698),
    one an instance of an instance; a dielectric sphere in front.
    Motion blur: `mesh_motion` is the base mesh's Mesh.motionBlur (local space, applied in the
    triangle test, boxes grown by it, RTContext.swift:283-289,480-481); `instance_motion` scales the
    MeshInstance.motionBlur vectors of the three instances (world space, applied to the ray origin before the
    instance transform while the instance boxes stay where they are, :437-455,658-661).
    Animation (8f row 4): `frame` moves, turns and scales the mesh, its instances and the sphere's placement —
    same objects, same BLASes, other transforms."""
    pos, tris = icosphere_mesh(2, radius=1.0, seed=7, disp=0.15)
    write_ply(os.path.join(out_dir, "inst_blob.ply"), pos, tris)
    sc = Scene(maxRecursionDepth=depth, backgroundColor=(12, 14, 20), ambient=(20, 20, 20))
    sc.materials = [_plain(1, (0.55, 0.55, 0.5), (0.1, 0.1, 0.1), 10.0),
                    _plain(2, (0.8, 0.3, 0.2), (0.5, 0.5, 0.5), 60.0),
                    _plain(3, (0.2, 0.5, 0.8), (0.4, 0.4, 0.4), 40.0),
                    Material(id=4, type=MAT_MIRROR, ambient=(0, 0, 0), diffuse=(0.05, 0.05, 0.05),
                             specular=(0.3, 0.3, 0.3), mirror=(0.85, 0.85, 0.85), phong=100.0),
                    Material(id=5, type=MAT_DIELECTRIC, ambient=(0, 0, 0), diffuse=(0.02, 0.02, 0.02), specular=(0.2, 0.2, 0.2),
                             mirror=(1, 1, 1), absorption=(0.02, 0.06, 0.04), phong=200.0, ior=1.5)]
    sc.vertexData = np.asarray([(0.0, 0.0, 0.0), (0.8, 1.1, 4.0)], np.float64)
    f = float(frame)
    Mm = mat_translate((-0.5 + 0.4 * f, 1.4, 0.3 * f)) @ mat_rotate((0.2, 1.0, 0.1), 20.0 + 35.0 * f) @ mat_scale((1.3, 1.3, 1.3))
    Ma = mat_translate((3.2 - 0.8 * f, 1.1 + 0.5 * f, -2.5)) @ mat_rotate((0.0, 0.0, 1.0), 50.0 * f) @ mat_scale((0.9, 1.1 + 0.2 * f, 0.9))
    Mb = mat_translate((-3.6, 1.2 + 0.6 * f, -1.5 + 1.2 * f)) @ mat_scale((-1.0, 1.0, 1.0))          # mirrored: det < 0
    Mc = mat_translate((0.5 + 1.5 * f, 3.6 - 0.7 * f, -3.0)) @ mat_rotate((1.0, 0.3, 0.0), -40.0 * f) @ mat_scale((0.6, 0.6, 0.6))
    Ms = mat_translate((-1.2 * f, 0.3 * f, 0.5 * f))                               # the sphere's object transform
    sc.objects = [
        SceneObject(OBJ_PLANE, id=1, material=1, centerIdx=1, normal=(0.0, 1.0, 0.0)),
        SceneObject(OBJ_SPHERE, id=2, material=5, centerIdx=2, radius=1.0, transform=colmajor(Ms)),
        SceneObject(OBJ_MESH, id=5, material=2, smooth=True, plyPath="inst_blob.ply", transform=colmajor(Mm),
                    motionBlur=tuple(mesh_motion)),
        SceneObject(OBJ_MESH_INSTANCE, id=6, material=4, baseMeshID=5, transform=colmajor(Ma),
                    motionBlur=(0.9 * instance_motion, 0.0, 0.3 * instance_motion)),
        SceneObject(OBJ_MESH_INSTANCE, id=7, material=None, baseMeshID=5, transform=colmajor(Mb),
                    motionBlur=(0.0, 0.7 * instance_motion, 0.0)),
        SceneObject(OBJ_MESH_INSTANCE, id=8, material=3, baseMeshID=6, transform=colmajor(Mc),
                    motionBlur=(-0.5 * instance_motion, -0.4 * instance_motion, 0.6 * instance_motion)),
    ]
    sc.pointLights = [PointLight((5.0, 10.0, 7.0), (22000, 22000, 22000)), PointLight((-6.0, 9.0, 2.0), (12000, 11000, 10000))]
    sc.cameras = [Camera(id=1, imageName="instances.png", width=width, height=height, numSamples=samples,
                         position=(0.5, 4.5, 12.0), gazePoint=(0.0, 1.2, 0.0), up=(0, 1, 0),
                         nearDistance=1.0, fovy=45.0, isLookAt=True)]
    path = os.path.join(out_dir, f"{name}.rtscene")
    write_rtscene(sc, path)
    return path


CONFIGS = {"c1": config_c1, "instances": config_instances, "c2": config_c2, "c3": config_c3, "c4": config_c4, "c5": config_c5}

if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description="write a synthetic config (.rtscene + PLY files)")
    ap.add_argument("config", choices=sorted(CONFIGS))
    ap.add_argument("out_dir")
    a = ap.parse_args()
    os.makedirs(a.out_dir, exist_ok=True)
    print(CONFIGS[a.config](a.out_dir))
