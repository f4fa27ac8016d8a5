"""ctypes binding of the CPU oracle (oracle/liboracle.so).  Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import struct
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
CPLY_SO = os.path.join(ROOT, "oracle", "_ref", "libcply.so")


class OrcInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("triangles", "spheres", "planes", "instances", "cameras", "materials",
                                         "pointLights", "areaLights", "blasNodesTotal", "tlasNodes", "maxBuildDepth")]
    _fields_ += [("plyViaReference", C.c_int32), ("buildMs", C.c_double)]


class OrcCameraInfo(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("numSamples", C.c_int32), ("sqrtSamples", C.c_int32)]


class OrcCounters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("raysPrimary", "raysShadow", "raysSecondary", "nodeVisits", "boxTests",
                                         "triTests", "sphereTests", "planeTests", "instEntries", "maxStack")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


PRIMARY_HIT_DTYPE = np.dtype([("inst", "<i4"), ("kind", "<i4"), ("prim", "<i8"), ("t", "<f8")])

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            raise RuntimeError(f"{ORACLE_SO} missing: run `make -C oracle` (or __graft_entry__.build())")
        L = C.CDLL(ORACLE_SO)
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_load.restype = C.c_void_p
        L.orc_scene_load.argtypes = [C.c_char_p, C.c_char_p]
        L.orc_scene_free.argtypes = [C.c_void_p]
        L.orc_scene_info.argtypes = [C.c_void_p, C.POINTER(OrcInfo)]
        L.orc_camera_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(OrcCameraInfo)]
        L.orc_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.POINTER(OrcCounters), C.POINTER(C.c_double)]
        L.orc_get_triangles.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_get_instances.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.orc_pcg32.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        L.orc_pixel_seed.restype = C.c_uint64
        L.orc_pixel_seed.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_jitter_table.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_fresnel_dielectric.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.orc_fresnel_conductor.restype = C.c_double
        L.orc_fresnel_conductor.argtypes = [C.c_double] * 3
        L.orc_orthonormal_basis.argtypes = [C.c_void_p] * 3
        L.orc_hit_aabb.restype = C.c_double
        L.orc_hit_aabb.argtypes = [C.c_void_p] * 4 + [C.c_double]
        L.orc_intersect_triangle.restype = C.c_double
        L.orc_intersect_triangle.argtypes = [C.c_void_p] * 3 + [C.c_double] * 3 + [C.c_void_p]
        L.orc_sphere.restype = C.c_double
        L.orc_sphere.argtypes = [C.c_int, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.orc_plane.restype = C.c_double
        L.orc_plane.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.orc_bvh_build.restype = C.c_int64
        L.orc_bvh_build.argtypes = [C.c_int64] + [C.c_void_p] * 5
        L.orc_aabb_transformed.argtypes = [C.c_void_p] * 3
        L.orc_ply_load.argtypes = [C.c_char_p, C.c_int, C.c_char_p] + [C.POINTER(C.c_int64)] * 3 + [C.c_void_p] * 3
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def d3(x):
    return np.ascontiguousarray(x, dtype=np.float64)


class OracleScene:
    """RTContext(scene:) + Renderer of the CPU oracle for one .rtscene file."""

    def __init__(self, path: str, use_reference_cply: bool = True):
        L = lib()
        cply = CPLY_SO.encode() if (use_reference_cply and os.path.exists(CPLY_SO)) else None
        self.h = L.orc_scene_load(path.encode(), cply)
        if not self.h:
            raise RuntimeError("oracle: " + L.orc_last_error().decode())

    def close(self):
        if self.h:
            lib().orc_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> OrcInfo:
        i = OrcInfo()
        lib().orc_scene_info(self.h, C.byref(i))
        return i

    def camera(self, idx=0) -> OrcCameraInfo:
        c = OrcCameraInfo()
        if lib().orc_camera_info(self.h, idx, C.byref(c)) != 0:
            raise IndexError(lib().orc_last_error().decode())
        return c

    def render(self, camera=0, threads=0, want_linear=False, want_ids=False, chunk_stride=1, chunk_offset=0):
        """Returns dict(rgba uint8 (H,W,4), linear?, ids?, counters, ms)."""
        c = self.camera(camera)
        W, H, n = c.width, c.height, c.sqrtSamples
        rgba = np.zeros((H, W, 4), np.uint8)
        linear = np.zeros((H, W, 3), np.float64) if want_linear else None
        ids = np.zeros((H, W, n * n), PRIMARY_HIT_DTYPE) if want_ids else None
        cnt = OrcCounters()
        ms = C.c_double(0)
        rc = lib().orc_render(self.h, camera, threads, chunk_stride, chunk_offset, _p(linear), _p(rgba), _p(ids),
                              C.byref(cnt), C.byref(ms))
        if rc != 0:
            raise RuntimeError("oracle render: " + lib().orc_last_error().decode())
        return {"rgba": rgba, "linear": linear, "ids": ids, "counters": cnt.as_dict(), "ms": ms.value}

    def triangles(self):
        n = self.info().triangles
        v = np.zeros((n, 3, 3)); nr = np.zeros((n, 3, 3))
        lib().orc_get_triangles(self.h, _p(v), _p(nr))
        return v, nr

    def instances(self):
        n = self.info().instances
        out = dict(l2w=np.zeros((n, 16)), w2l=np.zeros((n, 16)), nmat=np.zeros((n, 9)), bounds=np.zeros((n, 6)),
                   meta=np.zeros((n, 4), np.int64))
        lib().orc_get_instances(self.h, _p(out["l2w"]), _p(out["w2l"]), _p(out["nmat"]), _p(out["bounds"]), _p(out["meta"]))
        return out


def write_png(path: str, rgba: np.ndarray) -> None:
    """Minimal RGB PNG writer (alpha dropped) for eyeballing renders."""
    h, w = rgba.shape[:2]
    raw = b"".join(b"\x00" + rgba[y, :, :3].tobytes() for y in range(h))

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))
