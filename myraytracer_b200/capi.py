"""ctypes mirror of include/rtcuda.h and include/rtcuda_host.h (the C ABI of librtcuda.so).

There is no CPU fallback: if the shared library is missing, loading fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# RTCUDA_LIB: developer override to A/B kernel build variants (always a CUDA build of this same source)
LIB_PATH = os.environ.get("RTCUDA_LIB") or os.path.join(HERE, "librtcuda.so")

RT_ABI_VERSION = 2
RT_IPC_HANDLE_BYTES = 64
RT_OK, RT_ERR_INVALID, RT_ERR_CUDA, RT_ERR_UNSUPPORTED, RT_ERR_CANCELLED = 0, -1, -2, -3, -4
RT_ERR_BAD_CAMERA, RT_ERR_NO_SCENE, RT_ERR_NO_RENDERER = -10, -20, -21
RT_GEO_MESH, RT_GEO_TRIANGLE, RT_GEO_SPHERE, RT_GEO_PLANE = 0, 1, 2, 3
RT_HIT_NONE, RT_HIT_TRIANGLE, RT_HIT_SPHERE, RT_HIT_PLANE = 0, 1, 2, 3

D3 = C.c_double * 3


class rt_material(C.Structure):
    _fields_ = [("type", C.c_int32), ("reserved", C.c_int32), ("ambient", D3), ("diffuse", D3), ("specular", D3),
                ("mirror", D3), ("absorption", D3), ("phong", C.c_double), ("ior", C.c_double),
                ("absorptionIndex", C.c_double), ("roughness", C.c_double)]


class rt_point_light(C.Structure):
    _fields_ = [("position", D3), ("intensity", D3)]


class rt_area_light(C.Structure):
    _fields_ = [("position", D3), ("normal", D3), ("size", C.c_double), ("radiance", D3)]


class rt_sphere(C.Structure):
    _fields_ = [("center", D3), ("radius", C.c_double)]


class rt_plane(C.Structure):
    _fields_ = [("center", D3), ("normal", D3)]


class rt_prim(C.Structure):
    _fields_ = [("type", C.c_int32), ("smooth", C.c_int32), ("primitiveIndex", C.c_int64), ("materialID", C.c_int64)]


class rt_blas(C.Structure):
    _fields_ = [("firstPrim", C.c_int64), ("primCount", C.c_int64)]


class rt_instance(C.Structure):
    _fields_ = [("blas", C.c_int32), ("hasMaterialOverride", C.c_int32), ("materialOverride", C.c_int64),
                ("localToWorld", C.c_double * 16), ("worldToLocal", C.c_double * 16), ("normalMatrix", C.c_double * 9),
                ("worldBounds", C.c_double * 6), ("instanceMotion", D3)]


class rt_scene_desc(C.Structure):
    _fields_ = [("abiVersion", C.c_uint32), ("reserved", C.c_uint32),
                ("nTriangles", C.c_int64), ("triVertices", C.POINTER(C.c_double)), ("triNormals", C.POINTER(C.c_double)),
                ("triMotionBlur", C.POINTER(C.c_double)),
                ("nSpheres", C.c_int64), ("spheres", C.POINTER(rt_sphere)),
                ("nPlanes", C.c_int64), ("planes", C.POINTER(rt_plane)),
                ("nPrims", C.c_int64), ("prims", C.POINTER(rt_prim)),
                ("nBlas", C.c_int64), ("blas", C.POINTER(rt_blas)),
                ("nInstances", C.c_int64), ("instances", C.POINTER(rt_instance)),
                ("nMaterials", C.c_int64), ("materials", C.POINTER(rt_material)),
                ("nPointLights", C.c_int64), ("pointLights", C.POINTER(rt_point_light)),
                ("nAreaLights", C.c_int64), ("areaLights", C.POINTER(rt_area_light)),
                ("ambient", D3), ("background", D3),
                ("intersectionTestEpsilon", C.c_double), ("shadowRayEpsilon", C.c_double),
                ("maxRecursionDepth", C.c_int32), ("buildMode", C.c_int32)]


class rt_camera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("numSamples", C.c_int32), ("isLookAt", C.c_int32),
                ("hasFovy", C.c_int32), ("reserved", C.c_int32),
                ("position", D3), ("gaze", D3), ("gazePoint", D3), ("up", D3), ("nearPlane", C.c_double * 4),
                ("nearDistance", C.c_double), ("fovy", C.c_double), ("apertureSize", C.c_double),
                ("focusDistance", C.c_double)]


rt_progress_fn = C.CFUNCTYPE(C.c_int, C.c_int32, C.c_int32, C.c_void_p)


class rt_render_params(C.Structure):
    _fields_ = [("shardIndex", C.c_int32), ("shardCount", C.c_int32), ("collectCounters", C.c_int32),
                ("kernelVariant", C.c_int32), ("progress", rt_progress_fn), ("progressUser", C.c_void_p),
                ("scheduleMode", C.c_int32), ("shardMode", C.c_int32)]


class rt_stats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("meshes", "triangles", "spheres", "planes", "rays", "milliseconds",
                                         "raysPrimary", "raysShadow", "raysSecondary")]
    _fields_ += [("deviceMs", C.c_double), ("kernelLaunches", C.c_int64)]
    _fields_ += [(n, C.c_int64) for n in ("boxTests", "triTests", "sphereTests", "planeTests", "instEntries", "maxStack",
                                         "smoothHits", "tilesScheduled")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class rt_scene_info(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("triangles", "spheres", "planes", "instances", "blasCount", "blasNodes",
                                         "tlasNodes", "maxLeafSize", "maxDepth", "deviceBytes")]
    _fields_ += [("buildMs", C.c_double), ("uploadMs", C.c_double), ("device", C.c_int32), ("deviceBuiltBlas", C.c_int32),
                 ("deviceCount", C.c_int32), ("wideNodes", C.c_int32), ("replicateMs", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class rt_ply_layout(C.Structure):
    _fields_ = [("vertexCount", C.c_int64), ("faceCount", C.c_int64), ("vertexStride", C.c_int32), ("posOffset", C.c_int32 * 3),
                ("normalOffset", C.c_int32 * 3), ("faceLeadBytes", C.c_int32), ("faceCountBytes", C.c_int32),
                ("faceIndexBytes", C.c_int32), ("faceTrailBytes", C.c_int32), ("verticesPerFace", C.c_int32), ("reserved", C.c_int32)]


class rt_ply_stats(C.Structure):
    _fields_ = [("triangles", C.c_int64), ("kernelLaunches", C.c_int64), ("deviceMs", C.c_double)]


RT_PLY_STATIC_MAX_VERTICES, RT_PLY_STATIC_MAX_TRIANGLES, RT_PLY_STATIC_SCRATCH_BYTES = 65536, 131072, 10 << 20

HIT_ID_DTYPE = np.dtype([("inst", "<i4"), ("kind", "<i4"), ("prim", "<i8"), ("t", "<f8")])

# every symbol include/rtcuda.h and include/rtcuda_host.h declare
EXPORTS = ["rt_abi_version", "rt_device_count", "rt_last_error", "rt_scene_create", "rt_scene_destroy",
           "rt_scene_create_multi", "rt_scene_device_count", "rt_host_alloc", "rt_host_free", "rth_engine_create_multi",
           "rt_scene_get_info", "rt_scene_get_schedule", "rt_scene_read_blas_nodes", "rt_scene_read_wide_nodes", "rt_scene_update_instances", "rt_ply_ingest", "rt_ply_ingest_static", "rt_render", "rt_render_device", "rt_render_finish",
           "rt_shared_buffer_create", "rt_shared_buffer_open", "rt_shared_buffer_close",
           "rth_last_error", "rth_engine_create", "rth_engine_destroy", "rth_engine_ready", "rth_engine_camera_count",
           "rth_engine_camera", "rth_engine_scene", "rth_engine_describe", "rth_engine_instance_object_ids",
           "rth_engine_render", "rth_engine_inspect", "rth_engine_camera_name", "rth_engine_set_instance_transforms",
           "rth_image_write", "rth_context_create", "rth_context_destroy", "rth_context_describe"]

_lib = None


def load():
    """Load librtcuda.so and declare prototypes.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C myraytracer_b200/csrc` "
                           "(or __graft_entry__.build()); there is no CPU fallback")
    _lib = declare(C.CDLL(LIB_PATH))
    return _lib


def declare(L):
    """Set the prototypes of every C-ABI entry point on a loaded library object."""
    vp = C.c_void_p
    L.rt_abi_version.restype = C.c_int32
    L.rt_device_count.restype = C.c_int32
    L.rt_last_error.restype = C.c_char_p
    L.rt_scene_create.restype = vp
    L.rt_scene_create.argtypes = [C.POINTER(rt_scene_desc), C.c_int32]
    L.rt_scene_create_multi.restype = vp
    L.rt_scene_create_multi.argtypes = [C.POINTER(rt_scene_desc), C.POINTER(C.c_int32), C.c_int32]
    L.rt_scene_device_count.argtypes = [vp]
    L.rt_host_alloc.restype = vp
    L.rt_host_alloc.argtypes = [C.c_uint64]
    L.rt_host_free.argtypes = [vp]
    L.rt_host_free.restype = None
    L.rth_engine_create_multi.restype = vp
    L.rth_engine_create_multi.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.c_int32]
    L.rt_scene_destroy.argtypes = [vp]
    L.rt_scene_destroy.restype = None
    L.rt_scene_get_info.argtypes = [vp, C.POINTER(rt_scene_info)]
    L.rt_scene_get_schedule.argtypes = [vp, C.POINTER(C.c_int32), C.c_int64]
    L.rt_scene_get_schedule.restype = C.c_int64
    L.rt_scene_read_blas_nodes.argtypes = [vp, vp, C.c_int64]
    L.rt_scene_read_blas_nodes.restype = C.c_int64
    L.rt_scene_read_wide_nodes.argtypes = [vp, vp, C.c_int64]
    L.rt_scene_read_wide_nodes.restype = C.c_int64
    L.rt_ply_ingest.argtypes = [C.POINTER(rt_ply_layout), vp, vp, C.c_int32, C.c_int32, vp, vp, C.POINTER(rt_ply_stats)]
    L.rt_ply_ingest_static.argtypes = [C.POINTER(rt_ply_layout), vp, vp, C.c_int32, C.c_int32, vp, vp, vp, C.POINTER(rt_ply_stats)]
    L.rt_render.argtypes = [vp, C.POINTER(rt_camera), C.POINTER(rt_render_params), vp, vp, vp, C.POINTER(rt_stats)]
    L.rt_render_device.argtypes = [vp, C.POINTER(rt_camera), C.POINTER(rt_render_params), vp, vp, vp, vp]
    L.rt_render_finish.argtypes = [vp, C.POINTER(rt_stats)]
    L.rt_shared_buffer_create.argtypes = [C.c_int32, C.c_uint64, C.POINTER(vp), vp]
    L.rt_shared_buffer_open.argtypes = [C.c_int32, vp, C.POINTER(vp)]
    L.rt_shared_buffer_close.argtypes = [C.c_int32, vp, C.c_int32]
    L.rth_last_error.restype = C.c_char_p
    L.rth_engine_create.restype = vp
    L.rth_engine_create.argtypes = [C.c_char_p, C.c_int32]
    L.rth_engine_destroy.argtypes = [vp]
    L.rth_engine_destroy.restype = None
    L.rth_engine_ready.argtypes = [vp]
    L.rth_engine_camera_count.argtypes = [vp]
    L.rth_engine_camera.argtypes = [vp, C.c_int32, C.POINTER(rt_camera)]
    L.rth_engine_scene.restype = vp
    L.rth_engine_scene.argtypes = [vp]
    L.rth_engine_describe.argtypes = [vp, C.POINTER(rt_scene_desc)]
    L.rth_engine_instance_object_ids.argtypes = [vp, vp, C.c_int64]
    L.rth_engine_render.argtypes = [vp, C.c_int32, C.POINTER(rt_render_params), vp, vp, vp, C.POINTER(rt_stats)]
    L.rt_scene_update_instances.argtypes = [vp, C.POINTER(rt_instance), C.c_int64, vp]
    L.rth_engine_inspect.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.rth_engine_camera_name.argtypes = [vp, C.c_int32, C.c_char_p, C.c_int32, C.POINTER(C.c_int32)]
    L.rth_engine_set_instance_transforms.argtypes = [vp, C.POINTER(C.c_double), C.c_int64, C.c_int32]
    L.rth_image_write.argtypes = [C.c_char_p, vp, C.c_int32, C.c_int32]
    L.rth_context_create.restype = vp
    L.rth_context_create.argtypes = [C.c_char_p]
    L.rth_context_destroy.argtypes = [vp]
    L.rth_context_destroy.restype = None
    L.rth_context_describe.argtypes = [vp, C.POINTER(rt_scene_desc)]
    return L
