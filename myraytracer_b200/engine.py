"""Python face of the host mirror: RayTracerEngine / RenderStats over the C ABI.

Mirrors the reference's public entry points (Sources/RayTracer/RayTracer.swift:25-205):
``RayTracerEngine(from: url)``, ``render(format:cameraIndex:progress:)``, ``inspect``.
All work happens in librtcuda.so (C++ host mirror -> C ABI -> CUDA); this module only marshals.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Callable, Optional

import numpy as np

from . import capi


class RenderError(RuntimeError):
    """NSError(domain:code:) of the reference (RayTracer.swift:121-123,142-154)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"[{code}] {message}")
        self.code = code
        self.message = message


@dataclass
class RenderResult:
    rgba: np.ndarray                  # (H, W, 4) uint8, row 0 = top
    stats: dict                       # RenderStats + ray split + deviceMs
    linear: Optional[np.ndarray] = None   # (H, W, 3) float64: the [Vec3] Renderer.render returns
    ids: Optional[np.ndarray] = None      # (H, W, n*n) HIT_ID_DTYPE


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RayTracerEngine:
    def __init__(self, url: str, device: int = 0, devices=None):
        """devices: a list of CUDA device indices — the scene is built once and replicated onto every one of them
        (rt_scene_create_multi); render() then shards each frame across them inside the one rt_render call."""
        self._lib = capi.load()
        if devices is not None and len(devices) > 1:
            arr = (C.c_int32 * len(devices))(*devices)
            self._h = self._lib.rth_engine_create_multi(url.encode(), arr, len(devices))
        else:
            self._h = self._lib.rth_engine_create(url.encode(), devices[0] if devices else device)
        if not self._h:
            raise MemoryError(self._lib.rth_last_error().decode())
        self.load_error = "" if self.ready else self._lib.rth_last_error().decode()

    @property
    def ready(self) -> bool:
        return bool(self._h) and self._lib.rth_engine_ready(self._h) == 1

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rth_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- introspection ----------------------------------------------------
    def camera_count(self) -> int:
        return self._lib.rth_engine_camera_count(self._h)

    def camera(self, index: int = 0) -> capi.rt_camera:
        cam = capi.rt_camera()
        rc = self._lib.rth_engine_camera(self._h, index, C.byref(cam))
        if rc != 0:
            raise RenderError(rc, self._lib.rth_last_error().decode())
        return cam

    def scene_handle(self):
        """rt_scene_t* of the underlying Renderer (for direct rt_render_device calls)."""
        return self._lib.rth_engine_scene(self._h)

    def scene_info(self) -> dict:
        info = capi.rt_scene_info()
        if self._lib.rt_scene_get_info(self.scene_handle(), C.byref(info)) != 0:
            raise RenderError(-21, self._lib.rt_last_error().decode())
        return info.as_dict()

    def schedule(self) -> np.ndarray:
        """rt_scene_get_schedule: tile order learnt for the next frame (empty before the first adaptive frame)."""
        n = self._lib.rt_scene_get_schedule(self.scene_handle(), None, 0)
        if n < 0:
            raise RenderError(int(n), self._lib.rt_last_error().decode())
        out = np.zeros(n, np.int32)
        if n:
            self._lib.rt_scene_get_schedule(self.scene_handle(), out.ctypes.data_as(C.POINTER(C.c_int32)), n)
        return out

    def blas_nodes(self) -> np.ndarray:
        """rt_scene_read_blas_nodes: the mesh BVH node records on the device, (n, 16) uint32 words each."""
        n = self._lib.rt_scene_read_blas_nodes(self.scene_handle(), None, 0)
        if n < 0:
            raise RenderError(int(n), self._lib.rt_last_error().decode())
        out = np.zeros((n // 64, 16), np.uint32)
        if n:
            self._lib.rt_scene_read_blas_nodes(self.scene_handle(), _ptr(out), n)
        return out

    def wide_nodes(self) -> np.ndarray:
        """rt_scene_read_wide_nodes: the wide BVH nodes of the beam walk, (n, 16 children, 8) uint32 words:
        lo.xyz, hi.xyz (float bits), child reference, aux (child 0: sort axis)."""
        n = self._lib.rt_scene_read_wide_nodes(self.scene_handle(), None, 0)
        if n < 0:
            raise RenderError(int(n), self._lib.rt_last_error().decode())
        out = np.zeros((n // 512, 16, 8), np.uint32)
        if n:
            self._lib.rt_scene_read_wide_nodes(self.scene_handle(), _ptr(out), n)
        return out

    def describe(self) -> capi.rt_scene_desc:
        d = capi.rt_scene_desc()
        rc = self._lib.rth_engine_describe(self._h, C.byref(d))
        if rc != 0:
            raise RenderError(rc, self._lib.rth_last_error().decode())
        return d

    def instance_object_ids(self) -> np.ndarray:
        n = self._lib.rth_engine_instance_object_ids(self._h, None, 0)
        out = np.zeros(max(n, 0), np.int32)
        self._lib.rth_engine_instance_object_ids(self._h, _ptr(out), n)
        return out

    def inspect(self) -> dict:
        """RayTracerEngine.inspect(scene:format:) (RayTracer.swift:52-67): SceneInfo as a dict."""
        counts = (C.c_int64 * 4)()
        ncam = C.c_int32()
        rc = self._lib.rth_engine_inspect(self._h, counts, C.byref(ncam))
        if rc != 0:
            raise RenderError(rc, self._lib.rth_last_error().decode())
        cams = []
        for i in range(ncam.value):
            name = C.create_string_buffer(256)
            cid = C.c_int32()
            self._lib.rth_engine_camera_name(self._h, i, name, 256, C.byref(cid))
            cam = self.camera(i)
            cams.append(dict(index=i, id=cid.value, imageName=name.value.decode(), width=cam.width, height=cam.height))
        return dict(cameras=cams, meshes=counts[0], triangles=counts[1], spheres=counts[2], planes=counts[3])

    def render_all(self, progress: Optional[Callable[[float, str], bool]] = None, **kw):
        """RayTracerEngine.renderAll(progress:) (RayTracer.swift:70-103): every camera in order; the image name is
        the camera's imageName or "Camera-<index>".  progress(fraction, message) -> keep going?"""
        info = self.inspect()
        total = max(1, sum(max(1, c["height"]) for c in info["cameras"]))
        done_rows = 0
        out = []
        for c in info["cameras"]:
            h = max(1, c["height"])
            base = done_rows
            cb = None
            if progress is not None:
                def cb(done, rows, base=base, c=c, h=h):
                    y = min(h - 1, max(0, done - 1))
                    return progress((base + y + 1) / total, f"Camera {c['index'] + 1}/{len(info['cameras'])} — Row {y + 1}/{h}")
            res = self.render(c["index"], progress=cb, **kw)
            out.append((c["imageName"] or f"Camera-{c['index']}", res))
            done_rows += h
        if progress is not None:
            progress(1.0, "Done")
        return out

    # -- animation --------------------------------------------------------
    def set_instance_transforms(self, local_to_world, bounds_on_device: bool = True) -> None:
        """A new frame of an animation: one column-major 4x4 localToWorld per instance (RTContext order).  Replaces
        the per-frame engine rebuild of BatchRender.swift:26-40 by RTContext.refitTLAS semantics
        (RTContext.swift:883-928): instance records re-derived, TLAS boxes refitted on the device."""
        m = np.ascontiguousarray(np.asarray(local_to_world, np.float64).reshape(-1, 16))
        rc = self._lib.rth_engine_set_instance_transforms(self._h, m.ctypes.data_as(C.POINTER(C.c_double)), m.shape[0],
                                                          1 if bounds_on_device else 0)
        if rc != 0:
            raise RenderError(rc, self._lib.rth_last_error().decode())

    # -- render -----------------------------------------------------------
    def render(self, camera_index: int = 0, want_linear: bool = False, want_ids: bool = False,
               shard_index: int = 0, shard_count: int = 1, counters: bool = False,
               progress: Optional[Callable[[int, int], bool]] = None, variant: int = 0, schedule: int = 0,
               shard_mode: int = 0, out_rgba: Optional[np.ndarray] = None) -> RenderResult:
        """out_rgba: caller-owned (H, W, 4) uint8 buffer to render into, e.g. page-locked memory of host_buffer()
        (what a Swift caller passes to rt_render); default: a fresh pageable numpy array."""
        if not self._h or self._lib.rth_engine_camera_count(self._h) == 0 and not self.ready:
            pass  # let the C side produce the reference's error codes
        cam = None
        try:
            cam = self.camera(camera_index)
        except RenderError:
            cam = None
        W = max(1, cam.width) if cam else 1
        H = max(1, cam.height) if cam else 1
        n = int(np.floor(np.sqrt(max(1, cam.numSamples)))) if cam else 1
        rgba = np.zeros((H, W, 4), np.uint8) if out_rgba is None else out_rgba
        if rgba.shape != (H, W, 4) or rgba.dtype != np.uint8 or not rgba.flags["C_CONTIGUOUS"]:
            raise ValueError("out_rgba must be a C-contiguous (H, W, 4) uint8 array")
        linear = np.zeros((H, W, 3), np.float64) if want_linear else None
        ids = np.zeros((H, W, n * n), capi.HIT_ID_DTYPE) if want_ids else None
        prm = capi.rt_render_params()
        prm.shardIndex, prm.shardCount, prm.collectCounters = shard_index, shard_count, 1 if counters else 0
        prm.kernelVariant = variant
        prm.scheduleMode = schedule     # 0: longest tiles of the previous frame first, 1: spatial order
        prm.shardMode = shard_mode      # 0: 8-row chunks, 1: 8x4 tiles along diagonals
        cb = None
        if progress is not None:
            cb = capi.rt_progress_fn(lambda done, total, _u: 0 if progress(done, total) else 1)
            prm.progress = cb
        st = capi.rt_stats()
        rc = self._lib.rth_engine_render(self._h, camera_index, C.byref(prm), _ptr(rgba), _ptr(linear), _ptr(ids), C.byref(st))
        if rc != 0:
            raise RenderError(rc, self._lib.rth_last_error().decode())
        return RenderResult(rgba=rgba, stats=st.as_dict(), linear=linear, ids=ids)


    def render_device(self, camera_index: int = 0, d_rgba: int = 0, d_linear: int = 0, d_ids: int = 0, stream: int = 0,
                      shard_index: int = 0, shard_count: int = 1, counters: bool = False, variant: int = 0,
                      schedule: int = 0, shard_mode: int = 0) -> None:
        """rt_render_device: enqueue one frame on `stream` with caller-owned DEVICE buffers (raw pointers,
        e.g. torch.Tensor.data_ptr()); does not synchronise.  Pair with finish()."""
        cam = self.camera(camera_index)
        prm = capi.rt_render_params()
        prm.shardIndex, prm.shardCount, prm.collectCounters = shard_index, shard_count, 1 if counters else 0
        prm.kernelVariant = variant
        prm.scheduleMode = schedule
        prm.shardMode = shard_mode
        rc = self._lib.rt_render_device(self.scene_handle(), C.byref(cam), C.byref(prm), d_rgba or None, d_linear or None,
                                        d_ids or None, stream or None)
        if rc != 0:
            raise RenderError(rc, self._lib.rt_last_error().decode())

    def finish(self) -> dict:
        """rt_render_finish: wait for the frame enqueued by render_device; returns the stats dict."""
        st = capi.rt_stats()
        rc = self._lib.rt_render_finish(self.scene_handle(), C.byref(st))
        if rc != 0:
            raise RenderError(rc, self._lib.rt_last_error().decode())
        return st.as_dict()


class host_buffer:
    """Page-locked host memory from rt_host_alloc as a numpy array (`.array`); freed by close() / the context manager."""

    def __init__(self, shape, dtype=np.uint8):
        self._lib = capi.load()
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._p = self._lib.rt_host_alloc(n)
        if not self._p:
            raise MemoryError(self._lib.rt_last_error().decode())
        self.array = np.ctypeslib.as_array((C.c_uint8 * n).from_address(self._p)).view(dtype).reshape(shape)

    def close(self):
        if getattr(self, "_p", None):
            self.array = None
            self._lib.rt_host_free(self._p)
            self._p = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def write_image(path: str, rgba: np.ndarray) -> None:
    """Headless stand-in for makeCGImage + PNG output (Helpers/Image.swift): .png or .ppm by extension."""
    a = np.ascontiguousarray(rgba, np.uint8)
    if a.ndim != 3 or a.shape[2] != 4:
        raise ValueError("rgba must be (H, W, 4) uint8")
    rc = capi.load().rth_image_write(path.encode(), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0])
    if rc != 0:
        raise RenderError(rc, f"cannot write {path}")


class FlatContext:
    """Scene + RTContext(scene:) of the host mirror without a renderer (runs without a GPU)."""

    def __init__(self, url: str):
        self._lib = capi.load()
        self._h = self._lib.rth_context_create(url.encode())
        if not self._h:
            raise RuntimeError(self._lib.rth_last_error().decode())
        self.desc = capi.rt_scene_desc()
        self._lib.rth_context_describe(self._h, C.byref(self.desc))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rth_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def triangles(self):
        n = self.desc.nTriangles
        if n == 0:
            return np.zeros((0, 3, 3)), np.zeros((0, 3, 3))
        v = np.ctypeslib.as_array(self.desc.triVertices, shape=(n, 3, 3)).copy()
        nr = np.ctypeslib.as_array(self.desc.triNormals, shape=(n, 3, 3)).copy()
        return v, nr

    def instances(self):
        n = self.desc.nInstances
        out = dict(l2w=np.zeros((n, 16)), w2l=np.zeros((n, 16)), nmat=np.zeros((n, 9)), bounds=np.zeros((n, 6)),
                   blas=np.zeros(n, np.int64), mat=np.zeros(n, np.int64))
        for i in range(n):
            I = self.desc.instances[i]
            out["l2w"][i] = list(I.localToWorld)
            out["w2l"][i] = list(I.worldToLocal)
            out["nmat"][i] = list(I.normalMatrix)
            out["bounds"][i] = list(I.worldBounds)
            out["blas"][i] = I.blas
            out["mat"][i] = I.materialOverride
        return out
