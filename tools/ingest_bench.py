#!/usr/bin/env python
"""PLY -> flattened triangle arrays: host loader + flatten loops against the device ingest
(rt_ply_ingest), on the C4 mesh (999,698 triangles) and through whole-engine creation.  Prints one JSON line."""
import ctypes as C
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myraytracer_b200 import RayTracerEngine, capi, scenegen  # noqa: E402


def context_ms(L, path, mode, reps=3):
    os.environ["RTCUDA_INGEST"] = mode
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        h = L.rth_context_create(path.encode())
        best = min(best, (time.perf_counter() - t0) * 1e3)
        d = capi.rt_scene_desc()
        L.rth_context_describe(h, C.byref(d))
        n = d.nTriangles
        L.rth_context_destroy(h)
    return best, n


def main():
    L = capi.load()
    d = tempfile.mkdtemp(prefix="rtingest_")
    path = scenegen.config_c4(d, width=64, height=64, samples=1)
    L.rt_device_count()
    RayTracerEngine(scenegen.config_c1(d, width=16, height=16)).close()      # CUDA context up before anything is timed
    host_ms, n = context_ms(L, path, "host")
    dev_ms, _ = context_ms(L, path, "device")
    out = dict(workload="C4 mesh: 999,698-triangle binary PLY, smooth normals", triangles=int(n),
               flatten_host_ms=host_ms, flatten_device_ingest_ms=dev_ms)
    for mode in ("host", "device"):
        os.environ["RTCUDA_INGEST"] = mode
        t0 = time.perf_counter()
        e = RayTracerEngine(path)
        out[f"engine_create_{mode}_ingest_ms"] = (time.perf_counter() - t0) * 1e3
        info = e.scene_info()
        out["bvh_build_ms"] = info["buildMs"]
        e.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
