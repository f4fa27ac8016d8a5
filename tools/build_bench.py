#!/usr/bin/env python
"""Mesh BVH builders side by side: host binned SAH, device radix tree (LBVH), device binned SAH.
For each: scene build time (rt_scene_info.buildMs: BVH construction + flatten, before the upload), tree size and the
frame time it gives on the bench workload.  Prints one JSON line.   python tools/build_bench.py [c4|c2] [reps]"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myraytracer_b200 import RayTracerEngine, scenegen  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c4"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    d = tempfile.mkdtemp(prefix="rtbuild_")
    path = scenegen.config_c4(d) if which == "c4" else scenegen.config_c2(d)
    RayTracerEngine(scenegen.config_c1(d, width=16, height=16)).close()      # CUDA context up before anything is timed
    out = dict(workload=which, builders={})
    base = None
    for mode in ("host", "auto", "device", "device-sah"):
        os.environ["RTCUDA_BUILD"] = mode
        build, create, upload = [], [], []
        for _ in range(reps):
            t0 = time.perf_counter()
            eng = RayTracerEngine(path)
            create.append((time.perf_counter() - t0) * 1e3)     # the whole engine: file, flatten / device ingest, BVHs, upload
            assert eng.ready, eng.load_error
            info = eng.scene_info()
            build.append(info["buildMs"])
            upload.append(info["uploadMs"])
            if _ < reps - 1:
                eng.close()
        frames = []
        for k in range(8):                                                  # the first frames train the tile schedule
            r = eng.render(0, counters=(k == 7))
            frames.append(r.stats["deviceMs"])
        rgba = r.rgba
        if base is None:
            base = rgba
        out["builders"][mode] = dict(create_ms_min=min(create), create_ms_all=create, upload_ms_min=min(upload), build_ms_min=min(build), build_ms_all=build, blasNodes=info["blasNodes"], maxDepth=info["maxDepth"],
                                     maxLeafSize=info["maxLeafSize"], deviceBuiltBlas=info["deviceBuiltBlas"],
                                     frame_ms_min=min(frames[2:7]), boxTests=r.stats["boxTests"], triTests=r.stats["triTests"],
                                     image_equals_host=bool((rgba == base).all()))
        eng.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
