#!/usr/bin/env python
"""Animation path on one B200: per frame, new instance transforms on a resident scene
(rt_scene_update_instances: instance records + device bounds + device TLAS refit) against what the reference's batch
driver does per frame — a new engine: parse, flatten, every BVH, upload (BatchRender.swift:26-40).
Prints one JSON line.   python tools/anim_bench.py [--meshes 4] [--n 708] [--frames 12]"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myraytracer_b200 import RayTracerEngine, scenegen  # noqa: E402
from myraytracer_b200.engine import FlatContext  # noqa: E402


def rot_y(deg):
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    return np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1.0]])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--meshes", type=int, default=4)
    ap.add_argument("--n", type=int, default=708)
    ap.add_argument("--frames", type=int, default=12)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--samples", type=int, default=4)
    a = ap.parse_args()
    d = tempfile.mkdtemp(prefix="rtanim_")
    path = scenegen.config_c5(d, n=a.n, width=a.width, height=a.height, samples=a.samples, meshes=a.meshes)
    t0 = time.perf_counter()
    eng = RayTracerEngine(path)
    assert eng.ready, eng.load_error
    create_s = time.perf_counter() - t0          # what the reference pays per frame (plus its own, slower, BVH build)
    info = eng.scene_info()
    base = FlatContext(path).instances()["l2w"].reshape(-1, 4, 4).transpose(0, 2, 1)      # row-major 4x4 per instance
    eng.render(0)
    res = {}
    for mode in ("device", "host"):
        upd, ren = [], []
        for f in range(a.frames):
            M = np.stack([rot_y(0.7 * (f + 1) * (1 + 0.1 * i)) @ base[i] for i in range(len(base))])
            l2w = M.transpose(0, 2, 1).reshape(-1, 16)
            t1 = time.perf_counter()
            eng.set_instance_transforms(l2w, bounds_on_device=(mode == "device"))
            t2 = time.perf_counter()
            r = eng.render(0)
            upd.append((t2 - t1) * 1e3)
            ren.append(r.stats["deviceMs"])
        res[mode] = dict(update_ms=float(np.median(upd)), render_device_ms=float(np.median(ren)), rays=int(r.stats["rays"]))
    line = dict(workload=f"C5-style: {a.meshes} x {info['triangles'] // max(1, a.meshes)}-triangle terrains + 3 spheres, "
                         f"{a.width}x{a.height}, {a.samples} spp, depth 2; every instance turns a little per frame",
                instances=int(info["instances"]), triangles=int(info["triangles"]), frames=a.frames,
                engine_create_s=create_s, build_ms=info["buildMs"], upload_ms=info["uploadMs"],
                update_bounds_on_device=res["device"], update_bounds_on_host=res["host"])
    print(json.dumps(line))


if __name__ == "__main__":
    main()
