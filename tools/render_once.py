#!/usr/bin/env python
"""One frame of a bench config through the engine (for ncu captures): python tools/render_once.py c3area [frames]"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from myraytracer_b200 import RayTracerEngine  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "c3area"
e = RayTracerEngine(bench.make_scene(cfg, 0))
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 1):
    r = e.render(0)
print(cfg, r.stats["deviceMs"], r.stats["rays"])
