#!/usr/bin/env python
"""One pass over the kernels beyond the headline render kernel, for an ncu launch list (profiles/):
PLY ingest of the C4 mesh, an area-light frame (counting pass, prefix, shading pass), a motion-blur frame, an animation
update (instance bounds + TLAS refit) and its frame.   ncu --metrics gpu__time_duration.sum ... python tools/profile_ext.py"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myraytracer_b200 import RayTracerEngine, scenegen  # noqa: E402
from myraytracer_b200.engine import FlatContext  # noqa: E402

d = tempfile.mkdtemp(prefix="rtprof_")
os.environ["RTCUDA_INGEST"] = "device"
e = RayTracerEngine(scenegen.config_c4(d, width=960, height=540, samples=4))       # ingest kernels (999,698 triangles)
e.render(0)
l2w = FlatContext(scenegen.config_c4(d, width=960, height=540, samples=4)).instances()["l2w"]
e.set_instance_transforms(l2w)                                                       # instanceBounds + tlasRefit (1 instance: no TLAS nodes)
e.close()
a = RayTracerEngine(scenegen.config_c3(d, area_lights=2))                            # C3 + area lights, 1024x1024x4
a.render(0); a.render(0)
a.close()
m = RayTracerEngine(scenegen.config_instances(d, width=1024, height=768, samples=4, mesh_motion=(0.4, -0.3, 0.2), instance_motion=1.5))
m.render(0); m.render(0)
m.close()
k = RayTracerEngine(scenegen.config_c5(d, n=256, width=1280, height=720, samples=4, meshes=6))
base = FlatContext(scenegen.config_c5(d, n=256, width=1280, height=720, samples=4, meshes=6)).instances()["l2w"]
for f in range(2):
    t = base.copy()
    t[:, 12] += 0.3 * (f + 1)
    k.set_instance_transforms(t)
    k.render(0)
k.close()
print("done")
