"""myraytracer_b200 — B200-native render loop for MyRayTracer (one hot path, C ABI + CUDA sm_100a).

Layout:
  csrc/     CUDA kernels, GPU BVH builder, C-ABI implementation  -> librtcuda.so
  host/     C++ mirror of the reference's Swift interface (RTContext, Renderer, RayTracerEngine)
  capi.py   ctypes declarations of include/rtcuda.h + include/rtcuda_host.h
  engine.py Python face of RayTracerEngine
  scenegen.py seeded synthetic scenes (the five BASELINE.json configs)
"""
from . import capi  # noqa: F401
from .engine import FlatContext, RayTracerEngine, RenderError, RenderResult, write_image  # noqa: F401
