#!/usr/bin/env python
"""bench.py — Mrays/s (primary + shadow + secondary, device-timed) of the render hot path.

Workload (BASELINE.json configs[3], "C4"): synthetic 999,698-triangle terrain PLY, 1920x1080, 16 spp,
2 point lights with shadow rays; one "step" = one full frame through Renderer.render's replacement.
N GPUs: the frame's 8-row chunks are dealt round-robin to the ranks (strong scaling of ONE frame) and
the RGBA8 framebuffer is gathered to rank 0 over NVLink inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c4]

One JSON line on stdout (rank 0).  `value` = device-timed (CUDA events on the launch stream, inputs
resident in HBM); `e2e` = the same frame through the public call with the framebuffer copied back to
pinned host memory every step; `roofline` = algorithmic bytes of the render kernel / its duration
against the measured HBM peak; `cpu_baseline` = the CPU oracle (C++ FP64 restatement of the reference's
Swift loop, the reference itself cannot be built off macOS — DESIGN.md) on a bounded sample of the frame.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/sec (primary+secondary, device-timed)"
UNIT = "Mrays/s"
# algorithmic bytes per counted event
BYTES = dict(boxTests=32, triTests=36, sphereTests=16, planeTests=24, instEntries=48, smoothHits=36)

BENCH_CONFIGS = {
    # name: (scenegen function name, kwargs, description)
    "c4": ("config_c4", {}, "C4: 999,698-tri terrain PLY, 1920x1080, 16 spp, 2 point lights + shadow rays"),
    "c2": ("config_c2", {}, "C2: 69,938-tri terrain PLY, 1024x1024, 1 spp, 1 point light + shadow rays"),
    "c3": ("config_c3", {}, "C3: mirror/dielectric/conductor recursion depth 6, 1024x1024, 4 spp"),
    "c1": ("config_c1", {}, "C1: spheres + triangles, 2 point lights, 800x800, 1 spp"),
    "c5": ("config_c5", {}, "C5: 10 x 999,698-tri terrains + 3 reflective spheres, 3840x2160, 64 spp, depth 2"),
    # SURVEY.md 8f row 3 (beyond BASELINE.json's configs; general kernel): glossy recursion, area lights
    "c3glossy": ("config_c3", {"roughness": 0.12}, "C3 with glossy mirror/dielectric/conductor (roughness 0.12): RNG draws inside trace"),
    "c3area": ("config_c3", {"area_lights": 2}, "C3 with 1 point light + 2 area lights: counting pass + per-chunk prefix + shading pass"),
}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


_REAL_STDOUT = None


def capture_stdout():
    """Everything libraries print to fd 1 (NCCL's version banner, ...) goes to stderr; the contract's ONE JSON line is
    written to the real stdout by emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def make_scene(config: str, rank: int) -> str:
    """Write the synthetic scene for `config` (seeded, deterministic) into a per-rank scratch dir."""
    from myraytracer_b200 import scenegen
    fn, kw, _ = BENCH_CONFIGS[config]
    d = os.path.join(tempfile.gettempdir(), f"rtbench_{config}_r{rank}_{os.getpid()}")
    os.makedirs(d, exist_ok=True)
    return getattr(scenegen, fn)(d, **kw)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except Exception as e:  # nvidia-smi missing: report that, do not fail the bench
            log("clock sampler unavailable:", e)
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle (oracle/liboracle.so) timed on the host cores.  The only places bench.py touches
# oracle/: the `cpu_baseline` leg and `--impl reference`.  Never on the product path.
# --------------------------------------------------------------------------------------------------
def cpu_sample(path: str, target_s: float = 15.0, stride: int = 0, scene=None):
    """Render a bounded sample of the frame (every `stride`-th 8-row chunk, the reference's own unit of
    parallel work, Object+Extension.swift:75-82) with all host threads.  Returns (Mrays/s, info)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    o = scene if scene is not None else oracle_lib.OracleScene(path)
    cam = o.camera(0)
    nchunks = (cam.height + 7) // 8
    cores = os.cpu_count() or 1
    if stride <= 0:
        # size the sample: one chunk from the middle of the frame on one thread -> estimate of the full frame
        p = o.render(threads=1, chunk_stride=nchunks, chunk_offset=nchunks // 2)
        est_full_s = p["ms"] * 1e-3 * nchunks / min(cores, nchunks)
        stride = max(1, min(nchunks, int(round(est_full_s / target_s))))
    r = o.render(threads=0, chunk_stride=stride, chunk_offset=0)
    c = r["counters"]
    rays = c["raysPrimary"] + c["raysShadow"] + c["raysSecondary"]
    mrays = rays / (r["ms"] * 1e-3) / 1e6
    nsel = len(range(0, nchunks, stride))
    info = {"rays": rays, "ms": r["ms"], "cores": cores, "stride": stride, "chunks": nsel, "nchunks": nchunks,
            "sample": f"every {stride}-th 8-row chunk ({nsel} of {nchunks} chunks, {c['raysPrimary']} primary rays) of the "
                      f"{cam.width}x{cam.height}x{cam.numSamples}spp frame, {cores} threads, FP64, reference BVH (SAH 12 bins, maxLeaf 2)",
            "counters": c, "build_ms": o.info().buildMs}
    info["scene"] = o
    return mrays, info


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  The Swift
    sources cannot be built here or on the GPU box (no Swift toolchain; Apple-only simd/CoreGraphics;
    un-vendored ParsingKit — DESIGN.md), so this is the oracle port (`kind: port`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    path = make_scene(args.config, 0)
    # steps are bounded samples: size one step to ~6 s so warmup + steps end within a few minutes
    mr, info = cpu_sample(path, target_s=6.0)
    stride, scene = info["stride"], info["scene"]
    for _ in range(max(0, args.warmup - 1)):
        cpu_sample(path, stride=stride, scene=scene)
    tot_rays, tot_ms = 0, 0.0
    for _ in range(args.steps):
        _, i2 = cpu_sample(path, stride=stride, scene=scene)
        tot_rays += i2["rays"]; tot_ms += i2["ms"]
    value = tot_rays / (tot_ms * 1e-3) / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": BENCH_CONFIGS[args.config][2], "sample": info["sample"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["cores"], "kind": "port", "sample": info["sample"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def algorithmic_bytes(st: dict, pixels: int) -> int:
    """32 B per box test, 36 B per triangle test, 16/24 B per sphere/plane test, 48 B per
    instance entry, 36 B per smooth-shaded closest hit, 4 B per pixel written."""
    return sum(BYTES[k] * int(st[k]) for k in BYTES) + 4 * pixels


def run_ours(args):
    import torch
    import torch.distributed as dist
    from myraytracer_b200 import capi
    from myraytracer_b200.distributed import ShardedRenderer

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        log(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
    L = capi.load()  # raises if librtcuda.so is missing: no fallback
    if not torch.cuda.is_available() or L.rt_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the render path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def allreduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        return t.item()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    if args.build:
        os.environ["RTCUDA_BUILD"] = args.build        # rt_scene_desc.buildMode override (include/rtcuda.h)
    t0 = time.time()
    path = make_scene(args.config, rank)
    R = ShardedRenderer(path, device=local, rank=rank, world=world, variant=args.variant, gather=args.gather, shard=args.shard,
                        schedule=1 if args.schedule == "spatial" else 0)
    info = R.engine.scene_info()
    W, H = R.width, R.height
    log(f"[rank {rank}] scene ready in {time.time() - t0:.1f}s: {info}")

    # ---- counters pass (untimed, slower kernel variant): algorithmic bytes of this rank's launch
    R.enqueue(counters=True)
    cst = R.finish()
    n_side = max(1, int(max(1, int(R.engine.camera(0).numSamples)) ** 0.5))
    my_pixels = int(cst["raysPrimary"]) // (n_side * n_side)                 # pixels of this rank's shard
    my_bytes = algorithmic_bytes(cst, my_pixels)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        if not args.no_flush:
            flush.zero_()                 # L2 flush between timed iterations (outside the event pair)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        R.enqueue()
        e1.record()
        st = R.finish()
        e1.synchronize()
        return e0.elapsed_time(e1), st

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()          # every rank samples its own GPU (the ranks' kernel times differ by GPU); rank 0's is reported
    wall0 = time.perf_counter()
    ms_sum, kern_ms_sum, st = 0.0, 0.0, None
    for _ in range(args.steps):
        ms, st = step()
        ms_sum += ms; kern_ms_sum += st["deviceMs"]
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    log(f"[rank {rank}] step {ms_sum / args.steps:.3f} ms, kernel {kern_ms_sum / args.steps:.3f} ms, rays {st['rays']}, "
        f"clocks {clocks}")
    ms_step = allreduce(ms_sum / args.steps, dist.ReduceOp.MAX if world > 1 else None)
    rays = allreduce(float(st["rays"]), dist.ReduceOp.SUM if world > 1 else None)
    rp = allreduce(float(st["raysPrimary"]), dist.ReduceOp.SUM if world > 1 else None)
    rs = allreduce(float(st["raysShadow"]), dist.ReduceOp.SUM if world > 1 else None)
    value = rays / (ms_step * 1e-3) / 1e6
    kern_ms = kern_ms_sum / args.steps
    launches = int(allreduce(float(st["kernelLaunches"] * args.steps), dist.ReduceOp.SUM if world > 1 else None))
    scheduled = int(st["tilesScheduled"])

    # ---- the same frame with the tiles drawn top to bottom (what the first frame of a camera gets), for comparison
    spatial_ms = None
    if R.schedule == 0:
        R.schedule = 1
        step()
        barrier()
        sp_sum = sum(step()[0] for _ in range(min(5, args.steps)))
        spatial_ms = allreduce(sp_sum / min(5, args.steps), dist.ReduceOp.MAX if world > 1 else None)
        R.schedule = 0
        barrier()

    # ---- e2e: public call, framebuffer back in pinned host memory every step
    for _ in range(2):
        R.render()
    barrier()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()
        R.render()
    barrier()
    e2e_ms = allreduce((time.perf_counter() - t1) * 1e3 / args.steps, dist.ReduceOp.MAX if world > 1 else None)
    cam_bytes = ctypes_sizeof_camera()
    e2e = {"value": rays / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": cam_bytes * world, "d2h_bytes_per_step": W * H * 4 + 128 * world,
           "note": "scene/BVH resident (uploaded once at engine creation, as the reference builds RTContext before its "
                   "timer, RayTracer.swift:44,166); per step the camera/launch parameters go down and the RGBA8 frame "
                   "+ ray counters come back to pinned host memory"}

    R.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = my_bytes / (kern_ms * 1e-3) / 1e9
    general = args.variant == 2 or args.config in ("c3glossy", "c3area")
    kernel_name = ("renderKernelGeneral<false, true> (x2: counting + shading pass)" if args.config == "c3area" else
                   "renderKernelGeneral<false, false>") if general else "renderKernelFast<false>"
    roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": "measured" if peaks else "fallback",
                "algorithmic_bytes_per_launch": my_bytes, "kernel_ms": kern_ms,
                "counts": {k: int(cst[k]) for k in list(BYTES) + ["maxStack"]},
                "note": "algorithmic bytes are served by L1/L2 (traffic = ncu DRAM bytes of one launch), so a fraction above 1 "
                        "only says the working set is cache resident; the path is issue/latency bound — profiles/, "
                        "DESIGN.md section 5"}
    try:   # DRAM traffic of one launch from the committed `ncu --set full` capture of this workload (profiles/)
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = f"{args.config}_n{world}_v{args.variant}"
        if key in tr:
            roofline["traffic"] = tr[key]["dram_bytes_per_launch"]
            roofline["traffic_source"] = tr[key]["source"]
    except Exception:
        pass
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": BENCH_CONFIGS[args.config][2], "width": W, "height": H, "triangles": info["triangles"],
                       "rays_per_frame": int(rays), "rays_primary": int(rp), "rays_shadow": int(rs),
                       "rays_secondary": int(rays - rp - rs), "parallelism": f"tiles{world}", "gather": R.gather, "shard": R.shard, "kernel_variant": args.variant,
                       "schedule": ("tiles drawn longest-first from the per-tile cost recorded on the previous frame "
                                    f"({scheduled} tiles on rank 0); warm-up frames train it" if scheduled else "spatial order"),
                       "ms_per_step_spatial_order": spatial_ms,
                       "l2": ("NOT FLUSHED (diagnostic run); " if args.no_flush else "") + "256 MB buffer written between timed iterations (L2 flush); scene working set "
                             f"{info['deviceBytes'] / 1e6:.0f} MB"},
            "wall_ms_per_step": wall * 1e3 / args.steps, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "roofline": roofline, "build_ms": info["buildMs"], "upload_ms": info["uploadMs"],
            "bvh_build": ("host binned SAH" if not info["deviceBuiltBlas"] else
                          "device binned SAH" if os.environ.get("RTCUDA_BUILD") == "device-sah" else "device LBVH")}
    if world == 1 and not args.no_cpu:
        mr, ci = cpu_sample(path, target_s=args.cpu_seconds)
        line["cpu_baseline"] = {"value": mr, "unit": UNIT, "cores": ci["cores"], "kind": "port", "sample": ci["sample"],
                                "ms": ci["ms"], "per_core": mr / ci["cores"]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def ctypes_sizeof_camera() -> int:
    import ctypes
    from myraytracer_b200 import capi
    return ctypes.sizeof(capi.rt_camera) + ctypes.sizeof(capi.rt_render_params)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", choices=sorted(BENCH_CONFIGS), default="c4")
    ap.add_argument("--shard", choices=["tiles", "rows"], default=None,
                    help="N>1: 8-row chunks round-robin (default) or 8x4 tiles dealt along diagonals (p2p only)")
    ap.add_argument("--build", choices=["host", "device", "device-sah"], default=None,
                    help="who builds the mesh BVHs: host = multi-threaded binned SAH (default), device = CUDA LBVH, "
                         "device-sah = the host builder's tree made by CUDA kernels")
    ap.add_argument("--schedule", choices=["adaptive", "spatial"], default="adaptive",
                    help="rt_render_params.scheduleMode: tile order learnt from the previous frame, or top to bottom")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-flush", action="store_true", help="diagnostic only: skip the L2 flush between frames")
    ap.add_argument("--gather", choices=["p2p", "nccl"], default="p2p",
                    help="N>1 framebuffer exchange: kernel stores into rank 0's frame over NVLink (p2p) or NCCL gather")
    ap.add_argument("--variant", type=int, default=0, help="rt_render_params.kernelVariant (0 = automatic; A/B runs)")
    args = ap.parse_args()
    capture_stdout()
    if args.warmup < 3 and args.impl == "ours":
        log("note: timing rules ask for >= 3 warm-up steps")
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
