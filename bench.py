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
# oracle/: the `cpu_baseline` leg (+ the parity gate, which checks the GPU frame against it) and
# `--impl reference`.  Never on the product path.
# --------------------------------------------------------------------------------------------------
IDS_LIMIT_BYTES = 2 << 30      # hit ids (24 B per primary ray) are compared only when they fit comfortably in host memory


def cpu_sample(path: str, target_s: float = 15.0, stride: int = 0, scene=None, threads: int = 0, want_ids: bool = False):
    """Render a bounded sample of the frame (every `stride`-th 8-row chunk, the reference's own unit of
    parallel work, Object+Extension.swift:75-82) with `threads` host threads (0 = all).  Returns (Mrays/s, info);
    info["render"] keeps the oracle's frame (rgba, ids, counters) for the parity gate."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    o = scene if scene is not None else oracle_lib.OracleScene(path)
    cam = o.camera(0)
    nchunks = (cam.height + 7) // 8
    cores = os.cpu_count() or 1
    used = cores if threads <= 0 else threads
    if stride <= 0:
        # size the sample: one chunk from the middle of the frame on one thread -> estimate of the full frame
        p = o.render(threads=1, chunk_stride=nchunks, chunk_offset=nchunks // 2)
        est_full_s = p["ms"] * 1e-3 * nchunks / min(used, nchunks)
        stride = max(1, min(nchunks, int(round(est_full_s / target_s))))
    n2 = cam.sqrtSamples * cam.sqrtSamples
    ids_ok = want_ids and cam.width * cam.height * n2 * 24 <= IDS_LIMIT_BYTES
    r = o.render(threads=threads, chunk_stride=stride, chunk_offset=0, want_ids=ids_ok)
    c = r["counters"]
    rays = c["raysPrimary"] + c["raysShadow"] + c["raysSecondary"]
    mrays = rays / (r["ms"] * 1e-3) / 1e6
    nsel = len(range(0, nchunks, stride))
    info = {"rays": rays, "ms": r["ms"], "cores": used, "stride": stride, "chunks": nsel, "nchunks": nchunks,
            "sample": f"every {stride}-th 8-row chunk ({nsel} of {nchunks} chunks, {c['raysPrimary']} primary rays) of the "
                      f"{cam.width}x{cam.height}x{cam.numSamples}spp frame, {used} thread{'s' if used > 1 else ''}, FP64, "
                      "reference BVH (SAH 12 bins, maxLeaf 2)",
            "counters": c, "build_ms": o.info().buildMs, "render": r, "scene": o}
    return mrays, info


def workload_config(config: str, W: int, H: int, spp: int, triangles: int) -> dict:
    """`config` of the JSON line: the same keys from both arms."""
    return {"workload": BENCH_CONFIGS[config][2], "width": W, "height": H, "spp": spp, "triangles": triangles}


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
    cam, oi = scene.camera(0), scene.info()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.config, cam.width, cam.height, cam.numSamples, int(oi.triangles)),
            "sample": info["sample"],
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


def frame_hash(a) -> str:
    """64-bit BLAKE2b of the RGBA8 frame: one hash per frame, the same for every N (SURVEY.md §8(d) gate 4)."""
    import hashlib
    return hashlib.blake2b(a.tobytes(), digest_size=8).hexdigest()


def parity_gate(gpu, cpu_info) -> dict:
    """BASELINE.json's bars between the GPU frame of THIS run and the oracle frame of this run's cpu_baseline leg
    (tests/parity.py: compare), on the chunks the CPU sample rendered."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import parity
    ref, stride, nchunks = cpu_info["render"], cpu_info["stride"], cpu_info["nchunks"]
    H = ref["rgba"].shape[0]
    rows = np.zeros(H, bool)
    for c in range(0, nchunks, stride):
        rows[c * 8:(c + 1) * 8] = True
    g = {"rgba": gpu.rgba[rows], "ids": None if gpu.ids is None or ref["ids"] is None else gpu.ids[rows], "stats": None}
    r = {"rgba": ref["rgba"][rows], "ids": None if ref["ids"] is None else ref["ids"][rows]}
    rep = parity.compare(g, r)
    out = {k: rep[k] for k in ("px_within_1", "px_exact", "px_max_diff", "id_agree", "id_disagree", "id_bad", "t_bit_exact") if k in rep}
    c = ref["counters"]
    out["rows_compared"] = int(rows.sum())
    out["rays_ref"] = [int(c["raysPrimary"]), int(c["raysShadow"]), int(c["raysSecondary"])]
    if stride == 1:
        out["rays_gpu"] = [int(gpu.stats["raysPrimary"]), int(gpu.stats["raysShadow"]), int(gpu.stats["raysSecondary"])]
    ok = out["px_within_1"] >= 0.999 and (out.get("id_agree", 1.0) >= 0.9999) and out.get("id_bad", 0) == 0
    if stride == 1:
        ok = ok and out["rays_gpu"][0] == out["rays_ref"][0] and \
            all(abs(a - b) <= max(4, 1e-4 * b) for a, b in zip(out["rays_gpu"][1:], out["rays_ref"][1:]))
    out["pass"] = bool(ok)
    out["bars"] = "ids >= 99.99 % with disagreements only at t ties (1e-5 rel); 8-bit pixels within 1 LSB on >= 99.9 %"
    return out


def measured_roofs() -> dict:
    """Roofs measured on this pool's B200 by tools/microbench.cu (committed under profiles/) and by the driver."""
    out = {}
    try:
        out.update(json.load(open(os.path.join(ROOT, "profiles", "r2a_microbench.json"))))
    except Exception:
        pass
    try:
        out["driver"] = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    return out


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from myraytracer_b200 import RayTracerEngine, capi
    from myraytracer_b200.distributed import ShardedRenderer
    from myraytracer_b200.engine import host_buffer

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
    spp = int(R.engine.camera(0).numSamples)
    log(f"[rank {rank}] scene ready in {time.time() - t0:.1f}s: {info}")

    # ---- counters pass (untimed; the same traversal instantiated with counters): algorithmic bytes of this rank's launch
    R.enqueue(counters=True)
    cst = R.finish()
    n_side = max(1, int(max(1, spp) ** 0.5))
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

    # ---- the first frame of a camera (what one RayTracerEngine.render call gets: no learnt tile order yet)
    first_ms, _ = step()
    first_ms = allreduce(first_ms, dist.ReduceOp.MAX if world > 1 else None)
    for _ in range(max(0, args.warmup - 1)):
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

    # ---- the same frame with the tiles drawn top to bottom, for comparison
    spatial_ms = None
    if R.schedule == 0:
        R.schedule = 1
        step()
        barrier()
        sp_sum = sum(step()[0] for _ in range(min(5, args.steps)))
        spatial_ms = allreduce(sp_sum / min(5, args.steps), dist.ReduceOp.MAX if world > 1 else None)
        R.schedule = 0
        barrier()

    # ---- the multi-process public call: every rank renders its chunks into rank 0's frame over NVLink, rank 0 copies
    #      the frame to pinned host memory
    for _ in range(2):
        R.render()
    barrier()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()
        frame_ranks, _ = R.render()
    barrier()
    ranks_ms = allreduce((time.perf_counter() - t1) * 1e3 / args.steps, dist.ReduceOp.MAX if world > 1 else None)
    frame_hash_ranks = frame_hash(frame_ranks) if rank == 0 else None
    R.close()

    # ---- e2e: the call the Swift host makes — rt_render(scene, camera, params, hostRgba, ...) through the C ABI, from ONE
    #      process, host buffers, every device->host copy inside the timed region.  N > 1: a multi-device scene
    #      (rt_scene_create_multi) on all N GPUs driven by rank 0 alone; the other ranks wait on the store (no GPU work).
    e2e, parity, cpu_line, cpu1_line, frame_hash_abi = None, None, None, None, None
    store = dist.distributed_c10d._get_default_store() if world > 1 else None
    if rank == 0:
        eng = RayTracerEngine(path, devices=list(range(world)))
        assert eng.ready, eng.load_error
        flushes = [flush] + [torch.empty(256 << 20, dtype=torch.uint8, device=torch.device("cuda", k)) for k in range(1, world)]
        with host_buffer((H, W, 4)) as hb:
            for _ in range(max(2, args.warmup)):
                eng.render(0, out_rgba=hb.array)
            dts = []
            for _ in range(args.steps):
                for k, fb in enumerate(flushes):          # L2 flush on every GPU of the scene
                    fb.zero_()
                    torch.cuda.synchronize(torch.device("cuda", k))
                t1 = time.perf_counter()
                res = eng.render(0, out_rgba=hb.array)
                dts.append((time.perf_counter() - t1) * 1e3)
            e2e_ms = sum(dts) / len(dts)
            frame_hash_abi = frame_hash(hb.array)
            e2e_rays = int(res.stats["rays"])
            cam_bytes = ctypes_sizeof_camera()
            e2e = {"value": e2e_rays / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms, "ms_min": min(dts),
                   "h2d_bytes_per_step": cam_bytes * world, "d2h_bytes_per_step": W * H * 4 + 128 * world,
                   "call": "rt_render(scene, camera, params, rgbaOut = page-locked host buffer of rt_host_alloc, ...) from one process"
                           + (f" on a {world}-GPU scene (rt_scene_create_multi): every GPU copies its rows to the host buffer itself" if world > 1 else ""),
                   "device_ms": float(res.stats["deviceMs"]),
                   "note": "scene/BVH resident (built once at engine creation, as the reference builds RTContext before its "
                           "timer, RayTracer.swift:44,166); per step the camera/launch parameters go down and the RGBA8 frame "
                           "+ ray counters come back to host memory; wall clock around the blocking call",
                   "ranks_path": {"value": rays / (ranks_ms * 1e-3) / 1e6, "ms_per_step": ranks_ms,
                                  "call": "ShardedRenderer.render(): one process per GPU (torch.distributed), kernels store into rank 0's "
                                          "frame over NVLink, rank 0 copies it to pinned host memory"}}
        # ---- parity gate + CPU baseline (N = 1 only: the contract's cpu_baseline leg)
        if world == 1 and not args.no_cpu:
            mr, ci = cpu_sample(path, target_s=args.cpu_seconds, want_ids=True)
            want_ids = ci["render"]["ids"] is not None
            got = eng.render(0, want_ids=want_ids)
            parity = parity_gate(got, ci)
            parity["frame_equals_timed_frame"] = bool(frame_hash(got.rgba) == frame_hash_abi == frame_hash_ranks)
            cpu_line = {"value": mr, "unit": UNIT, "cores": ci["cores"], "kind": "port", "sample": ci["sample"], "ms": ci["ms"],
                        "counts_reference_order": {k: int(ci["counters"][k]) for k in ("nodeVisits", "boxTests", "triTests", "sphereTests",
                                                                                      "planeTests", "instEntries", "maxStack")}}
            m1, c1 = cpu_sample(path, target_s=min(10.0, args.cpu_seconds), scene=ci["scene"], threads=1)
            cpu1_line = {"value": m1, "unit": UNIT, "cores": 1, "kind": "port", "sample": c1["sample"], "ms": c1["ms"]}
        eng.close()
        if store is not None:
            store.set("bench_e2e_done", "1")
    elif store is not None:
        store.wait(["bench_e2e_done"])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    roofs = measured_roofs()
    drv = roofs.get("driver", {})
    hbm_peak = float(drv.get("hbm_gbs", 6650.0))
    general = args.variant == 2 or args.config in ("c3glossy", "c3area")
    kernel_name = ("renderKernelGeneral<false, true> (x2: counting + shading pass)" if args.config == "c3area" else
                   "renderKernelGeneral<false, false>") if general else "renderKernelFast<false>"
    # Which roof binds: warp instructions issued per second against SMs x 4 schedulers x the SM clock seen under load.
    # The instruction count of one launch is the ncu figure of the committed capture of this build (profiles/inst_counts.json;
    # smsp__inst_executed.sum — it does not depend on the run); the duration and the clock are this run's.
    inst, inst_src, traffic, traffic_src = None, None, None, None
    try:
        ic = json.load(open(os.path.join(ROOT, "profiles", "inst_counts.json")))
        key = f"{args.config}_n1_v{args.variant}"
        if key in ic:
            # N > 1: ncu must not wrap a multi-rank command, so rank 0's launch is the N = 1 capture scaled by its share of the rays
            share = float(st["rays"]) / float(ic[key]["rays_per_launch"]) if world > 1 else 1.0
            inst, inst_src = ic[key]["warp_inst_per_launch"] * share, ic[key]["source"]
            traffic, traffic_src = ic[key].get("dram_bytes_per_launch") * share, ic[key]["source"]
            if world > 1:
                inst_src = traffic_src = f"ESTIMATE: {share:.4f} (rank 0's share of the frame's rays) x the N = 1 capture; " + inst_src
    except Exception:
        pass
    sm_mhz = clocks.get("sm_mhz") or float(drv.get("sm_max_mhz", 1965.0))
    issue_peak = 148 * 4 * sm_mhz * 1e6 / 1e9                       # G warp-instructions / s
    alg_gbs = my_bytes / (kern_ms * 1e-3) / 1e9
    l2_peak = float(roofs.get("l2_read_gbs_32mb", 0.0)) or None
    roofline = {"bound": "issue", "kernel": kernel_name, "unit": "G warp-inst/s",
                "achieved": (inst / (kern_ms * 1e-3) / 1e9) if inst else None, "peak": issue_peak,
                "frac": (inst / (kern_ms * 1e-3) / 1e9 / issue_peak) if inst else None,
                "peak_source": f"148 SMs x 4 schedulers x {sm_mhz:.0f} MHz (clock sampled in this run); tools/microbench.cu reaches "
                               f"{roofs.get('int_ginst')} of {roofs.get('issue_roof_ginst')} G/s with independent integer ops",
                "warp_inst_per_launch": inst, "warp_inst_source": inst_src, "kernel_ms": kern_ms,
                "traffic": traffic, "traffic_source": traffic_src,
                "hbm": {"bound": "hbm", "achieved": alg_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": alg_gbs / hbm_peak,
                        "peak_source": "measured (MEASURED_PEAKS.json)" if drv else "fallback",
                        "algorithmic_bytes_per_launch": my_bytes,
                        "dram_frac": (traffic / (kern_ms * 1e-3) / 1e9 / hbm_peak) if traffic else None,
                        "note": "SURVEY.md 8(d) algorithmic bytes; they are served by L1/L2 (working set "
                                f"{info['deviceBytes'] / 1e6:.0f} MB against 126 MB of L2), so this fraction can exceed 1 and "
                                "HBM is not the roof that binds; dram_frac = DRAM bytes actually moved / peak"},
                "l2": {"achieved": alg_gbs, "peak": l2_peak, "unit": "GB/s", "frac": (alg_gbs / l2_peak) if l2_peak else None,
                       "peak_source": "tools/microbench.cu: float4 grid-stride reads over a 32 MB buffer (profiles/r2a_microbench.json)"},
                "fp64": {"peak_ginst": roofs.get("dadd_ginst"), "note": "DADD / DMUL / DFMA issue at half the integer rate (tools/microbench.cu)"},
                "counts": {k: int(cst[k]) for k in list(BYTES) + ["maxStack"]},
                "counts_note": "kernel instantiated with counters (same traversal code, COUNT = true), untimed pass"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "dtype_note": "everything that decides a result (ray generation, primitive tests, shading, sums) is FP64 "
                                          "in the reference's operation order; box culling is conservative FP32",
            "data": "synthetic",
            "config": workload_config(args.config, W, H, spp, int(info["triangles"])),
            "detail": {"rays_per_frame": int(rays), "rays_primary": int(rp), "rays_shadow": int(rs),
                       "rays_secondary": int(rays - rp - rs), "parallelism": f"tiles{world}", "gather": R.gather, "shard": R.shard,
                       "kernel_variant": args.variant,
                       "schedule": ("timed frames: tiles drawn longest-first from the per-tile cost recorded on the previous frame "
                                    f"({scheduled} tiles on rank 0; warm-up frames train it)" if scheduled else "spatial order"),
                       "ms_first_frame": first_ms, "ms_per_step_spatial_order": spatial_ms,
                       "l2": ("NOT FLUSHED (diagnostic run); " if args.no_flush else "") + "256 MB buffer written between timed iterations (L2 flush); scene working set "
                             f"{info['deviceBytes'] / 1e6:.0f} MB"},
            "wall_ms_per_step": wall * 1e3 / args.steps, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "frame_hash": frame_hash_abi, "frame_hash_ranks_path": frame_hash_ranks,
            "frame_hash_note": "BLAKE2b-64 of the RGBA8 frame rt_render returned / the multi-process path gathered: equal to each other and for every N",
            "roofline": roofline, "build_ms": info["buildMs"], "upload_ms": info["uploadMs"], "wide_nodes": info.get("wideNodes", 0),
            "bvh_build": ("host binned SAH" if not info["deviceBuiltBlas"] else
                          "device LBVH" if os.environ.get("RTCUDA_BUILD") == "device" else
                          "device binned SAH (the host builder's tree, node array identical)")}
    if parity is not None:
        line["parity"] = parity
    if cpu_line is not None:
        line["cpu_baseline"] = cpu_line
        line["cpu_baseline_1_thread"] = cpu1_line
        line["speedup_measured"] = {"e2e_vs_all_host_threads": e2e["value"] / cpu_line["value"], "host_threads": cpu_line["cores"],
                                    "e2e_vs_one_host_thread": e2e["value"] / cpu1_line["value"]}
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
    ap.add_argument("--build", choices=["auto", "host", "device", "device-sah"], default=None,
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
