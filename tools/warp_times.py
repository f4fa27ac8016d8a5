#!/usr/bin/env python
"""Per-warp timeline of one frame from a -DRT_WARPTIMES build of librtcuda (profiling only):

    make -C myraytracer_b200/csrc OUT=build_wt/librtcuda_wt.so BUILD=build_wt EXTRA=-DRT_WARPTIMES
    RTCUDA_LIB=myraytracer_b200/csrc/build_wt/librtcuda_wt.so python tools/warp_times.py [--config c4] [--shards 8]

Prints when warps start, finish their first unit, draw their last unit and exit, relative to the first warp start.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from myraytracer_b200.engine import RayTracerEngine  # noqa: E402


def pct(a):
    return " ".join(f"{np.percentile(a, q) / 1e3:8.1f}" for q in (0, 10, 50, 90, 99, 100))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c4")
    ap.add_argument("--shards", type=int, default=1)
    ap.add_argument("--shard", type=int, default=0)
    a = ap.parse_args()
    out = "/tmp/warp_times.bin"
    os.environ["RT_WARPTIMES_OUT"] = out
    E = RayTracerEngine(bench.make_scene(a.config, 0), device=0)
    cam = E.camera(0)
    fb = torch.zeros(((cam.height + 7) // 8 * 8, cam.width, 4), dtype=torch.uint8, device="cuda:0")
    st = torch.cuda.current_stream(0).cuda_stream
    for rep in range(4):
        E.render_device(0, d_rgba=fb.data_ptr(), stream=st, shard_index=a.shard, shard_count=a.shards)
        s = E.finish()
    raw = np.fromfile(out, dtype=np.uint64).astype(np.int64)
    info = E.scene_info()
    nw, raw = int(raw[0]), raw[1:]
    t = raw[: nw * 4].reshape(-1, 4)
    uall = raw[nw * 4:].reshape(-1, 4)
    ok = (uall[:, 0] > 0) & (uall[:, 1] > 0)
    u = uall[ok]
    t0 = t[:, 0].min()
    print(f"device {s['deviceMs']:.3f} ms, {len(t)} warps; microseconds after the first warp start, percentiles 0 10 50 90 99 100")
    print("  start        ", pct(t[:, 0] - t0))
    print("  first unit   ", pct(t[:, 1] - t[:, 0]), " (duration)")
    print("  last draw    ", pct(t[:, 2] - t0))
    print("  last unit    ", pct(t[:, 3] - t[:, 2]), " (duration)")
    print("  exit         ", pct(t[:, 3] - t0))
    end = (t[:, 3] - t0).max()
    for frac in (0.5, 0.9, 0.99):
        print(f"  {100 * frac:.0f}% of warps have exited {(end - np.percentile(t[:, 3] - t0, 100 * frac)) / 1e3:.1f} us before the end")


    if len(u):
        d = (u[:, 1] - u[:, 0]) / 1e3
        print(f"units {len(u)}: duration us mean {d.mean():.1f}  p50 {np.percentile(d, 50):.1f} p90 {np.percentile(d, 90):.1f} "
              f"p99 {np.percentile(d, 99):.1f} p99.9 {np.percentile(d, 99.9):.1f} max {d.max():.1f}")
        tilesX = (cam.width + 7) // 8
        idx = np.nonzero(ok)[0]
        visits, vlanes = u[:, 2] >> 32, u[:, 2] & 0xffffffff
        leaves, tests = u[:, 3] >> 32, u[:, 3] & 0xffffffff
        print(f"  per unit: node visits mean {visits.mean():.0f} max {visits.max()}, lanes/visit {vlanes.sum() / max(1, visits.sum()):.1f}; "
              f"leaf visits mean {leaves.mean():.0f} max {leaves.max()}, lane-tri tests mean {tests.mean():.0f} max {tests.max()}")
        A = np.stack([visits, leaves, tests, np.ones_like(visits)], 1).astype(np.float64)
        coef, *_ = np.linalg.lstsq(A, d, rcond=None)
        print(f"  least squares: us = {coef[0]:.4f}*visits + {coef[1]:.4f}*leaves + {coef[2]:.5f}*tests + {coef[3]:.2f}; "
              f"residual rms {np.sqrt(np.mean((A @ coef - d) ** 2)):.1f}")
        row = idx // 16 // tilesX          # tile row (4 pixel rows) within this shard's tile list
        print("  mean / max unit duration per band of 16 tile rows:")
        for b in range(0, row.max() + 1, 16):
            m = (row >= b) & (row < b + 16)
            print(f"    tile rows {b:4d}+  mean {d[m].mean():7.1f}  p99 {np.percentile(d[m], 99):7.1f}  max {d[m].max():7.1f}")
        late = u[:, 1] > (u[:, 1].max() - 150000)
        print(f"  units ending in the last 150 us: {late.sum()}, their durations: mean {d[late].mean():.1f} max {d[late].max():.1f}; "
              f"tile rows {row[late].min()}..{row[late].max()}")
        worst = np.argsort(-d)[:12]
        print("  slowest units (tile row, tile x, pass, us):", [(int(row[k]), int(idx[k] // 16 % tilesX), int(idx[k] % 16), round(float(d[k]), 1), int(visits[k]), int(leaves[k]), int(tests[k])) for k in worst])


if __name__ == "__main__":
    main()
