#!/usr/bin/env python
"""Time the render kernel of every shard of a frame on ONE GPU (rt_render_params.shardIndex/shardCount), to
separate what a rank of an N-GPU run costs by itself (launch ramp, tail, cold caches) from what running N GPUs
at once adds (clocks, the completion fence).

    python tools/shard_times.py [--config c4] [--counts 1,2,4,8] [--reps 5]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from myraytracer_b200.engine import RayTracerEngine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c4")
    ap.add_argument("--counts", default="1,2,4,8")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-flush", action="store_true")
    a = ap.parse_args()
    path = bench.make_scene(a.config, 0)
    E = RayTracerEngine(path, device=0)
    cam = E.camera(0)
    W, H = cam.width, cam.height
    dev = torch.device("cuda", 0)
    fb = torch.zeros(((H + 7) // 8 * 8, W, 4), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(0).cuda_stream
    for n in [int(x) for x in a.counts.split(",")]:
        times = []
        for r in range(n):
            best = []
            for rep in range(a.reps + 2):
                if not a.no_flush:
                    flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                E.render_device(0, d_rgba=fb.data_ptr(), stream=st, shard_index=r, shard_count=n)
                e1.record()
                E.finish()
                torch.cuda.synchronize()
                if rep >= 2:
                    best.append(e0.elapsed_time(e1))
            times.append(sum(best) / len(best))
        print(f"shards {n}: sum {sum(times):.3f} ms, max {max(times):.3f}, min {min(times):.3f}, mean {sum(times) / n:.3f}  "
              + " ".join(f"{t:.3f}" for t in times), flush=True)


if __name__ == "__main__":
    main()
