"""Run under torchrun with >= 2 GPUs: the sharded frame (both exchange modes) must equal the single-GPU frame
bit for bit (decomposition invariance: the per-pixel RNG depends only on (i, j), Object+Extension.swift:294).
Prints MGPU_OK on rank 0."""
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myraytracer_b200 import RayTracerEngine, scenegen  # noqa: E402
from myraytracer_b200.distributed import ShardedRenderer  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    d = tempfile.mkdtemp(prefix=f"mgpu_r{rank}_")
    path = scenegen.config_c3(d, width=200, height=150, samples=4, depth=3, ico_level=2)
    full = RayTracerEngine(path, device=local).render(0).rgba if rank == 0 else None
    for mode, shard in (("p2p", "tiles"), ("p2p", "rows"), ("nccl", "rows")):
        R = ShardedRenderer(path, device=local, rank=rank, world=world, gather=mode, shard=shard)
        for _ in range(3):
            frame, stats = R.render()
        if rank == 0:
            assert np.array_equal(frame, full), f"{mode}/{shard}: sharded frame differs from the single-GPU frame"
        tot = torch.tensor([stats["raysPrimary"]], dtype=torch.int64, device="cuda")
        dist.all_reduce(tot)
        assert tot.item() == 200 * 150 * 4
        R.close()
    # area lights + motion blur: the jitter index runs through whole 8-row chunks, so row shards reproduce the frame
    ext = scenegen.config_c3(d, width=200, height=150, samples=4, depth=3, ico_level=2, area_lights=2, roughness=0.1)
    full = RayTracerEngine(ext, device=local).render(0).rgba if rank == 0 else None
    for mode in ("p2p", "nccl"):
        R = ShardedRenderer(ext, device=local, rank=rank, world=world, gather=mode, shard="rows")
        for _ in range(2):
            frame, stats = R.render()
        if rank == 0:
            assert np.array_equal(frame, full), f"area lights, {mode}/rows: sharded frame differs from the single-GPU frame"
        R.close()
    # animation: whole frames per rank, no exchange
    from myraytracer_b200.distributed import AnimationRenderer, frames_of_rank
    from myraytracer_b200.engine import FlatContext
    tf = [FlatContext(scenegen.config_instances(d, width=96, height=64, samples=4, frame=f, name=f"f{f}")).instances()["l2w"]
          for f in (0.0, 0.5, 1.0)]
    A = AnimationRenderer(scenegen.config_instances(d, width=96, height=64, samples=4, name="base"), device=local, rank=rank, world=world)
    mine = A.render_frames(np.asarray(tf))
    assert sorted(mine) == list(frames_of_rank(3, rank, world))
    for f, res in mine.items():
        fresh = RayTracerEngine(scenegen.config_instances(d, width=96, height=64, samples=4, frame=(0.0, 0.5, 1.0)[f], name=f"g{f}"),
                                device=local).render(0).rgba
        assert np.array_equal(res.rgba, fresh), f"animation frame {f} on rank {rank}"
    st = A.gather_stats(device="cuda")
    assert st["rays"] > 0
    A.close()
    dist.barrier()
    if rank == 0:
        print("MGPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
