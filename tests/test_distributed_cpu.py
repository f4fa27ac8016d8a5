"""N>1 path on CPU: chunk ownership, pack/unpack and the rank-0 gather of myraytracer_b200.distributed over gloo
(world_size 2 and 3), plus the bench.py reference arm.  The per-rank "render" is simulated by filling the rows a
rank owns with a rank-independent pattern — the real kernel's output is itself independent of the sharding
(tests/test_gpu_parity.py::test_shards_are_bit_identical)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from myraytracer_b200 import distributed as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pattern(H, W):
    hp = D.n_chunks(H) * D.CHUNK_ROWS
    y = torch.arange(hp, dtype=torch.int64).view(hp, 1, 1)
    x = torch.arange(W, dtype=torch.int64).view(1, W, 1)
    c = torch.arange(4, dtype=torch.int64).view(1, 1, 4)
    return ((y * 31 + x * 7 + c * 3) % 251).to(torch.uint8)


@pytest.mark.parametrize("H,world", [(8, 1), (50, 2), (50, 3), (1080, 8), (7, 4)])
def test_chunk_ownership_partitions_the_frame(H, world):
    seen = []
    for r in range(world):
        seen += list(D.owned_chunks(H, r, world))
        assert len(D.owned_chunks(H, r, world)) <= D.max_owned(H, world)
    assert sorted(seen) == list(range(D.n_chunks(H)))          # every 8-row chunk rendered exactly once


@pytest.mark.parametrize("H,world", [(50, 2), (41, 3), (1080, 8)])
def test_pack_unpack_round_trip(H, world):
    W = 13
    full = _pattern(H, W)
    out = torch.zeros_like(full)
    for r in range(world):
        packed = D.pack_chunks(full, r, world)
        assert packed.shape == (D.max_owned(H, world), D.CHUNK_ROWS * W * 4)
        D.unpack_chunks(out, packed, r, world)
    assert torch.equal(out, full)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, H, W, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = _pattern(H, W)
    fb = torch.zeros_like(full)
    nc = D.n_chunks(H)
    for c in D.owned_chunks(H, rank, world):                   # "render" this rank's chunks only
        fb.view(nc, -1)[c] = full.view(nc, -1)[c]
    D.gather_chunks(fb, rank, world)
    if rank == 0:
        q.put(bool(torch.equal(fb, full)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,H", [(2, 50), (3, 41)])
def test_gather_to_rank0_gloo(world, H):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, H, 16, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_bench_reference_arm_prints_the_contract_line():
    """bench.py --impl reference: the CPU arm (oracle port) prints one JSON line with the agreed keys; a non-zero
    rank exits 0 without work."""
    env = dict(os.environ, RANK="0")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c1", "--steps", "1",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Mrays/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    env = dict(os.environ, RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c1", "--steps", "1",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


@pytest.mark.parametrize("shard", ["rows", "tiles"])
@pytest.mark.parametrize("H,W,world", [(50, 70, 2), (41, 33, 3), (1080, 1920, 8), (7, 5, 4)])
def test_owned_mask_partitions_the_frame(shard, H, W, world):
    total = sum(D.owned_mask(H, W, r, world, shard).astype(int) for r in range(world))
    assert (total == 1).all()                                   # every pixel rendered by exactly one rank
    if shard == "tiles" and H * W > 10000:                      # tile deal: shares equal to within a tile row
        share = [int(D.owned_mask(H, W, r, world, shard).sum()) for r in range(world)]
        assert max(share) - min(share) <= 32 * ((H + 3) // 4)


def test_device_tile_enumeration_matches_owned_mask(tmp_path):
    """ShardTiles (gpu_layout.h, shared by the host code and the kernels) enumerates exactly the tiles of owned_mask,
    each once, for both shard modes; tiles it adds beyond the image (padding of the tile grid) hold no pixel."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "probe.cpp"
    src.write_text('#include <cstdio>\n#include <cstdlib>\n#include "myraytracer_b200/csrc/gpu_layout.h"\n'
                   'int main(int c, char** v) { int W = atoi(v[1]), H = atoi(v[2]), r = atoi(v[3]), n = atoi(v[4]), m = atoi(v[5]);\n'
                   '  rtg::ShardTiles t(W, H, r, n, m);\n'
                   '  for (int k = 0; k < t.count; k++) { int x, y; t.origin(k, x, y); printf("%d %d\\n", x, y); } return 0; }\n')
    exe = tmp_path / "probe"
    subprocess.run(["g++", "-std=c++17", "-I", root, "-I", "/usr/local/cuda/include", str(src), "-o", str(exe)], check=True)
    for (W, H, n) in [(70, 50, 2), (33, 41, 3), (200, 120, 8), (1920, 1080, 8)]:
        for mode, shard in ((0, "rows"), (1, "tiles")):
            for r in range(n):
                out = subprocess.run([str(exe), str(W), str(H), str(r), str(n), str(mode)], capture_output=True, text=True, check=True).stdout
                tiles = [tuple(map(int, ln.split())) for ln in out.splitlines()]
                assert len(set(tiles)) == len(tiles)
                inside = {(x, y) for x, y in tiles if x < W and y < H}
                m = D.owned_mask(H, W, r, n, shard)
                want = {(x, y) for y in range(0, H, 4) for x in range(0, W, 8) if m[y, x]}
                assert inside == want
                assert all(x % 8 == 0 and y % 4 == 0 for x, y in tiles)


# ---------------------------------------------------------------------------------- frame-parallel animation
def _anim_worker(rank, world, port, scene_dir, q):
    """One rank of AnimationRenderer over gloo.  There is no GPU here, so the engine underneath is the emulation build
    of the product sources (tests/emu, test infrastructure) — the frame assignment, the per-frame instance update
    and the stats reduction are the product's."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import emu_lib
    from myraytracer_b200 import scenegen
    from myraytracer_b200.engine import FlatContext
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    kw = dict(width=40, height=32, samples=1)
    shared_dir = scene_dir
    scene_dir = os.path.join(shared_dir, f"rank{rank}")           # the generators write the PLY next to the scene
    os.makedirs(scene_dir, exist_ok=True)
    with emu_lib.use_emulator():
        frames = [0.0, 0.25, 0.5, 0.75, 1.0]
        tf = []
        for f in frames:
            tf.append(FlatContext(scenegen.config_instances(scene_dir, frame=f, name=f"r{rank}f{f}", **kw)).instances()["l2w"])
        A = D.AnimationRenderer(scenegen.config_instances(scene_dir, name=f"r{rank}base", **kw), rank=rank, world=world)
        out_dir = os.path.join(shared_dir, "frames")
        os.makedirs(out_dir, exist_ok=True)
        mine = A.render_frames(np.asarray(tf), out_dir=out_dir)
        stats = A.gather_stats()
        dist.barrier()
        if rank == 0:
            S = D.AnimationRenderer(scenegen.config_instances(scene_dir, name="single", **kw), rank=0, world=1)
            allf = S.render_frames(np.asarray(tf))
            ok = sorted(mine) == list(D.frames_of_rank(len(frames), 0, world))
            ok = ok and all(np.array_equal(mine[f].rgba, allf[f].rgba) for f in mine)
            ok = ok and stats["rays"] == sum(int(r.stats["rays"]) for r in allf.values())
            ok = ok and sorted(os.listdir(out_dir)) == [f"frame_{f:04d}.png" for f in range(len(frames))]
            q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_frames_of_rank_partition():
    for n, world in [(5, 2), (7, 3), (3, 8), (0, 2)]:
        seen = sorted(f for r in range(world) for f in D.frames_of_rank(n, r, world))
        assert seen == list(range(n))


def test_animation_frames_across_ranks_gloo(tmp_path):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_anim_worker, args=(r, world, port, str(tmp_path), q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
