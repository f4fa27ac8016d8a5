"""Tile-sharded rendering of one frame across the GPUs of a box: one process per GPU.

The reference parallelises a frame over 8-row chunks pulled by a CPU thread pool and merges them by plain
row copies (Extensions/Object+Extension.swift:75-82, 285-376).  Every pixel depends only on (i, j), the
camera and the read-only scene (per-pixel PCG32 seed, :294), so the same unit shards across GPUs with no
data-path collective: chunk c belongs to rank c % world (interleaved for load balance), the scene/BVH is
replicated, and the only exchange is the final gather of the RGBA8 rows to rank 0 over NVLink.

torch is plumbing here (device buffers, streams, torch.distributed/NCCL); the render itself is the CUDA
kernel behind the C ABI (rt_render_device).  The chunk bookkeeping (`owned_chunks`, `pack_chunks`,
`gather_chunks`) is backend-agnostic so the world_size>1 path is tested on CPU with gloo.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

CHUNK_ROWS = 8  # Object+Extension.swift:75


def n_chunks(height: int) -> int:
    return (height + CHUNK_ROWS - 1) // CHUNK_ROWS


def owned_chunks(height: int, rank: int, world: int) -> range:
    """Chunks rendered by `rank` (same rule as rt_render_params.shardIndex/shardCount)."""
    return range(rank, n_chunks(height), max(1, world))


def owned_mask(height: int, width: int, rank: int, world: int, shard: str = "rows"):
    """(H, W) bool array of the pixels `rank` renders — the rule of rt_render_params.shardMode (gpu_layout.h ShardTiles):
    "rows": 8-row chunks c with c % world == rank; "tiles": 8x4 tiles (tx, ty) with (tx + ty) % world == rank."""
    import numpy as np
    world = max(1, world)
    y, x = np.mgrid[0:height, 0:width]
    if shard == "tiles" and world > 1:
        return ((x // 8 + y // 4) % world) == rank
    return ((y // CHUNK_ROWS) % world) == rank


def max_owned(height: int, world: int) -> int:
    return (n_chunks(height) + world - 1) // max(1, world)


def pack_chunks(fb: torch.Tensor, rank: int, world: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fb: (n_chunks*8, W, 4) uint8 padded framebuffer -> (max_owned, 8*W*4) contiguous rows of this rank."""
    nc = fb.shape[0] // CHUNK_ROWS
    rows = fb.view(nc, -1)[rank::world]
    mo = (nc + world - 1) // world
    if out is None:
        out = torch.zeros((mo, rows.shape[1]), dtype=fb.dtype, device=fb.device)
    out[: rows.shape[0]].copy_(rows)
    return out


def unpack_chunks(fb: torch.Tensor, packed: torch.Tensor, rank: int, world: int) -> None:
    nc = fb.shape[0] // CHUNK_ROWS
    view = fb.view(nc, -1)[rank::world]
    view.copy_(packed[: view.shape[0]])


def gather_chunks(fb: torch.Tensor, rank: int, world: int, send: Optional[torch.Tensor] = None,
                  recv: Optional[list] = None, group=None) -> None:
    """Gather every rank's chunks into rank 0's `fb` (in place).  Other ranks' `fb` is left unchanged."""
    if world <= 1:
        return
    send = pack_chunks(fb, rank, world, send)
    if rank == 0:
        if recv is None:
            recv = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, gather_list=recv, dst=0, group=group)
        for r in range(1, world):
            unpack_chunks(fb, recv[r], r, world)
    else:
        dist.gather(send, gather_list=None, dst=0, group=group)


class ShardedRenderer:
    """RayTracerEngine on this rank's GPU + the framebuffer exchange.  rank/world default to the
    torch.distributed group (or 0/1 when it is not initialised).

    gather="p2p" (default for world > 1): rank 0 owns the RGBA8 frame (rt_shared_buffer_create) and every rank
    maps it over NVLink (CUDA IPC); each rank's render kernel stores its finished pixels straight into rank 0's
    frame, and one 4-byte all-reduce on the launch stream tells rank 0 that every rank's kernel has completed.
    gather="nccl": each rank renders into a private frame, packs its chunks and rank 0 gathers them (NCCL or gloo).

    shard="rows" (default; required by gather="nccl", whose exchange moves whole rows): rank r renders the 8-row
    chunks c with c % world == r.  shard="tiles" (gather="p2p" only): the 8x4-pixel tiles (tx, ty) with
    (tx + ty) % world == r, which equalises the ray counts of the ranks to 0.1 % — measured on 8 B200 it is not
    faster (the ranks' kernel times differ by GPU, not by content, and the interleave costs ~1 % of L2 locality).
    Within its share each GPU draws the tiles in the order learnt from the previous frame
    (rt_render_params.scheduleMode), which removes the tail of slow packets that would otherwise not shrink with
    the number of GPUs.  (A work counter shared by
    all GPUs over NVLink was tried instead of the static deal: the remote atomic's latency cost more than the
    imbalance it removed.)
    """

    def __init__(self, scene_url: str, device: int = 0, rank: Optional[int] = None, world: Optional[int] = None,
                 camera_index: int = 0, variant: int = 0, gather: str = "p2p", schedule: int = 0,
                 shard: Optional[str] = None):
        import ctypes as C
        from . import capi
        from .engine import RayTracerEngine
        if rank is None:
            rank = dist.get_rank() if dist.is_initialized() else 0
        if world is None:
            world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank, self.world, self.device, self.camera_index = rank, world, device, camera_index
        self.variant = variant  # rt_render_params.kernelVariant (0 = automatic)
        self.schedule = schedule  # rt_render_params.scheduleMode (0 = longest tiles of the previous frame first)
        self.gather = gather if world > 1 else "none"
        self.shard = shard or "rows"
        if self.gather == "nccl" and self.shard != "rows":
            raise ValueError("gather='nccl' exchanges whole rows: it needs shard='rows'")
        self.engine = RayTracerEngine(scene_url, device=device)
        if not self.engine.ready:
            raise RuntimeError("engine not ready: " + self.engine.load_error)
        cam = self.engine.camera(camera_index)
        self.width, self.height = max(1, cam.width), max(1, cam.height)
        dev = torch.device("cuda", device)
        hp = n_chunks(self.height) * CHUNK_ROWS
        self.send = self.recv = self.flag = None
        self._lib, self._shared_ptr, self._opened = capi.load(), None, 0
        self.frames = 0         # frames enqueued so far; the shared frame is double-buffered on its parity
        self._frame_bytes = hp * self.width * 4
        if self.gather == "p2p":
            # Two frames: while rank 0 still copies frame f to the host, the peers may already store pixels of frame
            # f + 1 into the other half; they cannot reach frame f + 2 before rank 0 has passed the fence of f + 1,
            # which its stream orders after that copy.
            ptrv, self._opened = self._share(2 * self._frame_bytes, dev, C, capi)
            ptr = C.c_void_p(ptrv)
            self._shared_ptr = ptr.value
            self.fb_ptr = ptr.value
            self.fb = None
            self.flag = torch.zeros(1, dtype=torch.int32, device=dev)
            if rank == 0:   # rank 0 reads the frames through torch: wrap the raw allocation without copying
                self._fb2 = _wrap_device_memory(ptr.value, (2, hp, self.width, 4), device)
                self.fb = self._fb2[0]
        else:
            self.fb = torch.zeros((hp, self.width, 4), dtype=torch.uint8, device=dev)  # padded to whole chunks
            self.fb_ptr = self.fb.data_ptr()
            if self.gather == "nccl":
                self.send = torch.zeros((max_owned(self.height, world), CHUNK_ROWS * self.width * 4), dtype=torch.uint8, device=dev)
                if rank == 0:
                    self.recv = [torch.empty_like(self.send) for _ in range(world)]
        self.host = torch.empty((self.height, self.width, 4), dtype=torch.uint8, pin_memory=True) if rank == 0 else None

    def _share(self, nbytes, dev, C, capi):
        """rank 0 allocates `nbytes` of zeroed device memory and every other rank maps it (CUDA IPC): (pointer, opened)."""
        handle = torch.zeros(capi.RT_IPC_HANDLE_BYTES, dtype=torch.uint8)
        ptr = C.c_void_p()
        if self.rank == 0:
            hb = (C.c_uint8 * capi.RT_IPC_HANDLE_BYTES)()
            rc = self._lib.rt_shared_buffer_create(self.device, nbytes, C.byref(ptr), hb)
            if rc != 0:
                raise RuntimeError("rt_shared_buffer_create: " + self._lib.rt_last_error().decode())
            handle = torch.tensor(list(hb), dtype=torch.uint8)
        hdev = handle.to(dev)
        dist.broadcast(hdev, src=0)
        opened = 0
        if self.rank != 0:
            hb = (C.c_uint8 * capi.RT_IPC_HANDLE_BYTES)(*hdev.cpu().tolist())
            rc = self._lib.rt_shared_buffer_open(self.device, hb, C.byref(ptr))
            if rc != 0:
                raise RuntimeError("rt_shared_buffer_open: " + self._lib.rt_last_error().decode())
            opened = 1
        return ptr.value, opened

    def close(self):
        if self._shared_ptr:
            torch.cuda.synchronize(self.device)
            if dist.is_initialized() and self.world > 1:
                dist.barrier()          # nobody unmaps / frees while a peer may still write
            self.fb = self._fb2 = None
            self._lib.rt_shared_buffer_close(self.device, self._shared_ptr, self._opened)
            self._shared_ptr = None
        self.engine.close()

    def enqueue(self, counters: bool = False) -> None:
        """Kernel for this rank's chunks + frame exchange towards rank 0, on torch's current stream; no host sync."""
        st = torch.cuda.current_stream(self.device).cuda_stream
        target = self.fb_ptr
        if self.gather == "p2p":
            half = self.frames & 1
            target = self._shared_ptr + half * self._frame_bytes
            if self.rank == 0:
                self.fb = self._fb2[half]           # the frame this call renders
        self.frames += 1
        self.engine.render_device(self.camera_index, d_rgba=target, stream=st, shard_index=self.rank,
                                  shard_count=self.world, counters=counters, variant=self.variant, schedule=self.schedule,
                                  shard_mode=1 if self.shard == "tiles" else 0)
        if self.gather == "p2p":
            dist.all_reduce(self.flag)          # completion fence: returns on-stream once every rank's kernel is done
        elif self.gather == "nccl":
            gather_chunks(self.fb, self.rank, self.world, self.send, self.recv)

    def finish(self) -> dict:
        return self.engine.finish()

    def render(self):
        """One frame end to end: returns (rank 0: (H, W, 4) uint8 numpy view of pinned host memory, stats)."""
        self.enqueue()
        if self.rank == 0:
            self.host.copy_(self.fb[: self.height], non_blocking=True)
        stats = self.finish()
        torch.cuda.current_stream(self.device).synchronize()
        return (self.host.numpy() if self.rank == 0 else None), stats

    @property
    def d2h_bytes(self) -> int:
        return self.height * self.width * 4 if self.rank == 0 else 0


def _wrap_device_memory(ptr: int, shape, device: int) -> torch.Tensor:
    """uint8 torch view of a raw device allocation (no copy, no ownership) via __cuda_array_interface__."""
    class _Raw:
        pass
    r = _Raw()
    r.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(r, device=torch.device("cuda", device))


# ---------------------------------------------------------------------------------------------------------
# Animation: frame-parallel rendering
# ---------------------------------------------------------------------------------------------------------
def frames_of_rank(n_frames: int, rank: int, world: int) -> range:
    """Frames rendered by `rank`: dealt round-robin, so neighbouring (similar-cost) frames land on different GPUs."""
    return range(rank, n_frames, max(1, world))


class AnimationRenderer:
    """Frames of an animation across the GPUs of a box: one process per GPU, WHOLE frames per rank.

    The reference's batch driver (BatchRender/BatchRender.swift:5-59) renders the frame files of a folder one after
    the other, building a new engine (scene parse, RTContext, every BVH) per frame.  Here the scene is resident on
    every GPU; a frame is a set of instance transforms (RayTracerEngine.set_instance_transforms: instance records +
    a device-side TLAS refit), frame f goes to rank f % world and is rendered by one launch on that GPU alone —
    no tile sharding, no framebuffer exchange, no collective on the data path ("replicas"): a 1.5 ms frame shard
    leaves ~0.15 ms of per-launch ramp and tail on the table at 8 GPUs, a whole frame per GPU does not.  Each rank
    writes its own frames (PNG / PPM, write_image) or returns them; `gather_stats` sums the ray counts on rank 0.
    """

    def __init__(self, scene_url: str, device: int = 0, rank: Optional[int] = None, world: Optional[int] = None,
                 camera_index: int = 0):
        from .engine import RayTracerEngine
        if rank is None:
            rank = dist.get_rank() if dist.is_initialized() else 0
        if world is None:
            world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank, self.world, self.camera_index = rank, world, camera_index
        self.engine = RayTracerEngine(scene_url, device=device)
        if not self.engine.ready:
            raise RuntimeError("engine not ready: " + self.engine.load_error)

    def render_frames(self, transforms, out_dir: Optional[str] = None, pattern: str = "frame_{:04d}.png", keep: bool = True):
        """transforms: (n_frames, n_instances, 16) column-major localToWorld.  Renders this rank's frames; returns
        {frame index: RenderResult} (empty when keep=False) and writes out_dir/pattern.format(f) when out_dir is given."""
        import os
        from .engine import write_image
        out = {}
        self.rays = 0
        self.device_ms = 0.0
        for f in frames_of_rank(len(transforms), self.rank, self.world):
            self.engine.set_instance_transforms(transforms[f])
            res = self.engine.render(self.camera_index)
            self.rays += int(res.stats["rays"])
            self.device_ms += float(res.stats["deviceMs"])
            if out_dir is not None:
                write_image(os.path.join(out_dir, pattern.format(f)), res.rgba)
            if keep:
                out[f] = res
        return out

    def gather_stats(self, device=None) -> dict:
        """Sum of rays and max of device time over the ranks (rank 0 gets the totals; works with gloo and NCCL)."""
        t = torch.tensor([float(self.rays), float(self.device_ms)], dtype=torch.float64, device=device or "cpu")
        if self.world > 1 and dist.is_initialized():
            s = t.clone()
            dist.all_reduce(s, op=dist.ReduceOp.SUM)
            m = t.clone()
            dist.all_reduce(m, op=dist.ReduceOp.MAX)
            return dict(rays=int(s[0].item()), device_ms_max=float(m[1].item()), device_ms_sum=float(s[1].item()))
        return dict(rays=int(t[0].item()), device_ms_max=float(t[1].item()), device_ms_sum=float(t[1].item()))

    def close(self):
        self.engine.close()
