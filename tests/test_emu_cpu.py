"""Kernel logic on the CPU: the product's CUDA sources compiled for the host against the SIMT emulation shim
(tests/emu: every CUDA thread a fiber, the lanes of a warp meeting at each *_sync intrinsic) and compared with the
oracle through the same C ABI and the same parity bars as tests/test_gpu_parity.py, on scenes small enough for
seconds.  This checks what a container without a GPU can check — arithmetic, control flow, warp-collective
protocol, work distribution, the tile schedule kernels — and is test infrastructure only: the emulation library is
never loaded by myraytracer_b200 (tests/emu_lib.py).  The -m gpu tests remain the parity tests proper."""
import numpy as np
import pytest

import emu_lib
import oracle_lib
import parity
from myraytracer_b200 import RayTracerEngine, scenegen


@pytest.fixture()
def emu():
    with emu_lib.use_emulator() as lib:
        yield lib


CASES = {
    # name: (generator, kwargs, kernel variants to run: 1 fast (warp packets), 2 general)
    "c1": (scenegen.config_c1, dict(width=72, height=56, samples=1), (1, 2)),
    "c1x4_ragged": (scenegen.config_c1, dict(width=50, height=37, samples=4), (1, 2)),      # image not a multiple of 8x4
    "c1x9": (scenegen.config_c1, dict(width=40, height=32, samples=9), (1, 2)),             # whole-tile units
    "c1x64": (scenegen.config_c1, dict(width=24, height=16, samples=64), (1,)),             # one-pixel units
    "c2": (scenegen.config_c2, dict(n=24, width=64, height=48), (1, 2)),
    "c3": (scenegen.config_c3, dict(width=64, height=48, samples=4, depth=4, ico_level=2), (1, 2)),
    "c3dof": (scenegen.config_c3, dict(width=48, height=40, samples=4, depth=3, ico_level=2, aperture=0.35, focus=13.0), (1, 2)),
    "c3glossy": (scenegen.config_c3, dict(width=48, height=40, samples=4, depth=3, ico_level=2, roughness=0.12), (2,)),
    # the warp's ray queue under every unit rule: 1 spp (one pass), 9 spp (whole tile = 9 passes, three batches), 16 spp
    # (four units of four passes per tile), 64 spp (one pixel = two passes)
    "c3x1": (scenegen.config_c3, dict(width=48, height=40, samples=1, depth=4, ico_level=2), (1,)),
    "c3x9": (scenegen.config_c3, dict(width=24, height=20, samples=9, depth=3, ico_level=1), (1,)),
    "c3x16": (scenegen.config_c3, dict(width=24, height=16, samples=16, depth=3, ico_level=1), (1,)),
    "c3x64": (scenegen.config_c3, dict(width=16, height=8, samples=64, depth=2, ico_level=1), (1,)),
    "c5": (scenegen.config_c5, dict(n=12, width=64, height=40, samples=4, meshes=3), (1, 2)),
    "instances": (scenegen.config_instances, dict(width=64, height=48, samples=4), (1, 2)),
    # area lights: counting pass, per-chunk prefix, shading pass (general kernel); ragged image, glossy recursion
    "c3area": (scenegen.config_c3, dict(width=64, height=48, samples=4, depth=3, ico_level=2, area_lights=2), (2,)),
    # motion blur: ray time down the path; mesh motion in the triangle test; instance motion through the exact walk of
    # the reference-topology TLAS
    "motion_mesh": (scenegen.config_instances, dict(width=64, height=48, samples=4, mesh_motion=(0.5, 0.2, -0.3)), (2,)),
    "motion_instances": (scenegen.config_instances, dict(width=64, height=48, samples=4, instance_motion=1.0), (2,)),
    "motion_both_ragged": (scenegen.config_instances, dict(width=50, height=37, samples=9, mesh_motion=(0.4, -0.3, 0.2),
                                                           instance_motion=1.5), (2,)),
    "c3area_glossy_ragged": (scenegen.config_c3, dict(width=50, height=37, samples=1, depth=4, ico_level=2, area_lights=3,
                                                      roughness=0.1), (2,)),
    # an area light of infinite radiance: the non-finite guard (:277-280) fires at most shading points; the discarded
    # levels' children still advance the chunk's jitter index and the pixel's RNG like the reference's do
    "c3area_glossy_inf": (scenegen.config_c3, dict(width=40, height=30, samples=4, depth=3, ico_level=1, area_lights=2,
                                                   roughness=0.15, inf_light=True), (2,)),
    "c3_inf": (scenegen.config_c3, dict(width=40, height=30, samples=4, depth=3, ico_level=1, inf_light=True), (0, 2)),
}


def check(got, ref):
    rep = parity.compare(got, ref)
    assert rep["alpha_ok"]
    assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0, rep
    assert rep["px_within_1"] >= 0.999, rep
    assert rep["rays_gpu"][0] == rep["rays_ref"][0], rep
    for a, b in zip(rep["rays_gpu"][1:], rep["rays_ref"][1:]):
        assert abs(a - b) <= max(4, 1e-4 * b), rep
    return rep


@pytest.mark.parametrize("name", sorted(CASES))
def test_emulated_kernels_vs_oracle(name, scene_dir, emu):
    fn, kw, variants = CASES[name]
    path = fn(scene_dir, **kw)
    eng = RayTracerEngine(path)
    assert eng.ready, eng.load_error
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    before = emu.rtemu_kernel_launches()
    for variant in variants:
        got = eng.render(0, want_ids=True, want_linear=True, variant=variant, counters=(variant == 2))
        rep = check(got, ref)
        print(name, variant, rep)
    assert emu.rtemu_kernel_launches() > before       # the kernels ran (on fibers), not some other path


def test_emulated_first_frame_takes_its_tile_order_from_probe_rays(scene_dir, emu, monkeypatch):
    """RTCUDA_PROBE=1 (opt-in), first frame of a geometry with >= 256 tiles: one probe ray per tile (tileProbeKernel)
    estimates the tile costs and the frame already draws expensive tiles first; the image is the spatial order's."""
    path = scenegen.config_c2(scene_dir, n=24, width=128, height=64)
    n_tiles = 16 * 16
    ref = RayTracerEngine(path).render(0, schedule=1)
    off = RayTracerEngine(path).render(0)
    assert off.stats["tilesScheduled"] == 0 and np.array_equal(off.rgba, ref.rgba)         # default: spatial order
    monkeypatch.setenv("RTCUDA_PROBE", "1")
    eng = RayTracerEngine(path)
    first = eng.render(0)
    assert first.stats["tilesScheduled"] == n_tiles and first.stats["kernelLaunches"] == 6
    assert np.array_equal(first.rgba, ref.rgba)
    again = eng.render(0)
    assert again.stats["tilesScheduled"] == n_tiles and again.stats["kernelLaunches"] == 3
    assert np.array_equal(again.rgba, ref.rgba)


@pytest.mark.parametrize("name", ["c3", "c5", "instances"])
def test_emulated_queue_rounds_walk_big_meshes_per_lane(name, scene_dir, emu, monkeypatch):
    """Child rays off the warp's queue walk meshes above kBigBlasTriangles per lane instead of as a packet
    (rt_packet.cuh: packetInstance); RTCUDA_BIG_BLAS=0 makes every mesh "big": same parity bars, same ray counts."""
    monkeypatch.setenv("RTCUDA_BIG_BLAS", "0")
    fn, kw, _ = CASES[name]
    path = fn(scene_dir, **kw)
    eng = RayTracerEngine(path)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    got = eng.render(0, want_ids=True, want_linear=True, variant=1)
    check(got, ref)
    monkeypatch.delenv("RTCUDA_BIG_BLAS")
    base = RayTracerEngine(path).render(0, want_linear=True, variant=1)
    assert (base.rgba == got.rgba).all(axis=-1).mean() >= 0.9999      # other visiting order inside a mesh: ties only


@pytest.mark.parametrize("name", ["c3", "c3x9", "c3x64", "c5", "instances"])
def test_emulated_two_launch_frames_defer_units_with_child_rays(name, scene_dir, emu, monkeypatch):
    """RTCUDA_SPLIT=1 (the default for scenes with recursion and big meshes): depth 0 of every unit in the lean instantiation,
    units whose primary hits spawn child rays deferred to a second launch of the queue instantiation.  Same frame bit for
    bit, same ray counts (a deferred unit's rays are counted once), one more kernel launch; under every unit rule."""
    fn, kw, _ = CASES[name]
    path = fn(scene_dir, **kw)
    monkeypatch.setenv("RTCUDA_SPLIT", "0")
    one = RayTracerEngine(path).render(0, want_ids=True, want_linear=True, variant=1, schedule=1, counters=True)
    monkeypatch.setenv("RTCUDA_SPLIT", "1")
    two = RayTracerEngine(path).render(0, want_ids=True, want_linear=True, variant=1, schedule=1, counters=True)
    assert np.array_equal(one.rgba, two.rgba) and np.array_equal(one.linear, two.linear) and np.array_equal(one.ids, two.ids)
    for k in ("raysPrimary", "raysShadow", "raysSecondary", "boxTests", "triTests", "sphereTests", "planeTests", "instEntries"):
        assert one.stats[k] == two.stats[k], k          # the tests of a deferred unit's first attempt are not counted either
    assert two.stats["kernelLaunches"] == one.stats["kernelLaunches"] + 1


def test_emulated_edge_cases(scene_dir, emu):
    """tests/edge_cases.py: empty scene (no TLAS: black, not background), one-instance TLAS, 1x1 / 9x5 / 33x17 images,
    sample counts that are not squares (n = floor(sqrt), sum divided by the count), depth 0 with a mirror — both kernels,
    linear colours identical to the oracle's."""
    import edge_cases
    for name in edge_cases.EDGE_CASES:
        path = edge_cases.make(scene_dir, name)
        eng = RayTracerEngine(path)
        assert eng.ready, (name, eng.load_error)
        ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
        for variant in (1, 2):
            got = eng.render(0, want_ids=True, want_linear=True, variant=variant)
            rep = check(got, ref)
            assert rep["px_exact"] == 1.0 and rep["rays_gpu"] == rep["rays_ref"], (name, variant, rep)
            assert np.allclose(got.linear, ref["linear"], rtol=1e-12, atol=1e-9), (name, variant)


def test_emulated_shards_and_schedule(scene_dir, emu):
    """Both shard rules partition the frame bit-identically; the learnt tile order does not change it."""
    from myraytracer_b200.distributed import owned_mask
    path = scenegen.config_c3(scene_dir, width=72, height=52, samples=4, depth=3, ico_level=2)
    eng = RayTracerEngine(path)
    full = eng.render(0, schedule=1)
    for mode in ("rows", "tiles"):
        acc = np.zeros_like(full.rgba)
        for k in range(3):
            part = eng.render(0, shard_index=k, shard_count=3, shard_mode=1 if mode == "tiles" else 0).rgba
            own = owned_mask(52, 72, k, 3, mode)
            assert (part[~own] == 0).all()
            acc[own] = part[own]
        assert np.array_equal(acc, full.rgba)
    for variant in (1, 2):
        first = eng.render(0, variant=variant)
        assert first.stats["tilesScheduled"] == 0
        again = eng.render(0, variant=variant)
        n_tiles = 9 * 7 * 2
        assert again.stats["tilesScheduled"] == n_tiles and again.stats["kernelLaunches"] == 3
        assert sorted(eng.schedule().tolist()) == list(range(n_tiles))
        assert np.array_equal(first.rgba, full.rgba) and np.array_equal(again.rgba, full.rgba)


@pytest.mark.parametrize("kw", [dict(width=64, height=48, samples=4, depth=3, ico_level=2, area_lights=2),
                                dict(width=50, height=37, samples=9, depth=4, ico_level=1, area_lights=3),
                                dict(width=40, height=30, samples=4, depth=3, ico_level=1, area_lights=2, inf_light=True)])
def test_emulated_area_light_counting_pass_on_the_fast_kernel(kw, scene_dir, emu, monkeypatch):
    """Pass 1 of an area-light frame (each pixel's number of jitter-table draws) runs on the fast kernel when no material is
    glossy and nothing moves (renderKernelFast<.., AREACOUNT>: packets + ray queue, no shadow rays); the general kernel's
    counting walk (RTCUDA_AREA_COUNT_GENERAL=1) must give the same frame bit for bit, and both the oracle's."""
    path = scenegen.config_c3(scene_dir, **kw)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    fast = RayTracerEngine(path).render(0, want_ids=True, want_linear=True)
    check(fast, ref)
    monkeypatch.setenv("RTCUDA_AREA_COUNT_GENERAL", "1")
    gen = RayTracerEngine(path).render(0, want_ids=True, want_linear=True)
    assert np.array_equal(fast.linear, gen.linear) and np.array_equal(fast.rgba, gen.rgba)
    assert fast.stats["rays"] == gen.stats["rays"]


def test_emulated_area_lights_shard_by_chunks(scene_dir, emu):
    """The jitter index of the area lights runs through an 8-row chunk (Object+Extension.swift:288): shards made of
    whole chunks reproduce the frame bit for bit; tile shards and the fast kernel are refused, not approximated."""
    from myraytracer_b200 import RenderError
    from myraytracer_b200.distributed import owned_mask
    path = scenegen.config_c3(scene_dir, width=40, height=44, samples=4, depth=2, ico_level=1, area_lights=2)
    eng = RayTracerEngine(path)
    full = eng.render(0, want_linear=True)
    assert full.stats["kernelLaunches"] == 5             # count pass, prefix, frame, two schedule kernels
    acc = np.zeros_like(full.linear)
    for k in range(3):
        part = eng.render(0, want_linear=True, shard_index=k, shard_count=3).linear
        own = owned_mask(44, 40, k, 3, "rows")
        acc[own] = part[own]
    assert np.array_equal(acc, full.linear)
    assert np.array_equal(eng.render(0, want_linear=True).linear, full.linear)       # learnt tile order: same frame
    for kw in (dict(shard_index=0, shard_count=2, shard_mode=1), dict(variant=1)):
        with pytest.raises(RenderError) as e:
            eng.render(0, **kw)
        assert e.value.code == -3


def test_emulated_motion_blur_changes_the_frame_and_refuses_the_fast_kernel(scene_dir, emu):
    from myraytracer_b200 import RenderError
    still = RayTracerEngine(scenegen.config_instances(scene_dir, width=48, height=36, samples=4)).render(0).rgba
    path = scenegen.config_instances(scene_dir, width=48, height=36, samples=4, instance_motion=1.5)
    eng = RayTracerEngine(path)
    moved = eng.render(0).rgba
    assert (moved != still).any(axis=-1).mean() > 0.02           # the motion is visible
    with pytest.raises(RenderError) as e:
        eng.render(0, variant=1)
    assert e.value.code == -3


def test_emulated_animation_refit_matches_rebuilt_scenes(scene_dir, emu):
    """rt_scene_update_instances: new instance transforms on a resident scene (records re-derived, TLAS refitted by the
    device kernel, topology kept) give the frame of a scene built from scratch with those transforms — and the
    oracle's — for every frame, and going back to frame 0 gives frame 0."""
    from myraytracer_b200.engine import FlatContext
    kw = dict(width=64, height=48, samples=4)
    eng = RayTracerEngine(scenegen.config_instances(scene_dir, **kw))
    base = eng.render(0, want_linear=True)
    for f in (0.5, 1.0, 0.0):
        pf = scenegen.config_instances(scene_dir, frame=f, name=f"frame{f}", **kw)
        want = FlatContext(pf).instances()
        # Instance.worldBounds of the meshes derived by the device kernel (default), then by the host loop of
        # refitTLAS step 1: the same doubles as a context flattened from scratch, either way
        for on_device in (True, False):
            eng.set_instance_transforms(want["l2w"], bounds_on_device=on_device)
            d = eng.describe()
            bounds = np.array([list(d.instances[i].worldBounds) for i in range(d.nInstances)])
            assert np.array_equal(bounds, want["bounds"]), on_device
        got = eng.render(0, want_ids=True, want_linear=True)
        check(got, oracle_lib.OracleScene(pf).render(want_ids=True, want_linear=True))
        fresh = RayTracerEngine(pf).render(0, want_linear=True)
        assert np.array_equal(fresh.linear, got.linear)
        assert (f == 0.0) == np.array_equal(got.linear, base.linear)
    with pytest.raises(Exception):
        eng.set_instance_transforms(np.zeros((2, 16)))          # the instance count of a scene is fixed


def test_image_writers_round_trip(tmp_path, emu):
    """rth_image_write: PNG (8-bit RGBA, stored deflate, CRCs, Adler-32) decodes to the frame; PPM is P6 RGB."""
    import struct
    import zlib
    from myraytracer_b200 import write_image
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, size=(37, 301, 4), dtype=np.uint8)          # > 65535 raw bytes: several stored blocks
    png, ppm = str(tmp_path / "a.png"), str(tmp_path / "a.ppm")
    write_image(png, img)
    write_image(ppm, img)
    raw = open(png, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, seen = 8, b"", []
    while pos < len(raw):
        n, kind = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(kind + body)
        seen.append(kind)
        if kind == b"IHDR":
            assert struct.unpack(">IIBBBBB", body) == (301, 37, 8, 6, 0, 0, 0)
        if kind == b"IDAT":
            idat += body
        pos += 12 + n
    assert seen == [b"IHDR", b"IDAT", b"IEND"]
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(37, 301 * 4 + 1)
    assert (rows[:, 0] == 0).all() and np.array_equal(rows[:, 1:].reshape(37, 301, 4), img)
    p = open(ppm, "rb").read()
    head = b"P6\n301 37\n255\n"
    assert p.startswith(head) and np.array_equal(np.frombuffer(p[len(head):], np.uint8).reshape(37, 301, 3), img[..., :3])
    with pytest.raises(Exception):
        write_image(str(tmp_path / "a.jpg"), img)


def test_emulated_ply_ingest_equals_host_flatten(tmp_path, emu, monkeypatch):
    """rt_ply_ingest (raw PLY blocks -> kernels -> per-triangle FP64 arrays) against the host loader + flatten loops of
    the mirror, which tests/test_host_cpu.py in turn holds against the oracle's: same vertices and normals, bit for bit
    (NaN normals of unused vertices included), for every layout the device path takes; the others go to the host."""
    import ply_cases
    from myraytracer_b200.engine import FlatContext
    for name, path, n_tris, device in ply_cases.make_cases(str(tmp_path)):
        monkeypatch.setenv("RTCUDA_INGEST", "host")
        hv, hn = FlatContext(path).triangles()
        before = emu.rtemu_kernel_launches()
        monkeypatch.setenv("RTCUDA_INGEST", "device")
        dv, dn = FlatContext(path).triangles()
        ran = emu.rtemu_kernel_launches() - before
        assert len(hv) == n_tris and len(dv) == n_tris, name
        assert np.array_equal(hv, dv) and np.array_equal(hn, dn, equal_nan=True), name
        assert (ran >= 2) == device, (name, ran)          # the kernels ran / the file was left to the host loader


def test_emulated_ingest_flatten_of_a_big_mesh_equals_host_flatten(tmp_path, emu, monkeypatch):
    """Above 32,768 triangles the flatten of a device-ingested mesh runs on all host threads over unwritten
    (default-initialised) arrays (rt_context.cpp: forParts, touchPages; raytracer_host.hpp: DefaultInitAllocator):
    every array of the description — vertices, normals, motion vectors, prim records — and the instance's matrices
    and world box (a threaded min / max reduction) equal the one-thread host flatten's, bit for bit, with a second
    mesh flattened on the host before and after it so the ingested block lands in the middle of the arrays."""
    import os
    import ctypes as C
    from myraytracer_b200.engine import FlatContext
    pos, tris = scenegen.terrain_mesh(131, 5, amp=1.3)                   # 2 * 130^2 = 33,800 triangles
    assert len(tris) > 32768
    scenegen.write_ply(os.path.join(str(tmp_path), "big.ply"), pos, tris)
    sp, st = scenegen.icosphere_mesh(1, radius=0.7)
    verts = [tuple(map(float, v)) for v in sp]
    sc = scenegen.Scene(maxRecursionDepth=0)
    sc.vertexData = verts
    sc.materials = [scenegen._plain(1, (0.5, 0.5, 0.5)), scenegen._plain(2, (0.2, 0.5, 0.8))]
    M = scenegen.mat_translate((1.5, -0.25, 3.0)) @ scenegen.mat_rotate((0.3, 1.0, 0.2), 37.0) @ scenegen.mat_scale((1.2, 0.8, 1.1))
    sc.objects = [
        scenegen.SceneObject(scenegen.OBJ_MESH, id=1, material=2, smooth=False, faces=(st + 1).astype(np.int32)),
        scenegen.SceneObject(scenegen.OBJ_MESH, id=2, material=1, smooth=True, plyPath="big.ply", transform=scenegen.colmajor(M),
                             motionBlur=(0.25, -0.5, 0.125)),
        scenegen.SceneObject(scenegen.OBJ_MESH, id=3, material=1, smooth=True, faces=(st + 1).astype(np.int32)),
    ]
    sc.pointLights = [scenegen.PointLight((8.0, 9.0, 8.0), (1000, 1000, 1000))]
    sc.cameras = [scenegen.Camera(id=1, imageName="big.png", width=16, height=16, numSamples=1, position=(0.0, 12.0, 19.0),
                                  gaze=(0.0, -0.68, -1.0), up=(0, 1, 0), nearPlane=(-0.4, 0.4, -0.4, 0.4), nearDistance=1.0)]
    path = os.path.join(str(tmp_path), "big.rtscene")
    scenegen.write_rtscene(sc, path)

    def arrays(ctx):
        d = ctx.desc
        v, n = ctx.triangles()
        mb = np.ctypeslib.as_array(d.triMotionBlur, shape=(d.nTriangles, 3)).copy()
        prims = np.frombuffer(C.string_at(d.prims, d.nPrims * C.sizeof(d.prims._type_)), np.uint8).copy()
        blas = [(d.blas[b].firstPrim, d.blas[b].primCount) for b in range(d.nBlas)]
        return v, n, mb, prims, blas, ctx.instances()

    monkeypatch.setenv("RTCUDA_INGEST", "host")
    hv, hn, hmb, hpr, hbl, hin = arrays(FlatContext(path))
    before = emu.rtemu_kernel_launches()
    monkeypatch.setenv("RTCUDA_INGEST", "device")
    dv, dn, dmb, dpr, dbl, din = arrays(FlatContext(path))
    assert emu.rtemu_kernel_launches() - before >= 2                     # the ingest kernels ran
    assert len(hv) == len(tris) + 2 * len(st) and len(dv) == len(hv)
    assert np.array_equal(hv, dv) and np.array_equal(hn, dn, equal_nan=True)
    assert np.array_equal(hmb, dmb) and np.array_equal(hpr, dpr) and hbl == dbl
    assert (dmb[len(st):len(st) + len(tris)] == (0.25, -0.5, 0.125)).all() and (dmb[:len(st)] == 0).all()
    for k in hin:
        assert np.array_equal(hin[k], din[k]), k
    assert np.isfinite(din["bounds"]).all()


def test_emulator_is_not_the_product(emu):
    """The product loader never hands out the emulation library, and the emulation build says what it is."""
    import os
    from myraytracer_b200 import capi
    assert os.path.basename(capi.LIB_PATH) == "librtcuda.so"
    assert "emu" not in capi.LIB_PATH
    assert hasattr(emu, "rtemu_kernel_launches")


SAH_CASES = {
    # 3,042 triangles: the top levels run the blocks-of-one-node path (shared-memory bins), the deep ones the mixed path
    "c2": (scenegen.config_c2, dict(n=40, width=64, height=48)),
    "c3": (scenegen.config_c3, dict(width=48, height=40, samples=4, depth=3, ico_level=2)),
    "c5": (scenegen.config_c5, dict(n=12, width=64, height=40, samples=4, meshes=3)),
    "instances": (scenegen.config_instances, dict(width=64, height=48, samples=4)),
    "motion_mesh": (scenegen.config_instances, dict(width=48, height=32, samples=4, mesh_motion=(0.5, 0.2, -0.3))),
}


@pytest.mark.parametrize("name", sorted(SAH_CASES))
def test_emulated_device_sah_build_makes_the_host_builders_tree(name, scene_dir, emu, monkeypatch):
    """RT_BUILD_DEVICE_SAH (csrc/bvh_build_sah_gpu.cu): the level-by-level device builder takes the host builder's
    decisions and numbers the nodes as it does, so the node array on the device is the host builder's byte for byte —
    the number of box tests of a frame is identical, the image is identical bit for bit — and only the triangle order
    inside a leaf (hence the any-hit triangle-test count of the
    shadow rays) may differ."""
    fn, kw = SAH_CASES[name]
    path = fn(scene_dir, **kw)
    monkeypatch.setenv("RTCUDA_BUILD", "host")
    host = RayTracerEngine(path)
    assert host.scene_info()["deviceBuiltBlas"] == 0
    a = host.render(0, want_ids=True, want_linear=True, variant=2, counters=True)
    monkeypatch.setenv("RTCUDA_BUILD", "device-sah")
    before = emu.rtemu_kernel_launches()
    dev = RayTracerEngine(path)
    assert emu.rtemu_kernel_launches() > before
    hi, di = host.scene_info(), dev.scene_info()
    assert di["deviceBuiltBlas"] >= 1
    for key in ("blasNodes", "tlasNodes", "maxLeafSize", "maxDepth", "triangles"):
        assert hi[key] == di[key], (key, hi, di)
    assert np.array_equal(host.blas_nodes(), dev.blas_nodes())       # boxes, child references and node numbers
    b = dev.render(0, want_ids=True, want_linear=True, variant=2, counters=True)
    assert a.stats["boxTests"] == b.stats["boxTests"] and a.stats["maxStack"] == b.stats["maxStack"]
    assert abs(a.stats["triTests"] - b.stats["triTests"]) <= 0.01 * a.stats["triTests"]
    assert np.array_equal(a.rgba, b.rgba) and np.array_equal(a.linear, b.linear)
    assert (a.ids["prim"] == b.ids["prim"]).mean() >= 0.9999
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    check(b, ref)
    if not name.startswith("motion"):        # the fast kernel does not carry a ray time
        check(dev.render(0, want_ids=True, want_linear=True, variant=1), ref)


def test_emulated_device_sah_build_degenerate_meshes(scene_dir, emu, monkeypatch):
    """Coincident centroids (a stack of identical triangles: no plane separates them, the list is halved until the
    leaves hold <= 4), two triangles, one triangle, and a mesh whose coordinates need FP64 records."""
    import ply_cases
    rng = np.random.default_rng(5)
    base = np.array([[-1.0, -1.0, 0.0], [1.0, -1.0, 0.0], [0.0, 1.0, 0.0]])
    meshes = {
        "stack": (np.tile(base, (11, 1)), np.arange(33).reshape(11, 3)),
        "pair": (np.array([[-1.0, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]]), np.array([[0, 1, 2], [0, 2, 3]])),
        "single": (base, np.array([[0, 1, 2]])),
        "scatter": (rng.uniform(-2, 2, (90, 3)), np.arange(90).reshape(30, 3)),
    }
    for name, (pos, tris) in meshes.items():
        scenegen.write_ply(f"{scene_dir}/{name}.ply", pos.astype(np.float32), tris.astype(np.int32))
        path = ply_cases._mesh_scene(scene_dir, name, f"{name}.ply", smooth=False)
        out = {}
        for mode in ("host", "device-sah"):
            monkeypatch.setenv("RTCUDA_BUILD", mode)
            eng = RayTracerEngine(path)
            assert eng.ready, (name, mode, eng.load_error)
            out[mode] = (eng.render(0, want_ids=True, want_linear=True, variant=2, counters=True), eng.scene_info(), eng.blas_nodes())
        (a, hi, _), (b, di, _) = out["host"], out["device-sah"]
        assert di["deviceBuiltBlas"] == (1 if len(tris) >= 2 else 0), (name, di)
        assert di["maxLeafSize"] <= 4
        assert np.array_equal(a.linear, b.linear), name
        # the same distance everywhere; which of several coincident triangles reports it may differ
        assert np.array_equal(a.ids["t"], b.ids["t"]) and np.array_equal(a.ids["inst"], b.ids["inst"]), name
        if name != "stack":
            assert a.stats["boxTests"] == b.stats["boxTests"], name
            assert np.array_equal(out["host"][2], out["device-sah"][2]), name
        check(b, oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)) if name != "stack" else None


def test_emulated_c_abi_error_behaviour(scene_dir, emu):
    """Every refusal is loud and leaves the handle usable: the reference's own error cases (RayTracer.swift:160-168:
    no renderer / no scene / invalid camera index) and this module's argument checks (include/rtcuda.h)."""
    import ctypes as C
    from myraytracer_b200 import capi
    from myraytracer_b200.engine import FlatContext, RenderError
    path = scenegen.config_c3(scene_dir, width=24, height=16, samples=1, depth=2, ico_level=1, roughness=0.1, area_lights=1)
    eng = RayTracerEngine(path)
    good = eng.render(0)

    def code(fn):
        with pytest.raises(RenderError) as e:
            fn()
        assert str(e.value), "an error without a message"
        return e.value.code

    assert code(lambda: eng.render(7)) == -10                                   # Invalid camera index
    assert code(lambda: eng.render(-1)) == -10
    assert code(lambda: eng.render(0, variant=9)) == -1
    assert code(lambda: eng.render(0, variant=1)) == -3                         # glossy / area lights on the fast kernel
    assert code(lambda: eng.render(0, schedule=5)) == -1
    assert code(lambda: eng.render(0, shard_index=1, shard_count=2, shard_mode=1)) == -3   # area lights, tile shards
    assert code(lambda: eng.render(0, shard_index=2, shard_count=2)) == -1      # shard index outside the count
    assert code(eng.finish) == -1                                               # no frame in flight
    assert np.array_equal(eng.render(0).rgba, good.rgba)                        # the handle still renders

    # descriptions that cannot be a scene
    ctx = FlatContext(path)
    lib = capi.load()
    for field, value, want in (("abiVersion", 999, "abiVersion"), ("buildMode", 7, "buildMode"), ("nMaterials", 0, "material")):
        d = capi.rt_scene_desc.from_buffer_copy(ctx.desc)
        setattr(d, field, value)
        assert not lib.rt_scene_create(C.byref(d), 0)
        assert want in lib.rt_last_error().decode()
    d = capi.rt_scene_desc.from_buffer_copy(ctx.desc)
    assert not lib.rt_scene_create(C.byref(d), 3)                               # device index outside the count
    assert not lib.rt_scene_create(None, 0)
    h = lib.rt_scene_create(C.byref(ctx.desc), 0)
    assert h
    info = capi.rt_scene_info()
    assert lib.rt_scene_get_info(h, C.byref(info)) == 0 and info.instances == ctx.desc.nInstances
    assert lib.rt_scene_update_instances(h, ctx.desc.instances, ctx.desc.nInstances + 1, None) == -1   # another instance count
    assert lib.rt_render(h, None, None, None, None, None, None) == -10
    cam = eng.camera(0)
    assert lib.rt_render(h, C.byref(cam), None, None, None, None, None) == -1   # nowhere to put the frame
    lib.rt_scene_destroy(h)
    lib.rt_scene_destroy(None)
    ctx.close()

    # an engine without a scene file: not ready, renders report the reference's message
    bad = RayTracerEngine(f"{scene_dir}/missing.rtscene")
    assert not bad.ready and bad.load_error
    assert code(lambda: bad.render(0)) in (-20, -21)


@pytest.mark.parametrize("seed", [1, 0xC0FFEE])
def test_emulated_results_do_not_depend_on_thread_order(seed, scene_dir, tmp_path, emu, monkeypatch):
    """On the device there is no thread order.  The emulator can impose different ones (rtemu_set_schedule_seed: blocks
    of a grid and runnable fibers of a block resumed in shuffled order, so atomics land differently); everything
    that goes through atomics must come out the same: the device SAH builder's node array, the PLY ingest's arrays
    (adjacency lists filled by atomic cursors, then sorted), instance bounds and TLAS refit, the area lights' counting
    pass and prefix, the frames."""
    import ply_cases
    from myraytracer_b200.engine import FlatContext
    path = scenegen.config_c3(scene_dir, width=40, height=24, samples=4, depth=3, ico_level=2, area_lights=2)
    mesh = scenegen.config_c2(scene_dir, n=24, width=40, height=24)
    inst = scenegen.config_instances(scene_dir, width=40, height=24, samples=1)

    def run():
        out = {}
        monkeypatch.setenv("RTCUDA_BUILD", "device-sah")
        e = RayTracerEngine(mesh)
        out["nodes"] = e.blas_nodes()
        out["mesh"] = e.render(0, want_linear=True).linear
        monkeypatch.setenv("RTCUDA_BUILD", "host")
        out["area"] = RayTracerEngine(path).render(0, want_linear=True).linear
        a = RayTracerEngine(inst)
        l2w = FlatContext(inst).instances()["l2w"].copy()
        l2w[:, 12] += 0.25                                   # every instance moves along x (column-major translation)
        a.set_instance_transforms(l2w, bounds_on_device=True)
        out["anim"] = a.render(0, want_linear=True).linear
        for name, p, n_tris, device in ply_cases.make_cases(str(tmp_path)):
            if device:
                monkeypatch.setenv("RTCUDA_INGEST", "device")
                v, nr = FlatContext(p).triangles()
                out["ply_" + name] = np.concatenate([v.ravel(), nr.ravel()])
        return out

    emu.rtemu_set_schedule_seed(0)
    base = run()
    emu.rtemu_set_schedule_seed(seed)
    shuffles = emu.rtemu_schedule_shuffles()
    again = run()
    assert emu.rtemu_schedule_shuffles() > shuffles + 100       # the other order was really imposed
    emu.rtemu_set_schedule_seed(0)
    assert base.keys() == again.keys() and len(base) > 6
    for key in base:
        assert np.array_equal(base[key], again[key], equal_nan=True), key


def _c_driver(tmp_path, lib_path):
    """Compiles tests/c_driver/driver.c with gcc (plain C, the two public headers only) against `lib_path`."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "driver")
    libdir, libname = os.path.split(lib_path)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(root, "include"),
                           os.path.join(root, "tests", "c_driver", "driver.c"), "-o", exe,
                           "-L", libdir, "-l:" + libname, "-Wl,-rpath," + libdir])
    return exe


def run_c_driver(exe, scene, image, devices=1, env=None):
    import subprocess
    r = subprocess.run([exe, scene, image] + ([str(devices)] if devices > 1 else []), stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    return r.stdout.split()


def fnv1a(data: bytes) -> int:
    h = 1469598103934665603
    for b in data:
        h = ((h ^ b) * 1099511628211) & ((1 << 64) - 1)
    return h


def test_c_abi_from_a_plain_c_program(scene_dir, tmp_path, emu):
    """The boundary without Python in between: a C99 program built from include/*.h alone creates
    the engine, renders, writes the image, then feeds the flattened scene to the render module directly; here linked
    against the emulation build, on the GPU box against librtcuda.so.  Its frame is the frame ctypes gets."""
    exe = _c_driver(tmp_path, emu_lib.EMU_PATH)
    path = scenegen.config_c3(scene_dir, width=40, height=24, samples=4, depth=3, ico_level=1)
    out = run_c_driver(exe, path, str(tmp_path / "c.ppm"))
    want = RayTracerEngine(path).render(0)
    assert out[0] == "frame" and (int(out[1]), int(out[2])) == (40, 24)
    assert out[3] == "fnv=%016x" % fnv1a(want.rgba.tobytes())
    assert out[4] == "rays=%d" % want.stats["rays"]
    assert out[6] == out[3].replace("fnv=", "fnv=") and out[5] == "module"
    ppm = open(tmp_path / "c.ppm", "rb").read()
    assert ppm.startswith(b"P6\n40 24\n255\n") and ppm[-40 * 24 * 3:] == want.rgba[..., :3].tobytes()
    assert out[7] == "progress" and out[9] == "last=24"


def test_c_abi_multi_device_scene(scene_dir, tmp_path, emu):
    """rt_scene_create_multi from plain C on three emulated devices (RTEMU_DEVICES): one build, replicas by device
    copies, one rt_render call shards the 8-row chunks over the devices' worker threads and every device copies its own
    rows into the caller's rt_host_alloc buffer — the frame and the ray total of the single-device scene.  41 rows:
    the ragged last chunk belongs to replica 2."""
    import os
    exe = _c_driver(tmp_path, emu_lib.EMU_PATH)
    path = scenegen.config_c3(scene_dir, width=40, height=41, samples=4, depth=3, ico_level=1)
    out = run_c_driver(exe, path, str(tmp_path / "m.ppm"), devices=3, env=dict(os.environ, RTEMU_DEVICES="3"))
    assert out[13] == "multi" and out[14] == "3", out
    assert out[15] == out[3] and out[16] == out[4], out


def test_multi_device_engine_matches_single(scene_dir, emu, monkeypatch):
    """RayTracerEngine(devices=[...]): linear colours, hit ids and RGBA8 of the sharded frame are the single-device
    frame's, bit for bit, for 2 and 3 (emulated) devices; progress is reported from the calling thread while the
    replicas' workers render; rt_scene_update_instances reaches every replica."""
    monkeypatch.setenv("RTEMU_DEVICES", "3")
    path = scenegen.config_instances(scene_dir, width=64, height=52, samples=4)
    one = RayTracerEngine(path)
    ref = one.render(0, want_linear=True, want_ids=True)
    for devs in ([0, 1], [0, 1, 2]):
        eng = RayTracerEngine(path, devices=devs)
        assert eng.ready, eng.load_error
        assert eng.scene_info()["deviceCount"] == len(devs)
        seen = []
        got = eng.render(0, want_linear=True, want_ids=True, progress=lambda done, total: seen.append((done, total)) or True)
        assert np.array_equal(got.rgba, ref.rgba) and np.array_equal(got.linear, ref.linear) and np.array_equal(got.ids, ref.ids)
        assert got.stats["rays"] == ref.stats["rays"] and got.stats["raysPrimary"] == ref.stats["raysPrimary"]
        assert seen[0] == (0, 52) and seen[-1] == (52, 52) and all(a[0] <= b[0] for a, b in zip(seen, seen[1:]))
        # animation: new transforms on every replica
        from myraytracer_b200.engine import FlatContext
        pf = scenegen.config_instances(scene_dir, frame=0.6, name="mframe", width=64, height=52, samples=4)
        want = FlatContext(pf).instances()
        eng.set_instance_transforms(want["l2w"])
        one.set_instance_transforms(want["l2w"])
        assert np.array_equal(eng.render(0).rgba, one.render(0).rgba)
        one.set_instance_transforms(FlatContext(path).instances()["l2w"])
        with pytest.raises(Exception):
            eng.render(0, shard_index=0, shard_count=2)          # a multi-device scene shards the frame itself
        eng.close()


def test_cancel_from_progress_callback(scene_dir, emu, monkeypatch):
    """A non-zero return of the progress callback stops the frame: RT_ERR_CANCELLED (-4), and the handle renders the
    next frame normally (two emulated devices: the callback runs while the workers' kernels are in flight)."""
    from myraytracer_b200 import RenderError
    monkeypatch.setenv("RTEMU_DEVICES", "2")
    path = scenegen.config_c1(scene_dir, width=96, height=96, samples=4)
    eng = RayTracerEngine(path, devices=[0, 1])
    ref = eng.render(0).rgba
    with pytest.raises(RenderError) as e:
        eng.render(0, progress=lambda done, total: False)        # cancelled before the launch
    assert e.value.code == -4
    calls = []
    def cb(done, total):
        calls.append(done)
        return not (0 < done < total)                            # cancel at the first in-flight report
    try:
        eng.render(0, progress=cb)
        cut = False
    except RenderError as err:
        assert err.code == -4
        cut = True
    assert cut == any(0 < c < 96 for c in calls)
    assert np.array_equal(eng.render(0).rgba, ref)


@pytest.mark.parametrize("rough,depth,variants,inf_light", [(0.0, 4, (1, 2), False), (0.15, 3, (2,), False),
                                                            (0.0, 2, (0, 2), True), (0.2, 3, (2,), True)])
def test_emulated_kernels_vs_independent_restatement(rough, depth, variants, inf_light, tmp_path, emu):
    """The CUDA sources against tests/reftrace.py directly — the pure-Python restatement of `trace` written from the
    Swift source, no BVH, no oracle in between: linear colours of every pixel to 1e-12 and the same ray counts."""
    import reftrace
    from test_oracle_trace_cpu import _build
    # inf_light: a light of infinite intensity -> non-finite Lo -> that level returns black (Object+Extension.swift:277-280)
    path, ref, cam = _build(str(tmp_path), rough, depth, W=24, H=16, inf_light=inf_light)
    want, counts = reftrace.render(ref, cam)
    eng = RayTracerEngine(path)
    for variant in variants:
        got = eng.render(0, want_linear=True, variant=variant)
        err = np.abs(got.linear - want) / np.maximum(1.0, np.abs(want))
        assert err.max() <= 1e-12, (variant, err.max())
        rays = (got.stats["raysPrimary"], got.stats["raysShadow"], got.stats["raysSecondary"])
        # inf_light too: the children of a level with a non-finite Lo are walked (and counted) like the reference
        # walks them, then discarded (PATH_DEAD, Object+Extension.swift:277-280)
        assert rays == (counts["primary"], counts["shadow"], counts["secondary"]), variant


def test_emulated_static_ingest_equals_rt_ply_ingest(emu):
    """rt_ply_ingest_static (no allocation: intermediates in a fixed array inside the library) on the emulation build,
    where device memory is host memory: bit-identical to rt_ply_ingest, call after call through the shared scratch,
    limits enforced.  The B200 test is tests/test_zz_no_alloc.py."""
    import ctypes as C
    import ply_cases
    from myraytracer_b200 import capi
    L = emu
    pos_a, tri_a = scenegen.terrain_mesh(40, 77)
    pos_b, tri_b = scenegen.icosphere_mesh(2, radius=1.5, seed=9, disp=0.2)
    quads = np.random.default_rng(5).integers(0, len(pos_b), size=(60, 4))
    for blocks, smooth in ((ply_cases.raw_blocks(pos_a, tri_a), 1), (ply_cases.raw_blocks(pos_b, tri_b), 0),
                           (ply_cases.raw_blocks(pos_b, quads, normals=pos_b * 1.7 + 0.1), 1), (ply_cases.raw_blocks(pos_a, tri_a), 1)):
        kw, vblock, fblock = blocks
        lay = ply_cases.make_layout(capi, kw)
        n = kw["faceCount"] * (2 if kw["verticesPerFace"] == 4 else 1)
        vb, fb = np.frombuffer(vblock, np.uint8), np.frombuffer(fblock, np.uint8)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        want_v, want_n, got_v, got_n = np.zeros((n, 9)), np.zeros((n, 9)), np.full((n, 9), np.nan), np.full((n, 9), np.nan)
        assert L.rt_ply_ingest(C.byref(lay), p(vb), p(fb), smooth, 0, p(want_v), p(want_n), None) == 0
        st = capi.rt_ply_stats()
        assert L.rt_ply_ingest_static(C.byref(lay), p(vb), p(fb), smooth, 0, p(got_v), p(got_n), None, C.byref(st)) == 0, L.rt_last_error()
        assert st.triangles == n
        assert np.array_equal(got_v, want_v) and np.array_equal(got_n, want_n, equal_nan=True)
    lay = ply_cases.make_layout(capi, dict(kw, vertexCount=capi.RT_PLY_STATIC_MAX_VERTICES + 1))
    assert L.rt_ply_ingest_static(C.byref(lay), p(vb), p(fb), 1, 0, p(got_v), p(got_n), None, None) == capi.RT_ERR_UNSUPPORTED
    assert b"65536" in L.rt_last_error()


# ---- beam walk of the wide BVH (rt_beam.cuh; built when RTCUDA_WIDE=1)

def _wide_leaves(wn, root):
    """Leaf references and boxes under wide node `root`, depth first; also the tree's depth."""
    leaves, depth, stack = [], 0, [(int(root), 1)]
    while stack:
        n, d = stack.pop()
        depth = max(depth, d)
        kids = wn[n]
        refs = kids[:, 6].view(np.int32)
        real = refs != 0x7fffffff
        assert real[:int(real.sum())].all(), "empty slots come last"
        assert int(real.sum()) >= 2
        axis = int(kids[0, 7])
        assert 0 <= axis <= 2
        box = kids[:, :6].view(np.float32)
        centre = box[real, axis].astype(np.float64) + box[real, 3 + axis]
        assert (np.diff(centre) >= 0).all(), "children sorted up the node's sort axis"
        assert (box[~real, :3] > box[~real, 3:]).all(), "empty slots carry an inverted box"
        for k in np.nonzero(real)[0]:
            if refs[k] >= 0:
                stack.append((int(refs[k]), d + 1))
            else:
                leaves.append((int(refs[k]), tuple(box[k])))
    return leaves, depth


def _binary_leaves(bn, root):
    leaves, stack = [], [int(root)]
    while stack:
        n = stack.pop()
        rec = bn[n]
        f = rec[:12].view(np.float32)
        boxes = [(f[0], f[1], f[2], f[3], f[4], f[5]), (f[6], f[7], f[8], f[9], f[10], f[11])]
        for ref, box in zip(rec[12:14].view(np.int32), boxes):
            if ref == 0x7fffffff:
                continue
            if ref >= 0:
                stack.append(int(ref))
            else:
                leaves.append((int(ref), tuple(box)))
    return leaves


@pytest.mark.parametrize("name", ["c2", "c5", "c3"])
def test_emulated_wide_tree_is_the_binary_tree_collapsed(name, scene_dir, emu, monkeypatch):
    """bvh_build.cpp: collapseWide — every leaf of the binary tree hangs under the wide tree exactly once with the same
    reference and the same conservative box; children sorted along the node's sort axis, empty slots last."""
    monkeypatch.setenv("RTCUDA_WIDE", "1")
    fn, kw, _ = CASES[name]
    eng = RayTracerEngine(fn(scene_dir, **kw))
    info = eng.scene_info()
    assert info["wideNodes"] > 0
    wn, bn = eng.wide_nodes(), eng.blas_nodes()
    assert wn.shape[0] == info["wideNodes"]
    # roots: a wide node that no other wide node refers to, one per mesh BLAS with >= 64 triangles
    refs = wn[:, :, 6].view(np.int32)
    inner = refs[(refs >= 0) & (refs != 0x7fffffff)]
    roots = sorted(set(range(wn.shape[0])) - set(inner.tolist()))
    brefs = bn[:, 12:14].view(np.int32)
    broots = sorted(set(range(bn.shape[0])) - set(brefs[(brefs >= 0) & (brefs != 0x7fffffff)].tolist()))
    wide_sets = [sorted(_wide_leaves(wn, r)[0]) for r in roots]
    bin_sets = [sorted(_binary_leaves(bn, r)) for r in broots]
    for ws in wide_sets:
        assert ws in bin_sets            # same leaves, same boxes, as one of the binary trees
    assert len(wide_sets) >= 1


@pytest.mark.parametrize("name", ["c2", "c5", "c3", "c3x16", "instances", "c1x4_ragged"])
def test_emulated_beam_walk_vs_oracle(name, scene_dir, emu, monkeypatch):
    """The fast kernel with the beam walk switched on (RTCUDA_WIDE=1): same parity bars against the oracle, and the
    frame of the packet walk bit for bit up to tie-breaking (the beam only culls; t, u, v come from the same FP64 test)."""
    fn, kw, _ = CASES[name]
    path = fn(scene_dir, **kw)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    base = RayTracerEngine(path).render(0, want_linear=True, variant=1, counters=True)
    monkeypatch.setenv("RTCUDA_WIDE", "1")
    eng = RayTracerEngine(path)
    got = eng.render(0, want_ids=True, want_linear=True, variant=1, counters=True)
    check(got, ref)
    assert (base.rgba == got.rgba).all(axis=-1).mean() >= 0.9999
    if eng.scene_info()["wideNodes"] > 0:
        assert got.stats["boxTests"] != base.stats["boxTests"]          # the beam walk ran
    assert got.stats["raysPrimary"] == base.stats["raysPrimary"] and got.stats["raysSecondary"] == base.stats["raysSecondary"]


@pytest.mark.parametrize("mode", ["2", "3"])
def test_emulated_beam_walk_by_packet_kind(mode, scene_dir, emu, monkeypatch):
    """RTCUDA_WIDE=2: only closest-hit packets walk as beams, =3: only any-hit (shadow) packets; the other kind takes the
    packet walk in the same frame.  Same image, and the box-test counts differ from both pure modes."""
    path = scenegen.config_c2(scene_dir, n=32, width=64, height=40)
    runs = {}
    for m in ("0", "1", mode):
        monkeypatch.setenv("RTCUDA_WIDE", m)
        runs[m] = RayTracerEngine(path).render(0, want_ids=True, variant=1, counters=True)
    assert np.array_equal(runs[mode].rgba, runs["0"].rgba) and np.array_equal(runs["1"].rgba, runs["0"].rgba)
    assert np.array_equal(runs[mode].ids["prim"], runs["0"].ids["prim"])
    assert len({runs[m].stats["boxTests"] for m in runs}) == 3


def test_emulated_beam_walk_budget_falls_back_to_the_packet_walk(scene_dir, emu, monkeypatch):
    """A low-resolution frame over a fine mesh makes fat beams (an 8x4-pixel packet spans hundreds of triangles): the
    visit budget runs out and the packet walk finishes — same image."""
    path = scenegen.config_c2(scene_dir, n=96, width=48, height=32)
    base = RayTracerEngine(path).render(0, want_ids=True, variant=1)
    monkeypatch.setenv("RTCUDA_WIDE", "1")
    got = RayTracerEngine(path).render(0, want_ids=True, variant=1)
    assert (base.rgba == got.rgba).all()
    assert (base.ids["prim"] == got.ids["prim"]).mean() >= 0.9999
