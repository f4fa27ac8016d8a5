"""GPU parity tests proper: CUDA path through the C ABI vs the CPU oracle on the same seeded scenes."""
import numpy as np
import pytest

import oracle_lib
import parity
from myraytracer_b200 import RayTracerEngine, scenegen

pytestmark = pytest.mark.gpu

CASES = {
    "c1": dict(fn=scenegen.config_c1, kw=dict()),                                   # full config: 800x800, 1 spp
    "c2": dict(fn=scenegen.config_c2, kw=dict(width=512, height=512)),              # 69,938 triangles
    "c2full": dict(fn=scenegen.config_c2, kw=dict()),                               # full config: 1024x1024, 1 spp
    "c3": dict(fn=scenegen.config_c3, kw=dict(width=256, height=256, samples=4)),   # depth 6 recursion
    "c3full": dict(fn=scenegen.config_c3, kw=dict()),                               # full config: 1024x1024, 4 spp, depth 6
    "c4s": dict(fn=scenegen.config_c4, kw=dict(n=260, width=480, height=270, samples=4)),
    "c5s": dict(fn=scenegen.config_c5, kw=dict(n=64, width=320, height=180, samples=4, meshes=4)),
    # SURVEY.md 8f row 3, beyond BASELINE.json's configs: thin-lens camera (fast kernel, 5 draws per sample),
    # glossy materials (general kernel: draws inside the recursion), both
    # the object kinds BASELINE.json's configs leave out: infinite plane, MeshInstances sharing a BLAS (material
    # override, mirrored transform, instance of an instance), depth 3
    "instances": dict(fn=scenegen.config_instances, kw=dict()),
    # 64 and 36 spp: units of one pixel (32 divides n*n) and of a whole tile (neither divides the other)
    "c1x64": dict(fn=scenegen.config_c1, kw=dict(width=96, height=80, samples=64)),
    "c1x36": dict(fn=scenegen.config_c1, kw=dict(width=96, height=80, samples=36)),
    # the warp's ray queue under every unit rule of the fast kernel (1, 9, 16, 64 spp)
    "c3x1": dict(fn=scenegen.config_c3, kw=dict(width=320, height=240, samples=1)),
    "c3x9": dict(fn=scenegen.config_c3, kw=dict(width=160, height=120, samples=9)),
    "c3x16": dict(fn=scenegen.config_c3, kw=dict(width=160, height=120, samples=16)),
    "c3x64": dict(fn=scenegen.config_c3, kw=dict(width=80, height=60, samples=64, depth=4)),
    "c3dof": dict(fn=scenegen.config_c3, kw=dict(width=192, height=192, samples=4, aperture=0.35, focus=13.0)),
    "c3glossy": dict(fn=scenegen.config_c3, kw=dict(width=192, height=192, samples=4, roughness=0.12)),
    "c3glossydof": dict(fn=scenegen.config_c3, kw=dict(width=160, height=160, samples=9, depth=4, roughness=0.3,
                                                       aperture=0.2, focus=14.0)),
    # area lights (8f row 3): light samples come from a 100-entry table through one running index per 8-row chunk —
    # counting pass, per-chunk prefix sum, shading pass (general kernel)
    "c3area": dict(fn=scenegen.config_c3, kw=dict(width=192, height=180, samples=4, area_lights=2)),
    # motion blur (8f row 3): Mesh.motionBlur and MeshInstance.motionBlur
    "motionmesh": dict(fn=scenegen.config_instances, kw=dict(mesh_motion=(0.5, 0.2, -0.3))),
    "motioninst": dict(fn=scenegen.config_instances, kw=dict(width=200, height=150, samples=9, instance_motion=1.2)),
    "motionboth": dict(fn=scenegen.config_instances, kw=dict(width=243, height=181, samples=4, mesh_motion=(0.4, -0.3, 0.2),
                                                            instance_motion=1.5)),
    # non-finite guard (a11, :277-280) with every side effect of the discarded children: area-light jitter index, glossy RNG
    "c3areaglossyinf": dict(fn=scenegen.config_c3, kw=dict(width=120, height=96, samples=4, depth=3, roughness=0.15, area_lights=2,
                                                          inf_light=True)),
    "c3inf": dict(fn=scenegen.config_c3, kw=dict(width=120, height=96, samples=4, depth=3, inf_light=True)),
    "c3areaglossy": dict(fn=scenegen.config_c3, kw=dict(width=150, height=117, samples=9, depth=4, roughness=0.2, area_lights=3)),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_parity_vs_oracle(name, scene_dir):
    case = CASES[name]
    path = case["fn"](scene_dir, **case["kw"])
    eng = RayTracerEngine(path)
    assert eng.ready, eng.load_error
    got = eng.render(0, want_ids=True, want_linear=True)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    rep = parity.compare(got, ref)
    print(name, rep)
    assert rep["alpha_ok"]
    assert rep["id_agree"] >= 0.9999, rep
    assert rep["id_bad"] == 0, rep                      # disagreements only at t ties (1e-5 relative)
    assert rep["px_within_1"] >= 0.999, rep
    assert rep["rays_gpu"][0] == rep["rays_ref"][0]     # primary rays: exact
    # shadow / secondary counts may differ only through tie-resolved paths
    for a, b in zip(rep["rays_gpu"][1:], rep["rays_ref"][1:]):
        assert abs(a - b) <= max(4, 1e-4 * b), rep


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("name", ["c1", "c3", "c5s", "instances"])
def test_kernel_variants_vs_oracle(name, variant, scene_dir):
    """Every kernel variant (fast: sample lanes + warp packets; general: pixel lanes, per-lane traversal) meets the parity bars, and the
    variants agree with each other except at hit-distance ties."""
    case = CASES[name]
    kw = dict(case["kw"])
    if name == "c1":
        kw.update(width=200, height=200, samples=9)      # 9 spp: the non-power-of-two ordered-sum path
    path = case["fn"](scene_dir, **kw)
    eng = RayTracerEngine(path)
    got = eng.render(0, want_ids=True, want_linear=True, variant=variant)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    rep = parity.compare(got, ref)
    print(name, variant, rep)
    assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0, rep
    assert rep["px_within_1"] >= 0.999, rep
    assert rep["rays_gpu"][0] == rep["rays_ref"][0]
    base = eng.render(0, want_linear=True, variant=2)
    same = (base.rgba == got.rgba).all(axis=-1).mean()
    assert same >= 0.9999, same


@pytest.mark.parametrize("name", ["c1_small", "c2_small", "c3_small", "c5_small", "c3_area_small", "motion_small"])
def test_gpu_matches_committed_golden(name, scene_dir):
    """CUDA path vs the frozen oracle renders under tests/golden/ (tools/make_golden.py)."""
    import os
    from golden_cases import GOLDEN_CASES
    fn, kw = GOLDEN_CASES[name]
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"{name}.npz"))
    got = RayTracerEngine(fn(scene_dir, **kw)).render(0, want_ids=True)
    d = np.abs(got.rgba[..., :3].astype(int) - g["rgba"][..., :3].astype(int)).max(axis=-1)
    assert (d <= 1).mean() >= 0.999, (d <= 1).mean()
    same = (got.ids["prim"] == g["prim"]) & (got.ids["kind"] == g["kind"])
    assert same.mean() >= 0.9999
    tg, tr = got.ids["t"][~same], g["t"][~same]
    assert np.all(np.abs(tg - tr) <= 1e-5 * np.abs(tr))          # disagreements only at hit-distance ties
    assert got.stats["raysPrimary"] == int(g["rays"][0])


@pytest.mark.parametrize("shard", ["rows", "tiles"])
def test_shards_are_bit_identical(shard, scene_dir):
    """Decomposition invariance: per-pixel RNG depends only on (i, j) (Object+Extension.swift:294).  Both
    rt_render_params.shardMode rules partition the frame; a shard leaves the other pixels untouched (zero here)."""
    from myraytracer_b200.distributed import owned_mask
    path = scenegen.config_c3(scene_dir, width=130, height=100, samples=4, depth=3, ico_level=2)
    eng = RayTracerEngine(path)
    full = eng.render(0).rgba
    acc = np.zeros_like(full)
    for n in (2, 3, 8):
        acc[:] = 0
        for k in range(n):
            part = eng.render(0, shard_index=k, shard_count=n, shard_mode=1 if shard == "tiles" else 0).rgba
            own = owned_mask(100, 130, k, n, shard)
            assert (part[~own] == 0).all()
            acc[own] = part[own]
        assert np.array_equal(acc, full)


@pytest.mark.parametrize("variant", [1, 2])
def test_tile_schedule_does_not_change_the_frame(variant, scene_dir):
    """rt_render_params.scheduleMode: the second frame draws its tiles in the cost order learnt from the first;
    pixels, linear colours and hit ids are those of the spatial order, whole frame and per shard."""
    path = scenegen.config_c3(scene_dir, width=200, height=120, samples=4, depth=3, ico_level=2)
    eng = RayTracerEngine(path)
    ref = eng.render(0, want_linear=True, want_ids=True, variant=variant, schedule=1)
    assert ref.stats["tilesScheduled"] == 0
    first = eng.render(0, want_linear=True, want_ids=True, variant=variant)
    assert first.stats["tilesScheduled"] == 0               # nothing learnt yet: spatial order
    assert sorted(eng.schedule().tolist()) == list(range(25 * 30))      # a permutation of the tiles
    for _ in range(3):
        again = eng.render(0, want_linear=True, want_ids=True, variant=variant)
        assert again.stats["tilesScheduled"] == 25 * 30      # 8x4 tiles of the 200x120 frame
        assert again.stats["kernelLaunches"] == 3
        for a in (first, again):
            assert np.array_equal(a.rgba, ref.rgba)
            assert np.array_equal(a.linear, ref.linear)
            assert np.array_equal(a.ids, ref.ids)
        assert again.stats["rays"] == ref.stats["rays"]
        assert sorted(eng.schedule().tolist()) == list(range(25 * 30))
    # another shard layout invalidates the order, then learns its own
    part = [eng.render(0, shard_index=1, shard_count=2, variant=variant) for _ in range(2)]
    assert part[0].stats["tilesScheduled"] == 0 and part[1].stats["tilesScheduled"] == 25 * 2 * 7
    assert np.array_equal(part[0].rgba, part[1].rgba)
    assert sorted(eng.schedule().tolist()) == list(range(25 * 2 * 7))
    rows = np.zeros(120, bool)
    for c in range(1, 15, 2):
        rows[c * 8:(c + 1) * 8] = True
    assert np.array_equal(part[1].rgba[rows], ref.rgba[rows]) and (part[1].rgba[~rows] == 0).all()


def test_first_frame_probe_orders_the_first_frame(scene_dir, monkeypatch):
    """RTCUDA_PROBE=1 (opt-in): the first frame of a geometry takes its tile order from one probe ray per tile
    (tileProbeKernel + the two schedule kernels); same image."""
    path = scenegen.config_c3(scene_dir, width=200, height=120, samples=4, depth=3, ico_level=2)
    ref = RayTracerEngine(path).render(0, schedule=1)
    monkeypatch.setenv("RTCUDA_PROBE", "1")
    eng = RayTracerEngine(path)
    first = eng.render(0)
    assert first.stats["tilesScheduled"] == 25 * 30 and first.stats["kernelLaunches"] == 6
    assert np.array_equal(first.rgba, ref.rgba)
    again = eng.render(0)
    assert again.stats["tilesScheduled"] == 25 * 30 and again.stats["kernelLaunches"] == 3
    assert np.array_equal(again.rgba, ref.rgba)


def test_tile_schedule_many_tiles(scene_dir):
    """More tiles than one block of the schedule kernels sorts (1024): still a permutation, same frame."""
    path = scenegen.config_c1(scene_dir, width=648, height=404)
    eng = RayTracerEngine(path)
    ref = eng.render(0, schedule=1).rgba
    n_tiles = 81 * 51 * 2                                      # 51 chunks of 8 rows, two tile rows each
    for k in range(3):
        out = eng.render(0)
        assert out.stats["tilesScheduled"] == (n_tiles if k else 0)
        assert np.array_equal(out.rgba, ref)
        order = eng.schedule()
        assert len(order) == n_tiles and np.array_equal(np.sort(order), np.arange(n_tiles))
    # tiles past the last image row (404 = 50.5 chunks) trace nothing: they are drawn towards the end
    slot = np.empty(n_tiles, np.int64)
    slot[order] = np.arange(n_tiles)
    assert np.median(slot[n_tiles - 81:]) > 0.75 * n_tiles


@pytest.mark.parametrize("name", ["c2", "c3", "c4s", "c5s"])
def test_device_built_bvh_meets_the_parity_bars(name, scene_dir, monkeypatch):
    """rt_scene_desc.buildMode = RT_BUILD_DEVICE_LBVH (here through RTCUDA_BUILD): the mesh BVHs come from the CUDA
    builder (Morton order + radix tree); the image must not depend on the tree."""
    case = CASES[name]
    path = case["fn"](scene_dir, **case["kw"])
    monkeypatch.setenv("RTCUDA_BUILD", "host")
    host = RayTracerEngine(path)
    assert host.scene_info()["deviceBuiltBlas"] == 0
    base = host.render(0, want_ids=True)
    monkeypatch.setenv("RTCUDA_BUILD", "device")
    eng = RayTracerEngine(path)
    info = eng.scene_info()
    assert info["deviceBuiltBlas"] >= 1 and info["maxLeafSize"] <= 2, info
    got = eng.render(0, want_ids=True, want_linear=True)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    rep = parity.compare(got, ref)
    print(name, info, rep)
    assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0, rep
    assert rep["px_within_1"] >= 0.999, rep
    assert rep["rays_gpu"][0] == rep["rays_ref"][0]
    # against the host-built tree of the same module: same hits except at equal-t ties
    assert (got.ids["prim"] == base.ids["prim"]).mean() >= 0.9999
    assert (got.rgba == base.rgba).all(axis=-1).mean() >= 0.9999


@pytest.mark.parametrize("name", ["c2", "c3", "c4s", "c5s"])
def test_device_sah_build_makes_the_host_builders_tree(name, scene_dir, monkeypatch):
    """rt_scene_desc.buildMode = RT_BUILD_DEVICE_SAH (csrc/bvh_build_sah_gpu.cu): same nodes, boxes and leaves as the
    host builder — identical box-test count and image — and the parity bars against the oracle."""
    case = CASES[name]
    path = case["fn"](scene_dir, **case["kw"])
    monkeypatch.setenv("RTCUDA_BUILD", "host")
    host = RayTracerEngine(path)
    a = host.render(0, want_ids=True, want_linear=True, counters=True)
    monkeypatch.setenv("RTCUDA_BUILD", "device-sah")
    dev = RayTracerEngine(path)
    hi, di = host.scene_info(), dev.scene_info()
    print(name, "host build ms", hi["buildMs"], "device-sah build ms", di["buildMs"])
    assert hi["deviceBuiltBlas"] == 0 and di["deviceBuiltBlas"] >= 1
    for key in ("blasNodes", "maxLeafSize", "maxDepth"):
        assert hi[key] == di[key], (key, hi, di)
    assert np.array_equal(host.blas_nodes(), dev.blas_nodes())       # the host builder's node array, byte for byte
    b = dev.render(0, want_ids=True, want_linear=True, counters=True)
    assert a.stats["boxTests"] == b.stats["boxTests"]
    assert np.array_equal(a.rgba, b.rgba)
    rep = parity.compare(b, oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True))
    assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0, rep
    assert rep["px_within_1"] >= 0.999, rep


@pytest.mark.parametrize("name", ["c2full", "c4s", "c5s", "c3", "instances"])
def test_beam_walk_meets_the_parity_bars(name, scene_dir, monkeypatch):
    """RTCUDA_WIDE=1: coherent packets walk the wide BVH as a beam (rt_beam.cuh: one lane per child box, per-ray tests at
    the leaves only).  Same parity bars against the oracle, and the packet walk's frame up to tie-breaking."""
    case = CASES[name]
    path = case["fn"](scene_dir, **case["kw"])
    base = RayTracerEngine(path).render(0, variant=1)
    monkeypatch.setenv("RTCUDA_WIDE", "1")
    eng = RayTracerEngine(path)
    assert eng.ready, eng.load_error
    assert eng.scene_info()["wideNodes"] > 0
    got = eng.render(0, want_ids=True, want_linear=True, variant=1)
    ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    rep = parity.compare(got, ref)
    print(name, "beam", rep)
    assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0, rep
    assert rep["px_within_1"] >= 0.999, rep
    assert rep["t_bit_exact"] == 1.0, rep
    assert (base.rgba == got.rgba).all(axis=-1).mean() >= 0.9999
    assert got.stats["raysPrimary"] == base.stats["raysPrimary"]


def test_default_build_mode_sends_big_meshes_to_the_device_builder(scene_dir, monkeypatch):
    """RT_BUILD_SAH (the default): a mesh of RT_BUILD_AUTO_DEVICE_TRIANGLES triangles or more is built by the device SAH
    builder — the host builder's node array byte for byte, so the frame is the host-built one — smaller meshes on the host."""
    monkeypatch.delenv("RTCUDA_BUILD", raising=False)
    big = RayTracerEngine(scenegen.config_c2(scene_dir))                    # 69,938 triangles
    small = RayTracerEngine(scenegen.config_c2(scene_dir, n=100))           # 19,602 triangles
    assert big.scene_info()["deviceBuiltBlas"] == 1 and small.scene_info()["deviceBuiltBlas"] == 0
    frame, nodes = big.render(0).rgba, big.blas_nodes()
    monkeypatch.setenv("RTCUDA_BUILD", "host")
    host = RayTracerEngine(scenegen.config_c2(scene_dir))
    assert host.scene_info()["deviceBuiltBlas"] == 0
    assert (host.blas_nodes() == nodes).all()
    assert (host.render(0).rgba == frame).all(axis=-1).mean() >= 0.9999      # the triangle order inside a leaf may differ: ties only


def test_c_abi_from_a_plain_c_program_on_the_gpu(scene_dir, tmp_path):
    """tests/c_driver/driver.c linked against librtcuda.so: engine -> frame -> image file, then the flattened scene
    straight into rt_scene_create / rt_render; same frame as through ctypes."""
    from myraytracer_b200 import capi
    from test_emu_cpu import _c_driver, fnv1a, run_c_driver
    exe = _c_driver(tmp_path, capi.LIB_PATH)
    path = scenegen.config_c3(scene_dir, width=160, height=96, samples=4, depth=4, ico_level=2)
    out = run_c_driver(exe, path, str(tmp_path / "c.png"))
    want = RayTracerEngine(path).render(0)
    assert out[3] == "fnv=%016x" % fnv1a(want.rgba.tobytes())
    assert out[6] == out[3]
    assert out[7] == "progress" and out[9] == "last=96"


def test_cancel_mid_flight_from_c(scene_dir, tmp_path):
    """RayTracer.swift:172-181 through the C ABI: the progress callback of a full-size C4 frame (1920x1080x16, ~10 ms)
    is called while the kernel runs and cancels at its first in-flight report; rt_render returns RT_ERR_CANCELLED
    within 2 ms of the request (the work counter is pushed past its end; warps finish the unit they hold), and the
    next frame on the same handle is complete and identical."""
    from myraytracer_b200 import capi
    from test_emu_cpu import _c_driver, run_c_driver
    exe = _c_driver(tmp_path, capi.LIB_PATH)
    out = run_c_driver(exe, scenegen.config_c4(scene_dir), str(tmp_path / "c4.ppm"))
    calls = int(out[8].split("=")[1])
    print(out[7:13])
    assert calls >= 5, out                                   # reports while in flight, not just at row 0 and row H
    assert out[10] == "cancel" and out[11] == "rc=-4", out
    assert 0 <= float(out[12].split("=")[1]) < 2.0, out


def test_multi_device_scene_from_c(scene_dir, tmp_path):
    """>= 2 GPUs only: rt_scene_create_multi + one rt_render call from plain C (SURVEY.md 8b/8e through the boundary the
    Swift host binds): same frame, same ray total as the single-device scene."""
    import torch
    from myraytracer_b200 import capi
    from test_emu_cpu import _c_driver, run_c_driver
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    exe = _c_driver(tmp_path, capi.LIB_PATH)
    path = scenegen.config_c3(scene_dir, width=320, height=203, samples=4, depth=4, ico_level=2)
    out = run_c_driver(exe, path, str(tmp_path / "m.png"), devices=n)
    assert out[13] == "multi" and out[14] == str(n) and out[15] == out[3] and out[16] == out[4], out


def test_multi_device_engine_on_the_gpus(scene_dir):
    """>= 2 GPUs only: RayTracerEngine(devices=[...]) — the frame of one rt_render call over all GPUs equals the
    single-GPU frame bit for bit (rgba, linear, ids), at C4's reduced size and on the recursive C3."""
    import torch
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    for name in ("c4s", "c3", "c3area"):
        case = CASES[name]
        path = case["fn"](scene_dir, **case["kw"])
        ref = RayTracerEngine(path).render(0, want_linear=True, want_ids=True)
        eng = RayTracerEngine(path, devices=list(range(n)))
        assert eng.ready, eng.load_error
        for _ in range(2):
            got = eng.render(0, want_linear=True, want_ids=True)
            assert np.array_equal(got.rgba, ref.rgba) and np.array_equal(got.linear, ref.linear) and np.array_equal(got.ids, ref.ids), name
            assert got.stats["rays"] == ref.stats["rays"], name
        eng.close()


def test_animation_refit_matches_rebuilt_scenes(scene_dir):
    """rt_scene_update_instances (8f row 4): instance records re-derived and the TLAS refitted by the device kernel,
    topology kept; every frame equals the frame of a scene built from scratch with those transforms and meets the
    parity bars against the oracle."""
    from myraytracer_b200.engine import FlatContext
    kw = dict(width=200, height=150, samples=4)
    eng = RayTracerEngine(scenegen.config_instances(scene_dir, **kw))
    base = eng.render(0, want_linear=True)
    for f in (0.3, 1.0, 0.0):
        pf = scenegen.config_instances(scene_dir, frame=f, name=f"frame{f}", **kw)
        want = FlatContext(pf).instances()
        eng.set_instance_transforms(want["l2w"])                # mesh bounds derived by the device kernel
        d = eng.describe()
        bounds = np.array([list(d.instances[i].worldBounds) for i in range(d.nInstances)])
        assert np.array_equal(bounds, want["bounds"])           # bit-identical to the host loop of refitTLAS step 1
        got = eng.render(0, want_ids=True, want_linear=True)
        rep = parity.compare(got, oracle_lib.OracleScene(pf).render(want_ids=True, want_linear=True))
        assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0 and rep["px_within_1"] >= 0.999, rep
        fresh = RayTracerEngine(pf).render(0, want_linear=True)
        assert (fresh.rgba == got.rgba).all(axis=-1).mean() >= 0.9999        # other TLAS topology: ties only
        assert (f == 0.0) == np.array_equal(got.linear, base.linear)


def test_ply_device_ingest_equals_host_flatten(tmp_path, scene_dir, monkeypatch):
    """rt_ply_ingest on the B200 (8f row 2): bit-identical to the host loader + flatten loops on every layout of
    tests/ply_cases.py and on the 999,698-triangle C4 mesh; engines ingest on the device by default."""
    import time
    import ply_cases
    from myraytracer_b200.engine import FlatContext
    cases = [(n, p, t) for n, p, t, _ in ply_cases.make_cases(str(tmp_path))]
    cases.append(("c4", scenegen.config_c4(scene_dir, width=64, height=64, samples=1), 999698))
    for name, path, n_tris in cases:
        monkeypatch.setenv("RTCUDA_INGEST", "host")
        t0 = time.perf_counter(); hv, hn = FlatContext(path).triangles(); t_host = time.perf_counter() - t0
        monkeypatch.setenv("RTCUDA_INGEST", "device")
        t0 = time.perf_counter(); dv, dn = FlatContext(path).triangles(); t_dev = time.perf_counter() - t0
        print(f"ingest {name}: {n_tris} triangles, host {t_host * 1e3:.1f} ms, device path {t_dev * 1e3:.1f} ms")
        assert len(hv) == n_tris and len(dv) == n_tris, name
        assert np.array_equal(hv, dv) and np.array_equal(hn, dn, equal_nan=True), name


def test_edge_cases_vs_oracle(scene_dir):
    """tests/edge_cases.py on the B200: empty scene, one-instance TLAS, tiny and ragged images, non-square sample
    counts, depth 0 with a mirror — both kernels."""
    import edge_cases
    for name in edge_cases.EDGE_CASES:
        path = edge_cases.make(scene_dir, name)
        eng = RayTracerEngine(path)
        assert eng.ready, (name, eng.load_error)
        ref = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
        for variant in (1, 2):
            got = eng.render(0, want_ids=True, want_linear=True, variant=variant)
            rep = parity.compare(got, ref)
            assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0 and rep["px_within_1"] >= 0.999, (name, variant, rep)
            assert rep["rays_gpu"] == rep["rays_ref"], (name, variant, rep)


@pytest.mark.parametrize("rough,depth,variants", [(0.0, 3, (0, 2)), (0.2, 3, (2,))])
def test_non_finite_light_level_is_discarded_like_the_reference(rough, depth, variants, tmp_path):
    """Row a11 on the hardware (Object+Extension.swift:277-280): a point light of infinite intensity makes Lo
    non-finite at every shading point it reaches; the reference still traces that level's children (their rays, and
    with glossy materials their RNG draws) and then returns black for the level.  Against the pure-Python restatement
    of `trace` (tests/reftrace.py) and against the oracle: same colours, same ray counts."""
    import reftrace
    from test_oracle_trace_cpu import _build
    path, ref, cam = _build(str(tmp_path), rough, depth, W=48, H=32, inf_light=True)
    want, counts = reftrace.render(ref, cam)
    orc = oracle_lib.OracleScene(path).render(want_ids=True, want_linear=True)
    eng = RayTracerEngine(path)
    from myraytracer_b200 import RenderError
    with pytest.raises(RenderError) as e:          # the fast kernel carries no guard: such scenes take the general kernel
        eng.render(0, variant=1)
    assert e.value.code == -3
    for variant in variants:
        got = eng.render(0, want_linear=True, want_ids=True, variant=variant)
        assert np.isfinite(got.linear).all()
        err = np.abs(got.linear - want) / np.maximum(1.0, np.abs(want))
        assert err.max() <= 1e-12, (variant, err.max())
        assert (got.stats["raysPrimary"], got.stats["raysShadow"], got.stats["raysSecondary"]) == \
            (counts["primary"], counts["shadow"], counts["secondary"]), variant
        rep = parity.compare(got, orc)
        assert rep["id_agree"] >= 0.9999 and rep["id_bad"] == 0 and rep["px_within_1"] >= 0.999, (variant, rep)
        assert rep["rays_gpu"] == rep["rays_ref"], (variant, rep)


def test_error_codes(scene_dir):
    path = scenegen.config_c1(scene_dir, width=32, height=32)
    eng = RayTracerEngine(path)
    from myraytracer_b200 import RenderError
    with pytest.raises(RenderError) as e:
        eng.render(5)
    assert e.value.code == -10                            # "Invalid camera index" (RayTracer.swift:152-155)
    empty = RayTracerEngine(path + ".missing")
    assert not empty.ready
    with pytest.raises(RenderError) as e:
        empty.render(0)
    assert e.value.code == -20                            # "No scene loaded" (:142-145)


def test_multi_gpu_frame_equals_single_gpu_frame():
    """>= 2 GPUs only: torchrun, one rank per GPU, P2P (kernel stores into rank 0's frame) and NCCL gather."""
    import os, subprocess, sys
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
                          "127.0.0.1", "--master-port", "29533", os.path.join(here, "mgpu_check.py")], capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0 and "MGPU_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
