"""rt_ply_ingest_static — the entry point that allocates nothing (its scratch is a fixed device array inside the
library; no counterpart in the reference) — against rt_ply_ingest, the
existing entry point of the same kernels, through the C ABI.  Collected last on purpose:
a failure here must not hide the parity tests."""
import ctypes as C

import numpy as np
import pytest

import ply_cases
from myraytracer_b200 import capi, scenegen

pytestmark = pytest.mark.gpu


def _ingest_existing(L, layout_kw, vblock, fblock, smooth):
    lay = ply_cases.make_layout(capi, layout_kw)
    n_tris = layout_kw["faceCount"] * (2 if layout_kw["verticesPerFace"] == 4 else 1)
    tv, tn = np.zeros((n_tris, 9)), np.zeros((n_tris, 9))
    vb, fb = np.frombuffer(vblock, np.uint8), np.frombuffer(fblock, np.uint8)
    rc = L.rt_ply_ingest(C.byref(lay), vb.ctypes.data_as(C.c_void_p), fb.ctypes.data_as(C.c_void_p), smooth, 0,
                         tv.ctypes.data_as(C.c_void_p), tn.ctypes.data_as(C.c_void_p), None)
    assert rc == 0, L.rt_last_error().decode()
    return tv, tn


def _ingest_static(L, layout_kw, d_vblock, d_fblock, smooth, d_tv, d_tn, stream=0):
    """Inputs and outputs are the caller's device tensors, allocated BEFORE the call; the call itself allocates nothing."""
    lay = ply_cases.make_layout(capi, layout_kw)
    st = capi.rt_ply_stats()
    rc = L.rt_ply_ingest_static(C.byref(lay), d_vblock.data_ptr(), d_fblock.data_ptr(), smooth, 0, d_tv.data_ptr(), d_tn.data_ptr(),
                                stream or None, C.byref(st))
    return rc, st


def test_static_ingest_equals_rt_ply_ingest_on_two_calls_in_a_row():
    import torch
    L = capi.load()
    assert L.rt_device_count() >= 1, "needs a CUDA device: there is no CPU fallback"
    dev = torch.device("cuda", 0)
    # three different meshes, back to back through the ONE shared scratch array: a big smooth one (the adjacency lists and
    # the scan are exercised), a small flat one whose intermediates sit on top of the first call's leftovers, quads with
    # normals in the file, and the first one again
    pos_a, tri_a = scenegen.terrain_mesh(150, 77)                              # 22,500 vertices, 44,402 triangles
    pos_b, tri_b = scenegen.icosphere_mesh(2, radius=1.5, seed=9, disp=0.2)    # 162 vertices, 320 triangles
    rng = np.random.default_rng(5)
    quads = rng.integers(0, len(pos_b), size=(60, 4))
    calls = [("terrain smooth", ply_cases.raw_blocks(pos_a, tri_a), 1),
             ("icosphere flat", ply_cases.raw_blocks(pos_b, tri_b), 0),
             ("quads, file normals", ply_cases.raw_blocks(pos_b, quads, normals=pos_b * 1.7 + 0.1), 1),
             ("terrain smooth again", ply_cases.raw_blocks(pos_a, tri_a), 1)]
    max_tris = max(kw["faceCount"] * (2 if kw["verticesPerFace"] == 4 else 1) for _, (kw, _, _), _ in calls)
    d_tv = torch.empty((max_tris, 9), dtype=torch.float64, device=dev)
    d_tn = torch.empty((max_tris, 9), dtype=torch.float64, device=dev)
    for name, (kw, vblock, fblock), smooth in calls:
        want_v, want_n = _ingest_existing(L, kw, vblock, fblock, smooth)
        d_vb = torch.frombuffer(bytearray(vblock), dtype=torch.uint8).to(dev)
        d_fb = torch.frombuffer(bytearray(fblock), dtype=torch.uint8).to(dev)
        d_tv.fill_(float("nan")); d_tn.fill_(float("nan"))
        torch.cuda.synchronize(dev)
        before = torch.cuda.memory_allocated(dev)
        rc, st = _ingest_static(L, kw, d_vb, d_fb, smooth, d_tv, d_tn)
        assert rc == 0, (name, L.rt_last_error().decode())
        assert torch.cuda.memory_allocated(dev) == before
        n = len(want_v)
        assert st.triangles == n and st.kernelLaunches >= 3
        got_v, got_n = d_tv[:n].cpu().numpy(), d_tn[:n].cpu().numpy()
        # the bar of the ingest tests (test_ply_device_ingest_equals_host_flatten): identical bit for bit
        assert np.array_equal(got_v, want_v), name
        assert np.array_equal(got_n, want_n, equal_nan=True), name
        assert np.isfinite(got_n).all() or name.startswith("quads"), name     # random quads may leave unused vertices (NaN normal)


def test_static_ingest_on_a_caller_stream_and_its_limits():
    import torch
    L = capi.load()
    dev = torch.device("cuda", 0)
    pos, tri = scenegen.icosphere_mesh(3, radius=2.0, seed=4, disp=0.1)
    kw, vblock, fblock = ply_cases.raw_blocks(pos, tri)
    want_v, want_n = _ingest_existing(L, kw, vblock, fblock, 1)
    d_vb = torch.frombuffer(bytearray(vblock), dtype=torch.uint8).to(dev)
    d_fb = torch.frombuffer(bytearray(fblock), dtype=torch.uint8).to(dev)
    d_tv = torch.zeros((len(tri), 9), dtype=torch.float64, device=dev)
    d_tn = torch.zeros((len(tri), 9), dtype=torch.float64, device=dev)
    s = torch.cuda.Stream(dev)
    torch.cuda.synchronize(dev)
    rc, _ = _ingest_static(L, kw, d_vb, d_fb, 1, d_tv, d_tn, stream=s.cuda_stream)
    assert rc == 0, L.rt_last_error().decode()
    assert np.array_equal(d_tv.cpu().numpy(), want_v) and np.array_equal(d_tn.cpu().numpy(), want_n, equal_nan=True)
    # beyond the static scratch: refused before anything runs, with the limits in the message
    big = dict(kw, vertexCount=capi.RT_PLY_STATIC_MAX_VERTICES + 1)
    rc, _ = _ingest_static(L, big, d_vb, d_fb, 1, d_tv, d_tn)
    assert rc == capi.RT_ERR_UNSUPPORTED and b"65536" in L.rt_last_error() and b"rt_ply_ingest" in L.rt_last_error()
    big = dict(kw, faceCount=capi.RT_PLY_STATIC_MAX_TRIANGLES + 1)
    rc, _ = _ingest_static(L, big, d_vb, d_fb, 1, d_tv, d_tn)
    assert rc == capi.RT_ERR_UNSUPPORTED and b"131072" in L.rt_last_error()
    # a face index out of range is caught on the device before it is dereferenced
    bad = np.array(tri, "<i4"); bad[7, 1] = len(pos)
    kwb, vb2, fb2 = ply_cases.raw_blocks(pos, bad)
    rc, _ = _ingest_static(L, kwb, d_vb, torch.frombuffer(bytearray(fb2), dtype=torch.uint8).to(dev), 1, d_tv, d_tn)
    assert rc == capi.RT_ERR_INVALID and b"out of range" in L.rt_last_error()
    # ... and the next call is clean (the error flag is one of the things a call clears)
    rc, _ = _ingest_static(L, kw, d_vb, d_fb, 1, d_tv, d_tn)
    assert rc == 0 and np.array_equal(d_tv.cpu().numpy(), want_v)
