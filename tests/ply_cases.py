"""PLY files for the ingest tests: the layouts the device path takes (rt_ply_ingest) and some it leaves to the host."""
import os

import numpy as np

from myraytracer_b200 import scenegen


def _mesh_scene(out_dir, name, ply, smooth):
    sc = scenegen.Scene(maxRecursionDepth=0, backgroundColor=(1, 2, 3), ambient=(10, 10, 10))
    sc.materials = [scenegen.Material(id=1)]
    sc.objects = [scenegen.SceneObject(scenegen.OBJ_MESH, id=1, material=1, smooth=smooth, plyPath=ply)]
    sc.pointLights = [scenegen.PointLight((0, 8, 8), (500, 500, 500))]
    sc.cameras = [scenegen.Camera(id=1, width=16, height=16, numSamples=1, position=(0, 0, 9), gazePoint=(0, 0, 0), up=(0, 1, 0),
                                  nearDistance=1.0, fovy=50.0, isLookAt=True)]
    path = os.path.join(out_dir, name + ".rtscene")
    scenegen.write_rtscene(sc, path)
    return path


def _write(path, header, vertex_bytes, face_bytes):
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode())
        f.write(vertex_bytes)
        f.write(face_bytes)


def make_cases(out_dir, seed=9):
    """-> list of (name, scene path, expected triangle count, taken by the device path?)"""
    rng = np.random.default_rng(seed)
    pos, tris = scenegen.icosphere_mesh(2, radius=1.5, seed=seed, disp=0.2)
    pos = pos.astype("<f4")
    tris = tris.astype("<i4")
    out = []

    def add(name, smooth, n_tris, device):
        out.append((name, _mesh_scene(out_dir, name, name + ".ply", smooth), n_tris, device))

    # plain triangles, smooth and flat
    scenegen.write_ply(os.path.join(out_dir, "tri_smooth.ply"), pos, tris)
    add("tri_smooth", True, len(tris), True)
    scenegen.write_ply(os.path.join(out_dir, "tri_flat.ply"), pos, tris)
    add("tri_flat", False, len(tris), True)
    # normals in the file (not unit length: the loader normalises)
    scenegen.write_ply(os.path.join(out_dir, "tri_normals.ply"), pos, tris, normals=pos * 1.7 + 0.1)
    add("tri_normals", True, len(tris), True)
    # quads (split 0,1,3 / 2,3,1), extra vertex properties between the coordinates, scalar properties around the list
    nq = 40
    quads = rng.integers(0, len(pos), size=(nq, 4)).astype("<i4")
    vrec = np.zeros(len(pos), dtype=[("x", "<f4"), ("r", "u1"), ("y", "<f4"), ("q", "<f8"), ("z", "<f4"), ("g", "<i2")])
    vrec["x"], vrec["y"], vrec["z"] = pos[:, 0], pos[:, 1], pos[:, 2]
    vrec["r"] = 7; vrec["q"] = 1.25; vrec["g"] = -3
    frec = np.zeros(nq, dtype=[("flag", "<i2"), ("n", "u1"), ("i", "<i4", (4,)), ("w", "<f4")])
    frec["flag"] = 5; frec["n"] = 4; frec["i"] = quads; frec["w"] = 0.5
    _write(os.path.join(out_dir, "quads_extra.ply"),
           ["ply", "format binary_little_endian 1.0", f"element vertex {len(pos)}", "property float x", "property uchar r",
            "property float y", "property double q", "property float z", "property short g", f"element face {nq}",
            "property short flag", "property list uchar int vertex_indices", "property float w", "end_header"],
           vrec.tobytes(), frec.tobytes())
    add("quads_extra", True, 2 * nq, True)
    # a vertex no face uses (its smooth normal is NaN in the reference too, and never read)
    scenegen.write_ply(os.path.join(out_dir, "unused_vertex.ply"), np.concatenate([pos, [[9, 9, 9]]]).astype("<f4"), tris)
    add("unused_vertex", True, len(tris), True)
    # left to the host loader: ascii; mixed triangles and quads
    with open(os.path.join(out_dir, "ascii.ply"), "w") as f:
        f.write("\n".join(["ply", "format ascii 1.0", f"element vertex {len(pos)}", "property float x", "property float y",
                           "property float z", f"element face {len(tris)}", "property list uchar int vertex_indices", "end_header"]) + "\n")
        for p in pos:
            f.write(f"{float(p[0])!r} {float(p[1])!r} {float(p[2])!r}\n")
        for t in tris:
            f.write(f"3 {t[0]} {t[1]} {t[2]}\n")
    add("ascii", True, len(tris), False)
    mixed = b"".join((np.uint8(3).tobytes() + tris[k].tobytes()) if k % 2 else (np.uint8(4).tobytes() + quads[k % nq].tobytes())
                     for k in range(30))
    _write(os.path.join(out_dir, "mixed.ply"),
           ["ply", "format binary_little_endian 1.0", f"element vertex {len(pos)}", "property float x", "property float y",
            "property float z", "element face 30", "property list uchar int vertex_indices", "end_header"], pos.tobytes(), mixed)
    add("mixed", False, 15 + 2 * 15, False)
    # an index past the vertex array: the mesh is dropped with a warning, by either path
    bad = tris.copy()
    bad[3, 1] = len(pos) + 5
    scenegen.write_ply(os.path.join(out_dir, "bad_index.ply"), pos, bad)
    add("bad_index", True, 0, True)
    return out


# ---- raw element blocks for rt_ply_ingest / rt_ply_ingest_static (no file in between) --------------------------------
def raw_blocks(pos, faces, normals=None):
    """(layout kwargs, vertex block bytes, face block bytes) of a binary-little-endian mesh: float32 x, y, z [nx, ny, nz]
    per vertex; `uchar count, int32 indices` per face (triangles or quads)."""
    pos = np.ascontiguousarray(pos, "<f4")
    faces = np.ascontiguousarray(faces, "<i4")
    k = faces.shape[1]
    if normals is None:
        vblock, stride, noff = pos.tobytes(), 12, (-1, -1, -1)
    else:
        vblock, stride, noff = np.hstack([pos, np.ascontiguousarray(normals, "<f4")]).astype("<f4").tobytes(), 24, (12, 16, 20)
    frec = np.zeros(len(faces), dtype=[("n", "u1"), ("i", "<i4", (k,))])
    frec["n"] = k
    frec["i"] = faces
    layout = dict(vertexCount=len(pos), faceCount=len(faces), vertexStride=stride, posOffset=(0, 4, 8), normalOffset=noff,
                  faceLeadBytes=0, faceCountBytes=1, faceIndexBytes=4, faceTrailBytes=0, verticesPerFace=k)
    return layout, vblock, frec.tobytes()


def make_layout(capi, kw):
    L = capi.rt_ply_layout()
    for name, v in kw.items():
        if isinstance(v, tuple):
            for a, x in enumerate(v):
                getattr(L, name)[a] = x
        else:
            setattr(L, name, v)
    return L
