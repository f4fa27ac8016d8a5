"""Small scenes at the edges of the render loop's input space (the reference has no tests; these follow its code paths:
sample-count rules Object+Extension.swift:298-306,354; empty TLAS :98-100 / RTContext.swift:416; single-instance TLAS;
image sizes that are not multiples of the 8-row chunk or the 8x4 tile)."""
import os

import numpy as np

from myraytracer_b200 import scenegen as G


def _base(width, height, samples, depth=0):
    sc = G.Scene(maxRecursionDepth=depth, backgroundColor=(30, 60, 90), ambient=(25, 25, 25))
    sc.materials = [G.Material(id=1, ambient=(0.5, 0.5, 0.5), diffuse=(0.7, 0.6, 0.5), specular=(0.3, 0.3, 0.3), phong=20.0),
                    G.Material(id=2, type=G.MAT_MIRROR, ambient=(0, 0, 0), diffuse=(0.1, 0.1, 0.1), specular=(0.2, 0.2, 0.2),
                               mirror=(0.8, 0.8, 0.8), phong=50.0)]
    sc.pointLights = [G.PointLight((4.0, 8.0, 6.0), (9000, 9000, 9000))]
    sc.cameras = [G.Camera(id=1, imageName="edge.png", width=width, height=height, numSamples=samples, position=(0.0, 2.0, 9.0),
                           gazePoint=(0.0, 0.5, 0.0), up=(0, 1, 0), nearDistance=1.0, fovy=45.0, isLookAt=True)]
    return sc


def _objects(sc, what):
    sc.vertexData = np.asarray([(0.0, 1.0, 0.0), (-6, 0, -6), (6, 0, -6), (6, 0, 6), (-6, 0, 6), (2.2, 0.8, 1.0)], np.float64)
    if what == "empty":
        sc.objects = []
    elif what == "sphere":
        sc.objects = [G.SceneObject(G.OBJ_SPHERE, id=1, material=1, centerIdx=1, radius=1.0)]
    else:
        sc.objects = [G.SceneObject(G.OBJ_SPHERE, id=1, material=2 if what == "mirror" else 1, centerIdx=1, radius=1.0),
                      G.SceneObject(G.OBJ_TRIANGLE, id=2, material=1, indices=(2, 4, 3)),
                      G.SceneObject(G.OBJ_TRIANGLE, id=3, material=1, indices=(2, 5, 4)),
                      G.SceneObject(G.OBJ_SPHERE, id=4, material=1, centerIdx=6, radius=0.8)]


EDGE_CASES = {
    # name: (objects, width, height, samples, depth)
    "empty_scene": ("empty", 24, 16, 4, 2),          # no instances: every ray returns the background
    "single_instance": ("sphere", 40, 24, 4, 0),     # the TLAS is one leaf
    "image_1x1": ("scene", 1, 1, 4, 0),
    "image_9x5": ("scene", 9, 5, 1, 0),              # smaller than one 8-row chunk / one tile in a dimension
    "image_33x17": ("scene", 33, 17, 4, 0),          # one pixel past a tile in both dimensions
    "spp2": ("scene", 32, 24, 2, 0),                 # n = 1: one ray, sum divided by 2 (:300-306,354)
    "spp8": ("scene", 32, 24, 8, 0),                 # n = 2: four rays, divided by 8
    "spp3_mirror": ("mirror", 32, 24, 3, 3),         # n = 1 with recursion
    "depth0_mirror": ("mirror", 32, 24, 4, 0),       # maxRecursionDepth 0: no secondary rays (:189)
}


def make(out_dir, name):
    what, w, h, spp, depth = EDGE_CASES[name]
    sc = _base(w, h, spp, depth)
    _objects(sc, what)
    path = os.path.join(out_dir, f"edge_{name}.rtscene")
    G.write_rtscene(sc, path)
    return path
