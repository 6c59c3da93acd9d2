#!/usr/bin/env python3
"""Generates tests/golden/sbvh_reference.npz: SBVHs built by the REAL reference.

Run in the build container only (needs oracle/_ref, see oracle/ref_recipe). For every shipped mesh
the `<obj>.bvh` cache is deleted, the reference's own binary (the -ffp-contract=off "_ids" build,
whose scalar builder code rounds like plain IEEE arithmetic) rebuilds it with its SBVH builder
(BottomLevelBVH.cpp:108-147) and the cache is read back: triangle vertices as the reference's
loader saw them are not in the file (it stores p0, p1-p0, p2-p0), so the fixture keeps the
vertices parsed from the OBJ (float32, fan triangulation) next to the reference's nodes + indices.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
from b200rt import scenes  # noqa: E402

DATA = os.path.join(REPO, "oracle", "_ref", "Data")
BIN = os.path.join(REPO, "oracle", "_ref", "bin", "ref_l8_b3_aniso_ids")
MESHES = ["Monkey", "Diamond", "icosphere", "Rock", "Torus", "Magnifier"]   # Concave.obj has polygon faces (tinyobj ear-clips them): not reproducible from a fan
NODE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("left_or_first", "<i4"), ("count", "<i4")])


def obj_triangles(path):
    v, tris = [], []
    for line in open(path):
        p = line.split()
        if not p:
            continue
        if p[0] == "v":
            v.append([np.float32(x) for x in p[1:4]])
        elif p[0] == "f":
            idx = [int(t.split("/")[0]) for t in p[1:]]
            idx = [i - 1 if i > 0 else len(v) + i for i in idx]
            for k in range(1, len(idx) - 1):
                tris.append([v[idx[0]], v[idx[k]], v[idx[k + 1]]])
    return np.array(tris, dtype=np.float32)


def read_cache(path):
    blob = open(path, "rb").read()
    n = int(np.frombuffer(blob[:4], "<i4")[0])
    o = 4
    hot = np.frombuffer(blob[o:o + n * 36], "<f4").reshape(n, 9); o += n * 100
    nn = int(np.frombuffer(blob[o:o + 4], "<i4")[0]); o += 4
    nodes = np.frombuffer(blob[o:o + nn * 32], NODE); o += nn * 32
    ni = int(np.frombuffer(blob[o:o + 4], "<i4")[0]); o += 4
    indices = np.frombuffer(blob[o:o + ni * 4], "<i4")
    return hot, nodes, indices


def main():
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name in MESHES:
            obj = os.path.join(DATA, name + ".obj")
            cache = obj + ".bvh"
            if os.path.exists(cache):
                os.remove(cache)
            s = scenes.SceneWriter(64, 64)
            s.camera((0.0, 1.0, -4.0))
            s.ambient((0.2, 0.2, 0.2))
            s.sky(os.path.join(DATA, "sky_256.float"))
            s.mesh(obj)
            s.point_light((10.0, 10.0, 10.0), (0.0, 5.0, -3.0))
            scene = s.write(os.path.join(tmp, name + ".scene"))
            subprocess.run([BIN, scene, "--frames", "1", "--width", "64", "--height", "64"], check=True, capture_output=True)
            hot, nodes, indices = read_cache(cache)
            tris = obj_triangles(obj)
            assert len(tris) == len(hot), (name, len(tris), len(hot))
            assert np.array_equal(tris[:, 0], hot[:, 0:3]) and np.array_equal(tris[:, 1] - tris[:, 0], hot[:, 3:6]), name
            out[name + "_vertices"] = tris
            out[name + "_nodes"] = nodes.copy()
            out[name + "_indices"] = indices.copy()
            print("%-10s %6d triangles %6d nodes %6d references" % (name, len(tris), len(nodes), len(indices)))
    np.savez_compressed(os.path.join(HERE, "sbvh_reference.npz"), **out)


if __name__ == "__main__":
    main()
