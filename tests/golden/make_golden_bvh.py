#!/usr/bin/env python3
"""Generates tests/golden/sbvh_reference.npz: SBVHs built by the REAL reference.

Run in the build container only (needs oracle/_ref, see oracle/ref_recipe). For every shipped mesh
the `<obj>.bvh` cache is deleted, the reference's own binary (the -ffp-contract=off "_ids" build,
whose scalar builder code rounds like plain IEEE arithmetic) rebuilds it with its SBVH builder
(BottomLevelBVH.cpp:108-147) and the cache is read back: triangle vertices as the reference's
loader saw them are not in the file (it stores p0, p1-p0, p2-p0), so the fixture keeps the
vertices parsed from the OBJ (float32; these meshes are triangles only) next to the reference's
nodes + indices. Polygon faces: polygon_obj() below, kept with the reference's whole cache file.
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
MESHES = ["Monkey", "Diamond", "icosphere", "Rock", "Torus", "Magnifier"]   # triangle meshes; polygon faces are covered by polygon_obj() below
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


def polygon_obj(seed=99):
    """An OBJ made of polygon faces only — convex, concave, clockwise, tilted, with collinear leading
    corners, skew quads, relative indices — to pin the loader's triangulation to tinyobjloader's
    ear clipping as the reference runs it (OBJLoader.cpp:97, tiny_obj_loader.h:1355-1557)."""
    rng = np.random.default_rng(seed)
    lines, counts = ["# polygon faces for the triangulation test"], [0, 0, 0]

    def basis(normal):
        n = np.asarray(normal, dtype=np.float64); n /= np.linalg.norm(n)
        a = np.cross(n, [0.0, 0.0, 1.0] if abs(n[2]) < 0.9 else [1.0, 0.0, 0.0]); a /= np.linalg.norm(a)
        return n, a, np.cross(n, a)

    def face(points2d, normal, origin, relative=False, skew=0.0):
        n, a, b = basis(normal)
        first = counts[0]
        for i, (x, y) in enumerate(points2d):
            p = np.asarray(origin) + a * x + b * y + n * (skew * (-1) ** i)
            lines.append("v %.6f %.6f %.6f" % tuple(p)); counts[0] += 1
            lines.append("vt %.6f %.6f" % (0.1 * x, 0.1 * y)); counts[1] += 1
        lines.append("vn %.6f %.6f %.6f" % tuple(n)); counts[2] += 1
        k = len(points2d)
        if relative:
            lines.append("f " + " ".join("%d/%d/-1" % (i - k, i - k) for i in range(k)))
        else:
            lines.append("f " + " ".join("%d/%d/%d" % (first + i + 1, first + i + 1, counts[2]) for i in range(k)))

    square = [(0, 0), (2, 0), (2, 2), (0, 2)]
    ell = [(0, 0), (3, 0), (3, 1), (1, 1), (1, 3), (0, 3)]
    star = [((2.0 if i % 2 == 0 else 0.8) * np.cos(np.pi * i / 5), (2.0 if i % 2 == 0 else 0.8) * np.sin(np.pi * i / 5)) for i in range(10)]
    dart = [(0, 0), (2, 1), (0, 2), (0.7, 1)]
    collinear = [(0, 0), (1, 0), (2, 0), (3, 0), (3, 2), (1.5, 0.5), (0, 2)]
    comb = [(0, 0), (8, 0)] + [p for i in range(4) for p in ((8 - 2 * i, 3), (7 - 2 * i, 1), (6 - 2 * i, 3))][:-1] + [(0, 3)]
    ring = [(2 * np.cos(2 * np.pi * i / 12), 2 * np.sin(2 * np.pi * i / 12)) for i in range(12)]
    shapes = [square, square[::-1], ell, ell[::-1], star, dart, dart[::-1], collinear, comb, comb[::-1], ring, [(0, 0), (1, 0), (0, 1)]]
    normals = [(0, 0, 1), (0, 1, 0), (1, 0, 0), (1, 1, 1), (0.2, -1, 0.3), (-1, 0.1, 0.05), (0.5, 0.5, -1)]
    at = 0
    for i, shape in enumerate(shapes):
        for j, normal in enumerate(normals):
            face(shape, normal, (10.0 * i, 10.0 * j, 3.0 * ((i + j) % 3)), relative=(i + j) % 5 == 0, skew=0.15 if (i * 7 + j) % 4 == 0 else 0.0)
            at += 1
    for i in range(40):   # star-shaped (hence simple) random polygons, random orientation
        k = int(rng.integers(4, 16))
        angles = np.sort(rng.uniform(0, 2 * np.pi, k))
        radii = rng.uniform(0.4, 3.0, k)
        pts = [(r * np.cos(t), r * np.sin(t)) for r, t in zip(radii, angles)]
        if i % 3 == 0:
            pts = pts[::-1]
        face(pts, rng.normal(size=3), rng.uniform(-40, 40, 3), skew=0.05 if i % 7 == 0 else 0.0)
    return "\n".join(lines) + "\n"


def polygon_case(tmp, out):
    text = polygon_obj()
    obj = os.path.join(tmp, "polygons.obj")
    with open(obj, "w") as f:
        f.write(text)
    s = scenes.SceneWriter(64, 64)
    s.camera((0.0, 1.0, -4.0))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(DATA, "sky_256.float"))
    s.mesh(obj)
    s.point_light((10.0, 10.0, 10.0), (0.0, 5.0, -3.0))
    scene = s.write(os.path.join(tmp, "polygons.scene"))
    subprocess.run([BIN, scene, "--frames", "1", "--width", "64", "--height", "64"], check=True, capture_output=True)
    blob = np.fromfile(obj + ".bvh", dtype=np.uint8)
    out["polygons_obj"] = np.frombuffer(text.encode(), dtype=np.uint8).copy()
    out["polygons_cache"] = blob
    faces = sum(1 for l in text.splitlines() if l.startswith("f "))
    print("polygons   %6d faces -> %d triangles by the reference's loader" % (faces, int(blob[:4].view("<i4")[0])))


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
        polygon_case(tmp, out)
    np.savez_compressed(os.path.join(HERE, "sbvh_reference.npz"), **out)


if __name__ == "__main__":
    main()
