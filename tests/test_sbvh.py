"""CPU: the host SBVH builder (cpu-raytracer_b200/host/SBVH.cpp) against SBVHs built by the REAL
reference's binary (tests/golden/sbvh_reference.npz, made by tests/golden/make_golden_bvh.py):
same node count, same node numbering, every reachable node the same 32 bytes, the same reference
list — so the `<obj>.bvh` written by this library is the reference's. Plus: the result does not
depend on the number of builder threads.
"""
import ctypes as C
import os

import numpy as np
import pytest

from b200rt import _ffi
from conftest import GOLDEN

NODE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("left_or_first", "<i4"), ("count", "<i4")])
MESHES = ["Monkey", "Diamond", "icosphere", "Rock", "Torus", "Magnifier"]


def build_sbvh(vertices, threads):
    lib = _ffi.host_lib()
    lib.b200rt_build_sbvh.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    v = np.ascontiguousarray(vertices, dtype=np.float32)
    nc, ic = C.c_int(), C.c_int()
    assert lib.b200rt_build_sbvh(v.ctypes.data, len(v), threads, None, 0, None, 0, C.byref(nc), C.byref(ic)) == 0
    nodes, indices = np.zeros(nc.value, NODE), np.zeros(ic.value, np.int32)
    assert lib.b200rt_build_sbvh(v.ctypes.data, len(v), threads, nodes.ctypes.data, nc.value, indices.ctypes.data, ic.value, C.byref(nc), C.byref(ic)) == 0
    return nodes, indices


def reachable(nodes):
    seen, stack = [], [0]
    while stack:
        i = stack.pop()
        seen.append(i)
        n = nodes[i]
        if (int(n["count"]) & 0x3FFFFFFF) == 0:
            stack += [int(n["left_or_first"]), int(n["left_or_first"]) + 1]
    return np.array(sorted(seen))


@pytest.fixture(scope="module")
def reference():
    with np.load(os.path.join(GOLDEN, "sbvh_reference.npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", MESHES)
def test_sbvh_is_the_references(reference, name):
    ref_nodes, ref_indices = reference[name + "_nodes"].view(NODE), reference[name + "_indices"]
    nodes, indices = build_sbvh(reference[name + "_vertices"], threads=1)
    assert len(nodes) == len(ref_nodes)
    assert np.array_equal(indices, ref_indices)
    live = reachable(ref_nodes)
    assert np.array_equal(live, reachable(nodes))
    # unreachable slots (node 1, the pairs reserved by SAH-terminated leaves) hold heap garbage in the reference's file
    assert nodes[live].tobytes() == ref_nodes[live].tobytes()
    if name == "Monkey":
        assert len(indices) > len(reference[name + "_vertices"])     # spatial splits duplicated references


def test_sbvh_does_not_depend_on_thread_count():
    rng = np.random.default_rng(7)
    n = 30000   # big enough for the task pool to split the work (subtrees > 4096 references)
    centres = rng.uniform(-50, 50, (n, 1, 3)).astype(np.float32)
    # long thin triangles: plenty of spatial splits
    tris = centres + rng.normal(0, 1, (n, 3, 3)).astype(np.float32) * np.array([8.0, 0.3, 0.3], dtype=np.float32)
    a_nodes, a_idx = build_sbvh(tris, threads=1)
    b_nodes, b_idx = build_sbvh(tris, threads=8)
    assert a_nodes.tobytes() == b_nodes.tobytes()
    assert np.array_equal(a_idx, b_idx)
    assert len(a_idx) > n
    # every triangle is referenced, every leaf range is inside the reference list
    assert len(np.unique(a_idx)) == n
    leaves = a_nodes[(a_nodes["count"] & 0x3FFFFFFF) > 0]
    assert ((leaves["left_or_first"] >= 0) & (leaves["left_or_first"] + leaves["count"] <= len(a_idx))).all()


def _split_cache(blob):
    n = int(blob[:4].view("<i4")[0])
    o = 4
    triangles = blob[o:o + n * 100]; o += n * 100          # TriangleHot[n] (36 B) then TriangleCold[n] (64 B)
    nn = int(blob[o:o + 4].view("<i4")[0]); o += 4
    nodes = blob[o:o + nn * 32].view(NODE); o += nn * 32
    ni = int(blob[o:o + 4].view("<i4")[0]); o += 4
    return n, triangles, nodes, blob[o:o + ni * 4].view("<i4")


def test_polygon_faces_are_triangulated_like_the_reference(reference, tmp_path):
    """An OBJ of polygon faces only (convex, concave, clockwise, tilted, collinear leading corners,
    skew, relative indices) through the host loader: the `<obj>.bvh` it writes holds the triangles —
    order, corner rotation, normals, texture coordinates — and the SBVH of the reference's own
    loader + builder (tinyobjloader's ear clipping, OBJLoader.cpp:97-175)."""
    from b200rt import scenes
    from b200rt.host import HostScene
    obj = tmp_path / "polygons.obj"
    obj.write_bytes(reference["polygons_obj"].tobytes())
    sky = scenes.write_sky(str(tmp_path / "sky.float"), 16)
    s = scenes.SceneWriter(64, 64)
    s.camera((0.0, 1.0, -4.0))
    s.sky(sky)
    s.mesh(str(obj))
    s.point_light((10.0, 10.0, 10.0), (0.0, 5.0, -3.0))
    HostScene(s.write(str(tmp_path / "polygons.scene")))

    count, triangles, nodes, indices = _split_cache(np.fromfile(str(obj) + ".bvh", dtype=np.uint8))
    ref_count, ref_triangles, ref_nodes, ref_indices = _split_cache(reference["polygons_cache"])
    assert count == ref_count and count > 700
    assert triangles.tobytes() == ref_triangles.tobytes()
    assert np.array_equal(indices, ref_indices)
    live = reachable(ref_nodes)
    assert len(nodes) == len(ref_nodes) and np.array_equal(live, reachable(nodes))
    assert nodes[live].tobytes() == ref_nodes[live].tobytes()
