"""CPU: the host library (libb200rt_host.so) — the reference's Scene surface re-hosted.

Pinned against the real reference: every golden fixture stores what the reference's own
Scene::update produced for that frame (instance matrices + inverses, TopLevelBVH nodes + indices,
rotated camera pyramid; dumped by oracle/ref_recipe/headless/Main.cpp). The pack next to it was
produced by OUR host library for the same scene file and frame number; the two must be identical
bytes (Scene.cpp:139-171, Mesh.cpp:9-15, Matrix4.h:88-138, TopLevelBVH.cpp:32-51, Camera.cpp:44-47).
The remaining tests build small procedural scenes (no reference assets needed).
"""
import os

import numpy as np
import pytest

from b200rt import _ffi, scenes, standin
from b200rt.host import HostError, HostScene, tile_count

NODE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("left_or_first", "<i4"), ("count", "<i4")])


def _parse_ref_frame(blob):
    n_inst, n_nodes = np.frombuffer(blob[:8], dtype="<i4")
    o = 8
    per = np.frombuffer(blob[o:o + n_inst * 38 * 4], dtype="<f4").reshape(n_inst, 38)
    o += n_inst * 38 * 4
    nodes = np.frombuffer(blob[o:o + n_nodes * 32], dtype=NODE)
    o += n_nodes * 32
    indices = np.frombuffer(blob[o:o + n_inst * 4], dtype="<i4")
    o += n_inst * 4
    camera = np.frombuffer(blob[o:o + 48], dtype="<f4")
    return dict(world=per[:, :16], inverse=per[:, 16:32], aabb=per[:, 32:], nodes=nodes, indices=indices, camera=camera)


def test_frame_inputs_equal_the_references(golden):
    ref = _parse_ref_frame(golden.ref_frame_inputs.tobytes())
    a = golden.pack.a
    inst = a["f_instances"].view("<f4").reshape(-1, 36)   # world[16], inverse[16], blas + reserved (as bits)
    assert np.array_equal(inst[:, :16].view("<u4"), ref["world"].view("<u4"))
    assert np.array_equal(inst[:, 16:32].view("<u4"), ref["inverse"].view("<u4"))
    assert np.array_equal(a["f_tlas_indices"].view("<i4"), ref["indices"])
    ours = a["f_tlas_nodes"].view(NODE)
    # node 1 is never written by the builder (BottomLevelBVH.cpp:92 / BVHBuilders.h): skip it
    keep = np.arange(len(ref["nodes"])) != 1
    assert len(ours) >= len(ref["nodes"])
    assert ours[:len(ref["nodes"])][keep].tobytes() == ref["nodes"][keep].tobytes()
    assert np.array_equal(a["f_camera"].view("<u4"), ref["camera"].view("<u4"))


@pytest.fixture(scope="module")
def small_scene(tmp_path_factory):
    d = tmp_path_factory.mktemp("scene")
    standin.lens(str(d / "lens.obj"), n=16)
    standin.mirror(str(d / "mirror.obj"), n=8)
    scenes.write_sky(str(d / "sky.float"), 32)
    s = scenes.SceneWriter(96, 64)
    s.camera((14.0, 3.0, 0.0), (0.0, -0.707107, 0.0, 0.707107))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(str(d / "sky.float"))
    s.sphere(1.0, (0.0, 1.0, 3.0), diffuse=(0.2, 0.2, 0.0), reflection=(0.6, 0.6, 0.0), transmittance=(0.6, 0.6, 0.6), ior=1.33)
    s.plane((0.0, -1.0, 0.0), reflection=(0.1, 0.1, 0.1))
    s.mesh(str(d / "lens.obj"), (5.0, 3.0, 0.0))
    s.mesh(str(d / "lens.obj"), (2.0, 3.0, -4.0))           # shares the BLAS of the first (BottomLevelBVH.cpp:16-22)
    s.mesh(str(d / "mirror.obj"), (-6.0, 0.0, 0.0))
    s.point_light((50.0, 50.0, 50.0), (6.0, 12.0, 6.0))
    s.directional_light((0.5, 0.5, 0.5), (-0.3, -1.0, 0.2))
    path = s.write(str(d / "small.scene"))
    return d, path


def _check_bvh(nodes, primitive_count, leaf_boxes=None):
    """Structural invariants of the reference's node array: root 0, node 1 unused, children adjacent."""
    seen = np.zeros(primitive_count, dtype=np.int32)
    stack = [0]
    while stack:
        n = nodes[stack.pop()]
        count = int(n["count"]) & 0x3FFFFFFF
        if count > 0:
            first = int(n["left_or_first"])
            assert 0 <= first and first + count <= primitive_count
            seen[first:first + count] += 1
        else:
            axis = (int(n["count"]) >> 30) & 3
            assert axis in (1, 2, 3)
            left = int(n["left_or_first"])
            assert left >= 2 and left + 1 < len(nodes)
            for c in (left, left + 1):
                assert (nodes[c]["min"] >= n["min"] - 1e-5).all() and (nodes[c]["max"] <= n["max"] + 1e-5).all()
                stack.append(c)
    assert (seen == 1).all()


def test_scene_load_shares_blas_and_builds_valid_bvhs(small_scene):
    _, path = small_scene
    scene = HostScene(path)
    scene.update(0.0, 0)
    pack = scene.pack()
    blas = pack.blas()
    assert len(blas) == 2                                    # two lens instances share one BLAS
    inst = pack.a["f_instances"].view("<i4").reshape(-1, 36)
    assert inst[:, 32].tolist() == [0, 0, 1]
    nodes = pack.nodes()
    for node_offset, node_count, tri_offset, tri_count, _ in blas:
        _check_bvh(nodes[node_offset:node_offset + node_count], tri_count)
    _check_bvh(pack.a["f_tlas_nodes"].view(NODE), 3)
    assert sorted(pack.a["f_tlas_indices"].view("<i4").tolist()) == [0, 1, 2]
    # materials: 0 = default black (Material.h:52-60), then sphere, plane, then OBJ materials in load order
    mats = pack.a["s_materials"].view("<f4").reshape(-1, 12)
    assert (mats[0, :3] == 0).all()
    assert np.allclose(mats[1, 4:7], (0.6, 0.6, 0.0)) and np.isclose(mats[1, 7], 1.33)
    scene.close()


def test_bvh_cache_roundtrip(small_scene):
    """`<obj>.bvh` (BottomLevelBVH.cpp:149-192): written on first load, read back to identical arrays."""
    d, path = small_scene
    HostScene(path).close()
    cache = d / "lens.obj.bvh"
    assert cache.exists()
    blob = cache.read_bytes()
    tri = int(np.frombuffer(blob[:4], "<i4")[0])
    o = 4 + tri * (36 + 64)
    nodes = int(np.frombuffer(blob[o:o + 4], "<i4")[0])
    o += 4 + nodes * 32
    idx = int(np.frombuffer(blob[o:o + 4], "<i4")[0])
    assert len(blob) == o + 4 + idx * 4 and idx >= tri       # headerless layout of SURVEY App. B
    first = HostScene(path).pack()                            # from cache
    mtime = cache.stat().st_mtime_ns
    second = HostScene(path).pack()
    assert cache.stat().st_mtime_ns == mtime                  # not rebuilt
    for key in ("s_nodes", "s_triangles_hot", "s_triangles_cold", "s_blas"):
        assert np.array_equal(first.a[key], second.a[key]), key


def test_original_index_map_is_a_cover(small_scene):
    _, path = small_scene
    scene = HostScene(path)
    pack = scene.pack()
    for b, (_, _, _, tri_count, _) in enumerate(pack.blas()):
        m = scene.blas_original_indices(b)
        assert len(m) == tri_count and m.min() == 0
        assert len(np.unique(m)) == m.max() + 1                # every source triangle is referenced


def test_update_is_deterministic_and_camera_resizes(small_scene):
    _, path = small_scene
    scene = HostScene(path)
    scene.update(0.0, 0)
    a = scene.pack()
    scene.update(0.0, 0)
    b = scene.pack()
    for key in a.a:
        assert np.array_equal(a.a[key], b.a[key]), key
    scene.resize(192, 128)
    scene.update(0.0, 0)
    c = scene.pack()
    assert c.width == 192 and c.height == 128
    # Camera::resize (Camera.cpp:5-16): the pyramid's x axis stays one pixel wide
    assert np.allclose(c.a["f_camera"][6:9], a.a["f_camera"][6:9])
    assert not np.allclose(c.a["f_camera"][3:6], a.a["f_camera"][3:6])


def test_errors_are_reported_not_aborted(tmp_path):
    with pytest.raises(HostError):
        HostScene(str(tmp_path / "missing.scene"))
    bad = tmp_path / "bad.scene"
    bad.write_text("resolution 64 64\nmesh /nonexistent/file.obj 0 0 0 0 0 0 1\n")
    with pytest.raises(HostError):
        HostScene(str(bad))


def test_tile_count_matches_window_cpp():
    # Window.cpp:11-12; SURVEY 8: 1280x720 -> 40x23, 1920x1080 -> 60x34, 3840x2160 -> 120x68
    assert tile_count(1280, 720) == (920, 40, 23)
    assert tile_count(1920, 1080) == (2040, 60, 34)
    assert tile_count(3840, 2160) == (8160, 120, 68)
    assert tile_count(33, 1) == (2, 2, 1)
