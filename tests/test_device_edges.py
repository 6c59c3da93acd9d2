"""GPU: edges of the C ABI on a live device.

* Frame sizes that are not tile multiples (the reference clips edge tiles, WorkerThread.cpp:57-61)
  down to a single pixel: the ray of pixel (x, y) depends on the camera vectors only
  (Raytracer.cpp:28-47), so a smaller frame with the same camera is the top-left crop of the larger
  one, bit for bit, and equals the reference's golden image there.
* Scenes stripped to the minimum: no lights, no spheres, no planes.
* Repeated frames are deterministic although tiles are handed out dynamically.
* Bad input is an error code + message (never an abort) and leaves the context usable.
"""
import numpy as np
import pytest

from conftest import Golden, unpack_rgb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dynamic():
    return Golden("dynamic_b3_aniso_f45")


def _renderer(golden, width=None, height=None, **kw):
    from b200rt.device import Renderer
    m = golden.meta
    return Renderer(width or m["width"], height or m["height"], kw.pop("bounces", m["bounces"]), m["filter"], **kw)


def _with(pack, **arrays):
    from b200rt.pack import ScenePack
    a = dict(pack.a)
    a.update(arrays)
    return ScenePack(a)


@pytest.mark.parametrize("size", [(1, 1), (7, 3), (33, 31), (100, 7), (479, 269), (480, 1)])
def test_smaller_frames_are_crops(dynamic, size):
    import oracle_ffi
    w, h = size
    full = _renderer(dynamic)
    full.upload_static(dynamic.pack)
    whole = full.render(dynamic.pack, want_ids=True)
    full.close()

    r = _renderer(dynamic, w, h)
    r.upload_static(dynamic.pack)
    part = r.render(dynamic.pack, want_ids=True)
    r.close()

    assert np.array_equal(part["frame_buffer"], whole["frame_buffer"][:h, :w])
    assert np.array_equal(part["ids"], whole["ids"][:h, :w])
    assert np.array_equal(part["ids"], dynamic.ref_ids[:h, :w])
    assert np.array_equal(part["hit"][..., 0].view(np.uint32), dynamic.ref_hit[:h, :w, 0].view(np.uint32))
    assert part["stats"]["primary_rays"] == w * h
    # FXAA wraps at the frame border (REPEAT sampler, Window.cpp:48-50), so the presented image is NOT a crop: check it against the oracle's pass
    assert np.array_equal(part["presented"], oracle_ffi.present(part["frame_buffer"], True))
    a, b = unpack_rgb(part["frame_buffer"]), unpack_rgb(dynamic.ref_frame_buffer[:h, :w])
    assert np.abs(a - b).max() <= 8 and (np.abs(a - b).max(-1) > 0).sum() <= max(1, w * h // 500)


@pytest.mark.parametrize("drop", ["lights", "spheres", "planes", "all"])
def test_minimal_scenes_equal_the_oracle(dynamic, drop):
    import oracle_ffi
    empty = np.zeros(0, dtype=np.uint8)
    gone = {"lights": ["f_lights"], "spheres": ["f_spheres"], "planes": ["f_planes"], "all": ["f_lights", "f_spheres", "f_planes"]}[drop]
    pack = _with(dynamic.pack, **{k: empty for k in gone})
    m = dynamic.meta
    r = _renderer(dynamic)
    r.upload_static(pack)
    out = r.render(pack, want_ids=True)
    r.close()
    ref = oracle_ffi.render(pack, m["bounces"], m["filter"])
    assert np.array_equal(out["ids"], ref["ids"])
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    assert (np.abs(a - b).max(-1) > 0).mean() <= 2e-3 and np.abs(a - b).max() <= 8
    if drop in ("lights", "all"):
        assert out["stats"]["shadow_rays"] == 0 == ref["stats"]["shadow_rays"]


def test_zero_bounces_is_direct_lighting_only(dynamic):
    import oracle_ffi
    r = _renderer(dynamic, bounces=0)
    r.upload_static(dynamic.pack)
    out = r.render(dynamic.pack)
    r.close()
    ref = oracle_ffi.render(dynamic.pack, 0, dynamic.meta["filter"])
    assert out["stats"]["reflection_rays"] == 0 and out["stats"]["refraction_rays"] == 0
    assert out["stats"]["shadow_rays"] == ref["stats"]["shadow_rays"]
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    assert (np.abs(a - b).max(-1) > 0).mean() <= 2e-3 and np.abs(a - b).max() <= 8


def test_repeated_frames_are_bit_identical(dynamic):
    r = _renderer(dynamic)
    r.upload_static(dynamic.pack)
    first = r.render(dynamic.pack, want_ids=True, want_float=True)
    for _ in range(4):                                  # the tile order changes from frame to frame (cost-ordered hand-out)
        again = r.render(dynamic.pack, want_ids=True, want_float=True)
        for key in ("frame_buffer", "presented", "ids"):
            assert np.array_equal(first[key], again[key]), key
        assert np.array_equal(first["colour"].view(np.uint32), again["colour"].view(np.uint32))
        assert np.array_equal(first["hit"].view(np.uint32), again["hit"].view(np.uint32))
        for key in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays"):
            assert first["stats"][key] == again["stats"][key]
    r.close()


def _corrupt(pack, what):
    a = {k: v.copy() for k, v in pack.a.items()}
    if what == "child index":
        nodes = a["s_nodes"].view("<i4").reshape(-1, 8)
        inner = np.flatnonzero((nodes[:, 7] & 0x3FFFFFFF) == 0)
        nodes[inner[0], 6] = len(nodes) + 5
    elif what == "leaf range":
        nodes = a["s_nodes"].view("<i4").reshape(-1, 8)
        leaf = np.flatnonzero((nodes[:, 7] & 0x3FFFFFFF) > 0)
        nodes[leaf[-1], 6] = 1 << 28
    elif what == "material id":
        a["s_triangles_cold"].view("<i4").reshape(-1, 16)[0, 15] = 100000
    elif what == "instance blas":
        a["f_instances"].view("<i4").reshape(-1, 36)[0, 32] = 77       # b200rt_instance: world[16], inverse[16], blas, reserved[3]
    elif what == "tlas child":
        nodes = a["f_tlas_nodes"].view("<i4").reshape(-1, 8)
        nodes[0, 6], nodes[0, 7] = 4096, nodes[0, 7] & ~0x3FFFFFFF
    elif what == "light order":
        lights = a["f_lights"].view("<i4").reshape(-1, 12)              # point, spot, directional -> directional, spot, point
        a["f_lights"] = lights[::-1].copy().view(np.uint8).reshape(-1)
    from b200rt.pack import ScenePack
    return ScenePack(a)


@pytest.mark.parametrize("what", ["child index", "leaf range", "material id"])
def test_bad_static_scene_is_refused(dynamic, what):
    from b200rt.device import DeviceError
    r = _renderer(dynamic)
    with pytest.raises(DeviceError) as e:
        r.upload_static(_corrupt(dynamic.pack, what))
    assert "out of" in str(e.value)
    r.upload_static(dynamic.pack)                       # the context is still usable
    assert np.array_equal(r.render(dynamic.pack, want_ids=True)["ids"], dynamic.ref_ids)
    r.close()


@pytest.mark.parametrize("what", ["instance blas", "tlas child", "light order"])
def test_bad_frame_is_refused(dynamic, what):
    from b200rt.device import DeviceError
    r = _renderer(dynamic)
    r.upload_static(dynamic.pack)
    with pytest.raises(DeviceError) as e:
        r.render(_corrupt(dynamic.pack, what))
    assert str(e.value)
    assert np.array_equal(r.render(dynamic.pack, want_ids=True)["ids"], dynamic.ref_ids)
    r.close()


def test_frame_before_static_scene_is_refused(dynamic):
    from b200rt.device import DeviceError
    r = _renderer(dynamic)
    with pytest.raises(DeviceError, match="upload_static"):
        r.render(dynamic.pack)
    r.close()


def test_leaves_larger_than_a_shadow_record(dynamic):
    """The shadow rays walk a threaded copy of every BLAS whose leaf records hold at most 63
    triangles (threaded_nodes.h: build_threaded_nodes); a larger leaf becomes a chain of records. Collapse
    two BLAS of the scene into a single root leaf each (514 and 24+ triangles) and compare with
    the oracle, which walks the reference's own node array."""
    import oracle_ffi
    a = {k: v.copy() for k, v in dynamic.pack.a.items()}
    nodes = a["s_nodes"].view("<i4").reshape(-1, 8)
    blas = dynamic.pack.blas()                       # node_offset, node_count, triangle_offset, triangle_count, material_offset
    collapsed = 0
    for node_offset, node_count, triangle_offset, triangle_count, _ in blas:
        if triangle_count in sorted(blas[:, 3])[-2:] or triangle_count < 64:
            nodes[node_offset, 6] = 0                # first triangle
            nodes[node_offset, 7] = triangle_count   # leaf: every triangle of the BLAS
            collapsed += 1
    assert collapsed >= 2 and blas[:, 3].max() > 63
    from b200rt.pack import ScenePack
    pack = ScenePack(a)
    m = dynamic.meta
    r = _renderer(dynamic)
    r.upload_static(pack)
    out = r.render(pack, want_ids=True)
    r.close()
    ref = oracle_ffi.render(pack, m["bounces"], m["filter"])
    assert np.array_equal(out["ids"], ref["ids"])
    assert out["stats"]["shadow_rays"] == ref["stats"]["shadow_rays"]
    x, y = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    assert (np.abs(x - y).max(-1) > 0).mean() <= 2e-3 and np.abs(x - y).max() <= 8
