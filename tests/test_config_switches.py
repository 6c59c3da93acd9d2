"""The reference's compile-time switches on the hot path, as run-time fields of b200rt_config:

    BVH_TRAVERSAL_STRATEGY   (Config.h; BVHNode.h:30-40)          b200rt_config.traversal_naive
    TEXTURE_SAMPLE_MODE      (Config.h; Texture.h:33-49)          b200rt_config.texture_sample_mode

The "switch_" fixtures hold what the real reference produced when it was compiled with the other
setting (tests/golden/make_golden.py; oracle/ref_recipe/build_ref.py variants _naive, _nearest,
_bilinear). The CPU tests pin the oracle to them, the GPU tests the device path: ids, hit distances
and the 8-bit image bit for bit, and the traversal counters against the oracle (the two orders visit
different numbers of nodes and triangles, so equal counters show the order itself is the reference's).
"""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, Golden

SWITCH_NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "switch_*.npz")))


@pytest.fixture(scope="module", params=SWITCH_NAMES)
def switched(request):
    return Golden(request.param)


def _oracle(g, **kw):
    import oracle_ffi
    m = g.meta
    return oracle_ffi.render(g.pack, m["bounces"], m["filter"], traversal=m["traversal"], texture_mode=m["texture_mode"], **kw)


def test_fixtures_cover_every_setting():
    metas = [Golden(n).meta for n in SWITCH_NAMES]
    assert {m["traversal"] for m in metas} >= {"naive"}
    assert {m["texture_mode"] for m in metas} >= {"nearest", "bilinear"}


def test_oracle_matches_reference(switched):
    out = _oracle(switched)
    assert np.array_equal(out["ids"], switched.ref_ids)
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), switched.ref_hit[..., 0].view(np.uint32))
    assert np.array_equal(out["frame_buffer"], switched.ref_frame_buffer)
    s = out["stats"]
    assert [s["primary_rays"], s["shadow_rays"], s["reflection_rays"], s["refraction_rays"]] == switched.ref_rays.tolist()


def test_the_switch_matters(switched):
    """The fixture's scene tells the settings apart: the default setting gives a different answer."""
    import oracle_ffi
    m = switched.meta
    default = oracle_ffi.render(switched.pack, m["bounces"], m["filter"])
    other = _oracle(switched)
    if m["texture_mode"] != "mipmap":
        assert not np.array_equal(default["frame_buffer"], switched.ref_frame_buffer)
    if m["traversal"] == "naive":
        assert other["stats"]["nodes_visited"] != default["stats"]["nodes_visited"]
        assert other["stats"]["triangles_tested"] != default["stats"]["triangles_tested"]


def test_bad_setting_is_refused():
    from b200rt import _ffi
    with pytest.raises(KeyError):
        _ffi.make_config(64, 64, 1, "tri", traversal="sideways")
    with pytest.raises(KeyError):
        _ffi.make_config(64, 64, 1, "tri", texture_mode="cubic")


@pytest.mark.gpu
def test_device_matches_reference_and_oracle(switched):
    from b200rt.device import Renderer
    m = switched.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], count_traversal=True,
                 traversal=m["traversal"], texture_mode=m["texture_mode"])
    r.upload_static(switched.pack)
    out = r.render(switched.pack, want_ids=True)
    r.close()
    assert np.array_equal(out["ids"], switched.ref_ids)
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), switched.ref_hit[..., 0].view(np.uint32))
    assert np.array_equal(out["frame_buffer"], switched.ref_frame_buffer)

    ref = _oracle(switched)
    assert np.array_equal(out["presented"], ref["presented"])
    for key in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays", "nodes_visited", "triangles_tested",
                "texels_fetched", "hits_accepted", "instances_entered"):
        assert out["stats"][key] == ref["stats"][key], key


@pytest.mark.gpu
def test_device_product_path_matches_reference(switched):
    """Without counters the shadow rays take the threaded any-hit traversal: same image."""
    from b200rt.device import Renderer
    m = switched.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], traversal=m["traversal"], texture_mode=m["texture_mode"])
    r.upload_static(switched.pack)
    out = r.render(switched.pack)
    r.close()
    assert np.array_equal(out["frame_buffer"], switched.ref_frame_buffer)


@pytest.mark.gpu
def test_device_refuses_bad_setting():
    from b200rt import _ffi
    import ctypes as C
    lib = _ffi.device_lib()
    for field in ("traversal_naive", "texture_sample_mode"):
        cfg = _ffi.make_config(64, 64, 1, "tri")
        setattr(cfg, field, 7)
        h = C.c_void_p()
        assert lib.b200rt_create(C.byref(cfg), C.byref(h)) != 0
        assert field in lib.b200rt_last_error(None).decode()
