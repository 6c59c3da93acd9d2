"""GPU, BASELINE.json's full sizes: configs[1] (Sponza 1920x1080, trilinear, 3 bounces) and configs[4]
(Sibenik 3840x2160, spot + point light) on the procedural stand-ins (the OBJs are absent from the
reference checkout). At these sizes the CUDA path is compared with the multi-threaded CPU oracle on
the same arrays, and through size-independent properties: tile shards of 2/4/8 "GPUs" OR together
to the single-GPU frame, FXAA of the device equals the oracle's FXAA of the same frame, the ray
counters add up.
"""
import numpy as np
import pytest

from conftest import psnr, unpack_rgb

pytestmark = pytest.mark.gpu


def _scene(kind):
    from b200rt import standin
    from b200rt.host import HostScene
    if kind == "sponza":
        scene = HostScene(standin.sponza_standin(width=1920, height=1080))
        cfg = (1920, 1080, 3, "tri")
    else:
        scene = HostScene(standin.sibenik_standin(width=3840, height=2160))
        cfg = (3840, 2160, 3, "aniso")
    scene.update(0.0, 0)
    return scene.pack(), cfg


@pytest.fixture(scope="module", params=["sponza", "sibenik"])
def full(request):
    import oracle_ffi
    from b200rt.device import Renderer
    pack, (w, h, bounces, filt) = _scene(request.param)
    r = Renderer(w, h, bounces, filt, count_traversal=True)
    r.upload_static(pack)
    out = r.render(pack, want_ids=True)
    r.close()
    ref = oracle_ffi.render(pack, bounces, filt)
    return dict(pack=pack, cfg=(w, h, bounces, filt), out=out, ref=ref)


def test_full_size_primary_hits_equal_oracle(full):
    out, ref = full["out"], full["ref"]
    same = (out["ids"] == ref["ids"]).all(-1)
    assert same.mean() >= 0.999
    assert same.all(), f"{(~same).sum()} pixels with a different primary hit"
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), ref["hit"][..., 0].view(np.uint32))


def test_full_size_image_equals_oracle(full):
    a, b = unpack_rgb(full["out"]["frame_buffer"]), unpack_rgb(full["ref"]["frame_buffer"])
    diff = np.abs(a - b)
    assert psnr(a, b) >= 50.0
    assert (diff.max(-1) > 0).mean() <= 2e-3, f"{(diff.max(-1) > 0).sum()} pixels differ"
    assert diff.max() <= 8


def test_full_size_counters(full):
    w, h, _, _ = full["cfg"]
    s, r = full["out"]["stats"], full["ref"]["stats"]
    assert s["primary_rays"] == w * h == r["primary_rays"]
    for key in ("shadow_rays", "reflection_rays", "refraction_rays", "nodes_visited", "triangles_tested"):
        assert abs(s[key] - r[key]) <= max(8, r[key] // 5000), key


def test_full_size_fxaa_equals_oracle(full):
    import oracle_ffi
    assert np.array_equal(full["out"]["presented"], oracle_ffi.present(full["out"]["frame_buffer"], True))


def test_full_size_tile_shards_reassemble(full):
    from b200rt.device import Renderer
    w, h, bounces, filt = full["cfg"]
    whole = full["out"]["frame_buffer"]
    for n in (2, 8):
        merged = np.zeros_like(whole)
        rays = 0
        for phase in range(n):
            r = Renderer(w, h, bounces, filt, tile_stride=n, tile_phase=phase)
            r.upload_static(full["pack"])
            part = r.render(full["pack"], want_presented=False)
            r.close()
            merged |= part["frame_buffer"]
            rays += part["stats"]["primary_rays"]
        assert np.array_equal(merged, whole), f"stride {n}"
        assert rays == w * h
