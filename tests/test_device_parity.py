"""GPU parity tests proper: the CUDA path, called through the C ABI (libb200rt.so), against
(1) the CPU oracle on the same inputs and (2) the golden outputs of the real reference renderer.

Bars (BASELINE.json north_star): primary-hit (instance, triangle) ids equal on >= 99.9 % of pixels,
final image PSNR >= 50 dB. What is actually observed is far tighter and asserted as such:
the CUDA kernels use the same operation order as the oracle (fused multiply-add only where the
reference calls madd/msub), so integer outputs are compared EXACTLY and the only tolerated
differences come from transcendentals (acosf/atan2f/expf/log2f of CUDA vs glibc, a few ulp).
"""
import numpy as np
import pytest

from conftest import Golden, psnr, unpack_rgb

pytestmark = pytest.mark.gpu

ID_AGREEMENT = 0.999   # north_star
PSNR_DB = 50.0         # north_star


def _render(golden, **kw):
    from b200rt.device import Renderer
    m = golden.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], count_traversal=kw.pop("count_traversal", False))
    r.upload_static(golden.pack)
    out = r.render(golden.pack, **kw)
    r.close()
    return out


def test_primary_hits_match_reference(golden):
    out = _render(golden, want_ids=True)
    same = (out["ids"] == golden.ref_ids).all(-1)
    assert same.mean() >= ID_AGREEMENT, f"id agreement {same.mean():.5f}"
    # in practice: bit-identical ids
    assert same.all(), f"{(~same).sum()} pixels with a different primary hit"
    # hit distance is free of transcendentals: bit-exact. material id too.
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), golden.ref_hit[..., 0].view(np.uint32))
    assert np.array_equal(out["hit"][..., 1], golden.ref_hit[..., 1])
    # u, v: exact for triangles and planes; spheres go through atan2f / acosf (libm vs libmvec vs CUDA)
    np.testing.assert_allclose(out["hit"][..., 2:], golden.ref_hit[..., 2:], rtol=0, atol=2e-6)


def test_image_matches_reference(golden):
    out = _render(golden)
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(golden.ref_frame_buffer)
    diff = np.abs(a - b)
    p = psnr(a, b)
    assert p >= PSNR_DB, f"PSNR {p:.2f} dB"
    # observed: a handful of pixels off by one level where a transcendental's last ulp decides a rounding
    assert (diff.max(-1) > 0).mean() <= 2e-3, f"{(diff.max(-1) > 0).sum()} pixels differ"
    assert diff.max() <= 8, f"max channel error {diff.max()}"


def test_image_and_counters_match_oracle(golden):
    import oracle_ffi
    m = golden.meta
    ref = oracle_ffi.render(golden.pack, m["bounces"], m["filter"])
    out = _render(golden, count_traversal=True)

    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    assert psnr(a, b) >= PSNR_DB
    assert (np.abs(a - b).max(-1) > 0).mean() <= 2e-3

    # ray counts: the reference's PerformanceStats, one count per traced ray
    for key in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays"):
        assert abs(out["stats"][key] - ref["stats"][key]) <= max(2, ref["stats"][key] // 10000), key
    # traversal work: same algorithm, same order => same node / triangle counts (texel counts may
    # differ where log2f picks a neighbouring mip level)
    for key in ("nodes_visited", "triangles_tested"):
        assert abs(out["stats"][key] - ref["stats"][key]) <= max(8, ref["stats"][key] // 5000), key


def test_fxaa_matches_oracle(golden):
    import oracle_ffi
    out = _render(golden)
    expect = oracle_ffi.present(out["frame_buffer"], enable_fxaa=True)
    # same table-driven arithmetic on both sides: exact
    assert np.array_equal(out["presented"], expect)


def test_identity_present_matches_oracle(golden):
    import oracle_ffi
    from b200rt.device import Renderer
    m = golden.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], enable_fxaa=False)
    r.upload_static(golden.pack)
    out = r.render(golden.pack)
    r.close()
    assert np.array_equal(out["presented"], oracle_ffi.present(out["frame_buffer"], enable_fxaa=False))


def test_tile_partition_reassembles_the_frame(golden):
    """Multi-GPU partition on one GPU: N contexts with interleaved tiles == one context (bit for bit)."""
    from b200rt.device import Renderer
    m = golden.meta
    full = _render(golden)["frame_buffer"]
    for n in (2, 4, 8):
        merged = np.zeros_like(full)
        for phase in range(n):
            r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], tile_stride=n, tile_phase=phase)
            r.upload_static(golden.pack)
            part = r.render(golden.pack, want_presented=False)["frame_buffer"]
            r.close()
            assert (merged[part != 0] == 0).all(), f"stride {n}: phase {phase} wrote pixels another phase had written"
            merged |= part
        assert np.array_equal(merged, full), f"stride {n}"


def test_page_locked_outputs_equal_pageable(golden):
    """b200rt_host_alloc buffers are filled by DMA directly; the images must be the same bytes."""
    from b200rt.device import Renderer
    m = golden.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
    r.upload_static(golden.pack)
    pageable = r.render(golden.pack)
    pinned = r.alloc_outputs()
    r.render(golden.pack, out=pinned)
    assert np.array_equal(pinned["frame_buffer"], pageable["frame_buffer"])
    assert np.array_equal(pinned["presented"], pageable["presented"])
    del pinned
    r.close()


def test_pipelined_frames_equal_synchronous_frames(golden):
    """b200rt_submit_frame / b200rt_wait_frame with two frames in flight deliver the same images and counters."""
    from b200rt.device import DeviceError, Renderer
    m = golden.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
    r.upload_static(golden.pack)
    sync = r.render(golden.pack)
    with pytest.raises(DeviceError, match="page-locked"):
        r.submit(golden.pack, {"frame_buffer": np.zeros_like(sync["frame_buffer"]), "presented": np.zeros_like(sync["presented"])})
    outs = [r.alloc_outputs(), r.alloc_outputs()]
    r.submit(golden.pack, outs[0])
    r.submit(golden.pack, outs[1])
    with pytest.raises(DeviceError, match="in flight"):
        r.submit(golden.pack, outs[0])
    for k in range(2):
        stats = r.wait()
        assert np.array_equal(outs[k]["frame_buffer"], sync["frame_buffer"])
        assert np.array_equal(outs[k]["presented"], sync["presented"])
        for key in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays"):
            assert stats[key] == sync["stats"][key]
    with pytest.raises(DeviceError, match="no frame in flight"):
        r.wait()
    del outs
    r.close()


def test_pipelined_animation_frames_match_the_reference():
    """Config 3 (dynamic scene): frames 0 and 45 of the animation — different instance matrices and a
    different per-frame TopLevelBVH — in flight together through b200rt_submit_frame; each must come
    out as the real reference rendered it (tests/golden/dynamic_b3_aniso_f0 / _f45)."""
    from conftest import Golden
    from b200rt.device import Renderer
    f0, f45 = Golden("dynamic_b3_aniso_f0"), Golden("dynamic_b3_aniso_f45")
    assert np.array_equal(f0.pack.a["s_nodes"], f45.pack.a["s_nodes"])            # same static scene ...
    assert not np.array_equal(f0.pack.a["f_instances"], f45.pack.a["f_instances"])  # ... different frame data
    m = f0.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
    r.upload_static(f0.pack)
    outs = [r.alloc_outputs(), r.alloc_outputs()]
    for round_ in range(3):
        r.submit(f0.pack, outs[0])
        r.submit(f45.pack, outs[1])
        r.wait()
        r.wait()
        for out, g in ((outs[0], f0), (outs[1], f45)):
            diff = np.abs(unpack_rgb(out["frame_buffer"]) - unpack_rgb(g.ref_frame_buffer))
            assert (diff.max(-1) > 0).mean() <= 2e-3 and diff.max() <= 8
    del outs
    r.close()


@pytest.mark.gpu
def test_a_frame_the_queue_gave_up_is_rendered_again_depth_first(monkeypatch):
    """FrameCounters::q_abort (the queue kernel's watchdog) marks a frame as incomplete: b200rt_render_frame renders it
    again depth-first, the asynchronous entry points report it, and the context stays depth-first afterwards."""
    from b200rt.device import DeviceError, Renderer
    monkeypatch.setenv("B200RT_TEST_QUEUE_STALL", "1")      # the first queued frame of a context is marked as given up
    g = Golden("dielectrics_b8_ewa")
    m = g.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], deep_queue="all")
    r.upload_static(g.pack)
    first = r.render(g.pack, want_presented=False)
    second = r.render(g.pack, want_presented=False)
    r.close()
    assert np.array_equal(first["frame_buffer"], second["frame_buffer"])
    assert first["stats"]["queue_nodes"] == 0 and second["stats"]["queue_nodes"] == 0
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], deep_queue="all")
    r.upload_static(g.pack)
    outs = r.alloc_outputs()
    r.submit(g.pack, outs)
    with pytest.raises(DeviceError, match="stalled"):
        r.wait()
    r.submit(g.pack, outs)                                   # again, as the message says
    assert r.wait()["queue_nodes"] == 0
    assert np.array_equal(outs["frame_buffer"], first["frame_buffer"])
    r.close()
