"""The kernels' SOURCE run on the host (tests/emu: one fibre per CUDA thread) against the reference's stored frames.

No GPU is involved and nothing is proven about the GPU's memory model or speed — the `-m gpu` tests do that through the
real library. What this catches, in the container where the code is written: a wrong frame-stack walk, a deep-pixel queue
whose countdowns, parent links, overflow fall-backs or termination protocol are wrong (a hang shows as a timeout), a
threaded / wide layout that loses a triangle, an FXAA kernel that differs from the oracle. libm stands in for CUDA's
math functions — the functions the oracle calls — so the emulated frames must equal the reference's bit for bit.
"""
import os
import subprocess

import numpy as np
import pytest

import oracle_ffi
from conftest import REPO, Golden, psnr, unpack_rgb

EMU = os.path.join(REPO, "tests", "emu")


@pytest.fixture(scope="module")
def emulated():
    """Builds tests/emu/libb200rt_emu.so and loads it IN PLACE of libb200rt.so for this module (B200RT_LIB)."""
    subprocess.run(["make", "-C", EMU, "-s"], check=True)
    from b200rt import _ffi
    saved, saved_env = _ffi._device, os.environ.get("B200RT_LIB")
    _ffi._device = None
    os.environ["B200RT_LIB"] = os.path.join(EMU, "libb200rt_emu.so")
    from b200rt.device import Renderer
    yield Renderer
    _ffi._device = saved
    if saved_env is None:
        del os.environ["B200RT_LIB"]
    else:
        os.environ["B200RT_LIB"] = saved_env


def render(Renderer, g, **settings):
    m = g.meta
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], **settings)
    r.upload_static(g.pack)
    out = r.render(g.pack, want_ids=True)
    r.close()
    return out


def rays(stats):
    return [int(stats[k]) for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays")]


@pytest.mark.timeout(300)
@pytest.mark.parametrize("name,queue", [("dielectrics_b8_ewa", "off"), ("dynamic_b3_aniso_f45", "off"), ("dynamic_b3_tri_f0", "glass"), ("monkey_b1_aniso", "all")])
def test_emulated_kernels_give_the_references_frame(emulated, name, queue):
    g = Golden(name)
    out = render(emulated, g, deep_queue=queue)
    assert np.array_equal(out["ids"], g.ref_ids)
    assert np.array_equal(out["frame_buffer"], g.ref_frame_buffer)
    assert rays(out["stats"]) == [int(x) for x in g.ref_rays]
    assert np.array_equal(out["presented"], oracle_ffi.present(out["frame_buffer"], True))       # fxaa_kernel == oracle_present
    if queue != "off":
        assert out["stats"]["queue_nodes"] > 0 and out["stats"]["queue_fallbacks"] == 0


@pytest.mark.timeout(300)
def test_emulated_queue_with_exhausted_pools_falls_back_and_still_matches(emulated, monkeypatch):
    """B200RT_QUEUE_CAPACITY caps the three pools: most nodes and items find no slot and are walked depth-first instead."""
    monkeypatch.setenv("B200RT_QUEUE_CAPACITY", "2000")
    g = Golden("monkey_b1_aniso")
    out = render(emulated, g, deep_queue="all")
    assert out["stats"]["queue_fallbacks"] > 1000
    assert np.array_equal(out["frame_buffer"], g.ref_frame_buffer)
    assert rays(out["stats"]) == [int(x) for x in g.ref_rays]


@pytest.mark.timeout(300)
def test_emulated_fast_arithmetic_stays_inside_north_stars_tolerance(emulated):
    g = Golden("monkey_b1_aniso")
    out = render(emulated, g, arithmetic="fast")
    assert (out["ids"] == g.ref_ids).all(-1).mean() >= 0.999
    assert psnr(unpack_rgb(out["frame_buffer"]), unpack_rgb(g.ref_frame_buffer)) >= 50.0


@pytest.mark.timeout(600)
def test_emulated_auto_queue_times_both_walks_and_every_frame_matches(emulated):
    """deep_queue = AUTO: warm-up frames depth-first, then alternating timed frames, then the faster walk — same bits throughout."""
    g = Golden("monkey_b1_aniso")
    m = g.meta
    r = emulated(m["width"], m["height"], m["bounces"], m["filter"], deep_queue="auto")
    r.upload_static(g.pack)
    queued = []
    for _ in range(10):
        out = r.render(g.pack, want_presented=False)
        assert np.array_equal(out["frame_buffer"], g.ref_frame_buffer)
        assert rays(out["stats"]) == [int(x) for x in g.ref_rays]
        queued.append(out["stats"]["queue_nodes"] > 0)
    r.close()
    assert queued[:8] == [False, False, False, True, False, True, False, True]     # AutoQueue::WARM = 2, PAIRS = 3
    assert queued[8] == queued[9]                                                   # decided: no more alternation


@pytest.mark.timeout(300)
def test_emulated_no_alloc_entry_point_gives_the_same_frame_twice(emulated):
    """b200rt_render_resident_noalloc on the host build (tests/test_zz_no_alloc.py is the same test on the GPU)."""
    import ctypes as C
    from b200rt import _ffi
    g = Golden("monkey_b1_aniso")
    m = g.meta
    lib = _ffi.device_lib()
    r = emulated(m["width"], m["height"], m["bounces"], m["filter"])
    r.upload_static(g.pack)
    expected = r.render(g.pack, want_presented=False)["frame_buffer"]
    for _ in range(2):
        ptr, handle = C.c_uint64(), _ffi.PeerHandle()
        r._check(lib.b200rt_peer_alloc(r._h, expected.nbytes, C.byref(ptr), C.byref(handle)))
        r.render_resident_noalloc(ptr.value)
        got = np.zeros_like(expected)
        r._check(lib.b200rt_peer_read(r._h, ptr, got.ctypes.data, got.nbytes))
        r._check(lib.b200rt_peer_free(r._h, ptr))
        assert np.array_equal(got, expected)
    r.close()


@pytest.mark.timeout(300)
def test_emulated_queue_that_gave_a_frame_up_is_rendered_again_depth_first(emulated, monkeypatch):
    """The queue kernel's watchdog (FrameCounters::q_abort) marks a frame as incomplete: b200rt_render_frame renders it again
    depth-first and the context stays depth-first. B200RT_TEST_QUEUE_STALL marks the first queued frame that way."""
    monkeypatch.setenv("B200RT_TEST_QUEUE_STALL", "1")
    g = Golden("monkey_b1_aniso")
    m = g.meta
    r = emulated(m["width"], m["height"], m["bounces"], m["filter"], deep_queue="all")
    r.upload_static(g.pack)
    first = r.render(g.pack, want_presented=False)
    second = r.render(g.pack, want_presented=False)
    r.close()
    for out in (first, second):
        assert np.array_equal(out["frame_buffer"], g.ref_frame_buffer)
        assert out["stats"]["queue_nodes"] == 0            # the frame that counts was walked depth-first
        assert rays(out["stats"]) == [int(x) for x in g.ref_rays]


@pytest.mark.timeout(600)
@pytest.mark.parametrize("keep", ["0", "1", "2"])
def test_emulated_queue_gives_the_same_frame_whichever_child_stays_in_the_thread(emulated, monkeypatch, keep):
    """An exported node's thread goes on with one child and queues the other (B200RT_QUEUE_KEEP_REFRACTION: the reflection
    ray, the refraction ray, or the one that stays inside the medium — the shipped setting): the same bits every time."""
    monkeypatch.setenv("B200RT_QUEUE_KEEP_REFRACTION", keep)
    g = Golden("dielectrics_b8_ewa")
    out = render(emulated, g, deep_queue="all")
    assert np.array_equal(out["frame_buffer"], g.ref_frame_buffer)
    assert rays(out["stats"]) == [int(x) for x in g.ref_rays]
    assert out["stats"]["queue_nodes"] > 0 and out["stats"]["queue_bounce_items"] > 0 and out["stats"]["queue_fallbacks"] == 0
