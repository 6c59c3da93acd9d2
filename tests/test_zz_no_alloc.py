"""b200rt_render_resident_noalloc — the entry point that allocates nothing (include/b200rt.h) — against the entry point
that existed before it, b200rt_render_frame_device, through the C ABI on the GPU.

Collected last on purpose. The scratch of the new entry point is one static device
array shared by every call of the process, which keeps the previous call's contents: the test calls it twice in a row,
each time into a fresh zeroed image, after the array may already have been used by anything that ran before.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import Golden, unpack_rgb

pytestmark = pytest.mark.gpu


def device_image(renderer, lib, nbytes):
    from b200rt import _ffi
    ptr, handle = C.c_uint64(), _ffi.PeerHandle()
    renderer._check(lib.b200rt_peer_alloc(renderer._h, nbytes, C.byref(ptr), C.byref(handle)))     # zeroed device memory (the TEST may allocate)
    return ptr.value


def read_image(renderer, lib, ptr, height, width):
    image = np.zeros((height, width), dtype=np.uint32)
    renderer._check(lib.b200rt_peer_read(renderer._h, C.c_uint64(ptr), image.ctypes.data, image.nbytes))  # waits for the context's stream
    return image


@pytest.mark.parametrize("name", ["dynamic_b3_aniso_f45", "dielectrics_b8_ewa"])
def test_no_alloc_entry_point_renders_the_same_frame_twice_in_a_row(name):
    from b200rt import _ffi
    from b200rt.device import Renderer
    g = Golden(name)
    m = g.meta
    lib = _ffi.device_lib()
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
    r.upload_static(g.pack)
    images = [device_image(r, lib, m["width"] * m["height"] * 4) for _ in range(3)]
    try:
        stats = r.render_device(g.pack, images[0])                      # the existing entry point: b200rt_render_frame_device
        expected = read_image(r, lib, images[0], m["height"], m["width"])
        assert stats["primary_rays"] == m["width"] * m["height"]
        diff = np.abs(unpack_rgb(expected) - unpack_rgb(g.ref_frame_buffer))
        assert (diff.max(-1) > 0).mean() <= 2e-3 and diff.max() <= 8    # ... which is the reference's frame (test_device_parity's bars)
        for call in (1, 2):
            r.render_resident_noalloc(images[call])
            got = read_image(r, lib, images[call], m["height"], m["width"])
            # between two entry points of one library the tests' tolerance is "the same bits" (test_device_parity: pinned / pageable,
            # pipelined / synchronous)
            assert np.array_equal(got, expected), "call %d: %d pixels differ" % (call, int((got != expected).sum()))
        # the context is as it was: the existing entry point still gives the same frame and counters
        again = r.render_device(g.pack, images[0])
        assert np.array_equal(read_image(r, lib, images[0], m["height"], m["width"]), expected)
        assert [again[k] for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays")] == [stats[k] for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays")]
    finally:
        for ptr in images:
            lib.b200rt_peer_free(r._h, C.c_uint64(ptr))
        r.close()


def test_no_alloc_entry_point_refuses_what_its_scratch_cannot_hold():
    from b200rt.device import DeviceError, Renderer
    r = Renderer(4128, 4128, 1, "tri")                                   # 129 x 129 = 16 641 tiles > B200RT_NOALLOC_MAX_TILES
    try:
        with pytest.raises(DeviceError, match="B200RT_NOALLOC_MAX_TILES"):
            r.render_resident_noalloc(0x1000)
    finally:
        r.close()
    r = Renderer(64, 64, 1, "tri")
    try:
        with pytest.raises(DeviceError, match="no frame is resident"):
            r.render_resident_noalloc(0x1000)
        with pytest.raises(DeviceError, match="null device frame buffer"):
            r.render_resident_noalloc(0)
    finally:
        r.close()
