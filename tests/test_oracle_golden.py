"""CPU: pins the oracle (oracle/whitted_oracle.c) against the golden outputs of the REAL reference.

The fixtures under tests/golden/ were produced by tests/golden/make_golden.py from the reference's
own sources (oracle/ref_recipe), run per ray through its Raytracer::bounce. The oracle is a
restatement of the same arithmetic in C; here it has to reproduce the reference bit for bit:
frame buffer, primary-hit ids, hit records and the four ray counters. No GPU involved.
"""
import numpy as np

import oracle_ffi
from conftest import unpack_rgb


def _oracle(golden, **kw):
    m = golden.meta
    return oracle_ffi.render(golden.pack, m["bounces"], m["filter"], **kw)


def test_oracle_frame_buffer_is_the_references(golden):
    out = _oracle(golden)
    assert np.array_equal(out["frame_buffer"], golden.ref_frame_buffer)


def test_oracle_primary_hits_are_the_references(golden):
    out = _oracle(golden)
    assert np.array_equal(out["ids"], golden.ref_ids)
    # distance and material id carry no transcendental: bit-exact
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), golden.ref_hit[..., 0].view(np.uint32))
    assert np.array_equal(out["hit"][..., 1], golden.ref_hit[..., 1])
    np.testing.assert_allclose(out["hit"][..., 2:], golden.ref_hit[..., 2:], rtol=0, atol=2e-6)


def test_oracle_ray_counters_are_the_references(golden):
    s = _oracle(golden, want_ids=False)["stats"]
    got = [s["primary_rays"], s["shadow_rays"], s["reflection_rays"], s["refraction_rays"]]
    assert got == golden.ref_rays.tolist()


def test_oracle_is_thread_count_independent(golden):
    a = _oracle(golden, threads=1, want_ids=False)
    b = _oracle(golden, threads=5, want_ids=False)
    assert np.array_equal(a["frame_buffer"], b["frame_buffer"])
    assert a["stats"]["nodes_visited"] == b["stats"]["nodes_visited"]


def test_oracle_tile_partition_reassembles(golden):
    """tile t -> shard t % N. Disjoint shards OR together to the full frame."""
    full = _oracle(golden, want_ids=False)
    for n in (2, 3, 8):
        merged = np.zeros_like(full["frame_buffer"])
        rays = 0
        for phase in range(n):
            part = _oracle(golden, want_ids=False, tile_stride=n, tile_phase=phase)
            assert not (merged.astype(bool) & part["frame_buffer"].astype(bool)).any()
            merged |= part["frame_buffer"]
            rays += part["stats"]["primary_rays"]
        assert np.array_equal(merged, full["frame_buffer"])
        assert rays == full["stats"]["primary_rays"]


def test_present_identity_is_gamma_of_the_8bit_image(golden):
    """fragment_identity.glsl: pow(texel, 1/2.2) written to an 8-bit target."""
    fb = golden.ref_frame_buffer
    table = oracle_ffi.gamma_table()
    expect = np.rint(np.clip(table[unpack_rgb(fb)], 0, 1) * 255.0).astype(np.uint8)
    assert np.array_equal(oracle_ffi.present(fb, enable_fxaa=False), expect)


def test_present_fxaa_flat_regions_and_wrap():
    """FXAA leaves a constant image alone; borders wrap (GL_REPEAT, Window.cpp:48-50)."""
    flat = np.full((16, 24), 0x00804020, dtype=np.uint32)
    out = oracle_ffi.present(flat, True)
    assert (out == oracle_ffi.present(flat, False)).all()
    # a vertical edge on the left border must see the right border through the wrap
    img = np.zeros((16, 24), dtype=np.uint32)
    img[:, 0] = 0x00FFFFFF
    a = oracle_ffi.present(img, True)
    rolled = oracle_ffi.present(np.roll(img, 5, axis=1), True)
    assert np.array_equal(np.roll(a, 5, axis=1), rolled)
