"""The two restatements of Data/Shaders/fragment_fxaa.glsl held against each other (CPU).

oracle_present (C, pixel loop, 256-entry gamma table) is what fxaa_kernel is compared with on the GPU; tests/fxaa_uv.py is
a second restatement written separately (numpy, vectorised, pow() per texel, GL's sampler rule spelled out). Neither can
be pinned to the shader (no graphics stack: unavailable on this box). Both follow the GLSL in uv space; what is left
between them is numpy's float32 pow against libm's powf (90 of the 256 table entries differ in the last 1-3 ulp), which
moves a pixel by one level where a channel sits on a rounding edge. profiles/r02_fxaa_crosscheck.jsonl has the counts at
1080p and 4K (0 and 4 pixels), profiles/r02_fxaa_crosscheck_pixel_space_round1.jsonl what round 1's pixel-space form gave
(60 and 133 pixels, up to 21 levels)."""
import numpy as np

import oracle_ffi
from fxaa_uv import fxaa_uv


def test_uv_space_restatement_agrees_with_the_oracle(golden):
    fb = golden.ref_frame_buffer
    a, b = oracle_ffi.present(fb, True), fxaa_uv(fb, True)
    diff = np.abs(a.astype(np.int32) - b.astype(np.int32)).max(-1)
    assert (diff > 0).sum() <= 2 and diff.max() <= 1
    assert np.array_equal(oracle_ffi.present(fb, False), fxaa_uv(fb, False))     # fragment_identity.glsl: gamma only


def test_taps_that_sit_on_a_texel_border():
    """A frame whose edges make the blur direction hit exact fractions: both forms must pick the same texels."""
    rng = np.random.default_rng(3)
    fb = (rng.integers(0, 2, size=(64, 96)).astype(np.uint32) * 0xFFFFFF)
    fb[::2, ::3] = 0x808080
    a, b = oracle_ffi.present(fb, True), fxaa_uv(fb, True)
    diff = np.abs(a.astype(np.int32) - b.astype(np.int32)).max(-1)
    assert (diff > 1).sum() == 0 and (diff > 0).sum() <= 4
