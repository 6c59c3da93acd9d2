"""TEST INFRASTRUCTURE — a second, independently written restatement of Data/Shaders/fragment_fxaa.glsl.

The shader works in normalised texture coordinates: in_uv = ((px + .5) / W, (py + .5) / H) (vertex.glsl:7-11), offsets
multiplied by inv_screen_size = (1 / W, 1 / H) (Window.cpp:57), and the NEAREST / REPEAT sampler of Window.cpp:48-50
picks texel floor(u * W) mod W. This module follows the GLSL line by line in float32 numpy — written separately from
oracle_present (oracle/whitted_oracle.c) and fxaa_kernel (csrc/kernels.cuh), vectorised over the image instead of
pixel by pixel, pow() instead of a 256-entry table — so that the restatements can be held against each other. Round 1's
oracle and kernel stepped their taps in PIXEL space (cx + dir * k, then floor): a tap within rounding distance of a
texel border then picked a different texel than the uv arithmetic does — 60 of 2 073 600 pixels at 1080p, 133 of
8 294 400 at 4K (profiles/r02_fxaa_crosscheck.jsonl) — and both were rewritten in uv space. The shader itself cannot
run here (no graphics stack: unavailable on this box), so this is a cross-check between restatements, not a pin.
"""
import numpy as np

F = np.float32


def _texture_nearest_repeat(image_rgb, u, v):
    """texture2D on a NEAREST, GL_REPEAT sampler (OpenGL 4.2 spec 3.9.11): texel floor(u * W) mod W."""
    h, w, _ = image_rgb.shape
    i = np.mod(np.floor(u * F(w)).astype(np.int64), w)
    j = np.mod(np.floor(v * F(h)).astype(np.int64), h)
    return image_rgb[j, i], i, j


def fxaa_uv(frame_buffer, enable_fxaa=True, return_taps=False):
    """frame_buffer: uint32 [H][W] 0x00RRGGBB (Window::plot). Returns uint8 [H][W][3] as the shader would write it."""
    fb = np.asarray(frame_buffer, dtype=np.uint32)
    h, w = fb.shape
    # RGBA8 UNORM texture -> float: c / 255; get_screen(): pow(rgb, 1 / 2.2)
    texel = np.stack([(fb >> 16) & 255, (fb >> 8) & 255, fb & 255], -1).astype(F) / F(255.0)
    screen = np.power(texel, F(1.0 / 2.2), dtype=F)

    inv = np.array([F(1.0) / F(w), F(1.0) / F(h)], dtype=F)                       # uniform inv_screen_size
    px, py = np.meshgrid(np.arange(w, dtype=F), np.arange(h, dtype=F))
    in_u = (px + F(0.5)) / F(w)
    in_v = (py + F(0.5)) / F(h)

    def get_screen(u, v):
        return _texture_nearest_repeat(screen, u.astype(F), v.astype(F))

    if not enable_fxaa:
        out = get_screen(in_u, in_v)[0]
    else:
        luma_w = np.array([0.299, 0.587, 0.114], dtype=F)

        def luma(c):
            return (luma_w[0] * c[..., 0] + luma_w[1] * c[..., 1]) + luma_w[2] * c[..., 2]

        # vec2(-inv.x, -inv_screen_size): the second argument is a vec2, its .x fills the second component (:25-28)
        tl = luma(get_screen(in_u - inv[0], in_v - inv[0])[0])
        tr = luma(get_screen(in_u + inv[0], in_v - inv[0])[0])
        bl = luma(get_screen(in_u - inv[0], in_v + inv[0])[0])
        br = luma(get_screen(in_u + inv[0], in_v + inv[0])[0])
        m = luma(get_screen(in_u, in_v)[0])

        luma_min = np.minimum(np.minimum(np.minimum(tl, tr), np.minimum(bl, br)), m)
        luma_max = np.maximum(np.maximum(np.maximum(tl, tr), np.maximum(bl, br)), m)

        dir_x = (bl + br) - (tl + tr)
        dir_y = (tl + bl) - (tr + br)
        reduce = np.maximum(F(1.0 / 128.0), (tl + tr + bl + br) * F(0.25) * F(1.0 / 8.0))
        adjust = F(1.0) / (np.minimum(np.abs(dir_x), np.abs(dir_y)) + reduce)
        dir_x = np.clip(dir_x * adjust, F(-8.0), F(8.0)) * inv[0]
        dir_y = np.clip(dir_y * adjust, F(-8.0), F(8.0)) * inv[1]

        taps = []
        samples = []
        for k in (F(1.0) / F(3.0) - F(0.5), F(2.0) / F(3.0) - F(0.5), F(0.0) / F(3.0) - F(0.5), F(3.0) / F(3.0) - F(0.5)):   # float constants, as in the GLSL
            c, i, j = get_screen(in_u + dir_x * k, in_v + dir_y * k)
            samples.append(c)
            taps.append((i, j))
        result_a = F(0.5) * (samples[0] + samples[1])
        result_b = F(0.5) * (samples[2] + samples[3])
        result = F(0.5) * (result_a + result_b)
        luma_result = luma(result)
        outside = (luma_result < luma_min) | (luma_result > luma_max)
        out = np.where(outside[..., None], result_a, result)
    rgb8 = np.rint(np.clip(out, 0.0, 1.0) * F(255.0)).astype(np.uint8)           # UNORM8 colour attachment: round to nearest
    if return_taps and enable_fxaa:
        return rgb8, taps, (dir_x * F(w), dir_y * F(h))
    return rgb8
