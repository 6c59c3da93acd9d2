#!/usr/bin/env python3
"""Where do the two restatements of fragment_fxaa.glsl disagree? (development tool; CPU only)

oracle_present / fxaa_kernel step their taps in pixel space; tests/fxaa_uv.py follows the shader in uv space. Compares
both on the golden fixtures (480x270) and on full-size frames rendered by the C oracle (Sponza stand-in 1080p, Sibenik
stand-in 4K); one JSON line per frame, written to stdout (profiles/r02_fxaa_crosscheck.jsonl is a copy)."""
import json, os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python")); sys.path.insert(0, os.path.join(REPO, "tests")); sys.path.insert(0, os.path.join(REPO, "oracle"))
import numpy as np
import conftest, oracle_ffi
from fxaa_uv import fxaa_uv


def report(name, fb):
    a = oracle_ffi.present(fb, True)
    b = fxaa_uv(fb, True)
    diff = np.abs(a.astype(np.int32) - b.astype(np.int32)).max(-1)
    ident = bool(np.array_equal(oracle_ffi.present(fb, False), fxaa_uv(fb, False)))
    h, w = fb.shape
    print(json.dumps({"frame": name, "size": [w, h], "pixels": int(w * h), "pixels_that_differ": int((diff > 0).sum()), "max_channel_difference": int(diff.max()),
                      "identity_pass_equal": ident, "where": [[int(x), int(y)] for y, x in zip(*np.nonzero(diff))][:8]}), flush=True)


for name in conftest.golden_names():
    report(name, conftest.Golden(name).ref_frame_buffer)
if "--full" in sys.argv:
    from b200rt import standin
    from b200rt.host import HostScene
    for label, fn, w, h, bounces, filt in (("sponza_standin_1080p", standin.sponza_standin, 1920, 1080, 3, "tri"), ("sibenik_standin_4k", standin.sibenik_standin, 3840, 2160, 3, "aniso")):
        scene = HostScene(fn(width=w, height=h)); scene.update(0.0, 0)
        out = oracle_ffi.render(scene.pack(), bounces, filt, want_ids=False)
        report(label, out["frame_buffer"])
