#!/usr/bin/env python3
"""Where the end-to-end frame time goes (development tool): the pipelined loop of bench.py with and
without the per-frame host work, against the device time the library reports for the same frames."""
import ctypes as C
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
sys.path.insert(0, REPO)
import bench  # noqa: E402
from b200rt import _ffi  # noqa: E402
from b200rt.device import Renderer  # noqa: E402


def main():
    frames = int(os.environ.get("FRAMES", "200"))
    _, scene = bench.build_scene()
    pack = scene.pack()
    r = Renderer(bench.WIDTH, bench.HEIGHT, bench.BOUNCES, bench.FILTER)
    r.upload_static(pack)
    outs = [r.alloc_outputs(), r.alloc_outputs()]
    res = {}

    def loop(n, prep):
        device = 0.0
        parts = [0.0, 0.0]
        for i in range(n):
            if prep:
                scene.update(0.0, 0)
                p = pack.with_frame(scene.frame_view())
            else:
                p = pack
            r.submit(p, outs[i & 1])
            if i >= 1:
                st = r.wait(); device += st["device_ms_total"]; parts[0] += st["device_ms_render"]; parts[1] += st["device_ms_fxaa"]
        st = r.wait(); device += st["device_ms_total"]; parts[0] += st["device_ms_render"]; parts[1] += st["device_ms_fxaa"]
        return device / n, parts[0] / n, parts[1] / n

    for name, prep in (("with_host_prep", True), ("submit_wait_only", False)):
        loop(5, prep)
        t0 = time.perf_counter()
        device, render, fxaa = loop(frames, prep)
        res[name] = {"wall_ms_per_frame": (time.perf_counter() - t0) * 1e3 / frames, "device_begin_to_copied_ms": device, "device_ms_render": render, "device_ms_fxaa_and_schedule": fxaa}

    # raw ctypes loop: one prebuilt frame struct, no Python objects created per frame
    lib = r._lib
    frame = pack.frame_view()
    views = [_ffi.Outputs(o["frame_buffer"].ctypes.data, o["presented"].ctypes.data, None, None, None, None) for o in outs]
    stats = _ffi.Stats()
    for n in (5, frames):
        t0 = time.perf_counter()
        for i in range(n):
            assert lib.b200rt_submit_frame(r._h, C.byref(frame), C.byref(views[i & 1])) == 0
            if i >= 1:
                assert lib.b200rt_wait_frame(r._h, C.byref(stats)) == 0
        assert lib.b200rt_wait_frame(r._h, C.byref(stats)) == 0
        wall = (time.perf_counter() - t0) * 1e3 / n
    res["raw_ctypes"] = {"wall_ms_per_frame": wall}
    ms_render, ms_fxaa = r.time_resident(20, 0)
    res["resident_warm_l2"] = {"ms_render": ms_render, "ms_fxaa": ms_fxaa}
    t0 = time.perf_counter()
    for _ in range(200):
        scene.update(0.0, 0)
        pack.with_frame(scene.frame_view())
    res["host_prep_ms_per_frame"] = (time.perf_counter() - t0) * 1e3 / 200
    print(json.dumps(res))


if __name__ == "__main__":
    main()
