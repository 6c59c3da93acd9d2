#!/usr/bin/env python3
"""Config 3 (BASELINE.json configs[2]): the shipped SCENE_DYNAMIC animated for 240 frames at dt = 1/60,
1920x1080, 3 bounces, anisotropic filter — the per-frame TopLevelBVH rebuild on the host
(Scene.cpp:139-171) pipelined against the GPU (b200rt_submit_frame / b200rt_wait_frame), next to
the reference's own binary on the same scene file. Needs the reference's assets (oracle/_ref/Data,
staged by oracle/ref_recipe/stage_assets.py; they travel to the GPU box). One JSON line.
"""
import json
import os
import subprocess
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))

from b200rt import scenes  # noqa: E402
from b200rt.device import Renderer  # noqa: E402
from b200rt.host import HostScene  # noqa: E402

W, H, FRAMES, DT = 1920, 1080, 240, 1.0 / 60.0
DATA = os.path.join(REPO, "oracle", "_ref", "Data")


def main():
    with tempfile.TemporaryDirectory() as tmp:
        scene_file = scenes.dynamic(DATA, os.path.join(tmp, "dynamic.scene"), W, H)
        scene = HostScene(scene_file)
        scene.update(DT, 1)
        pack = scene.pack()
        r = Renderer(W, H, 3, "aniso")
        r.upload_static(pack)
        outs = [r.alloc_outputs(), r.alloc_outputs()]
        rays = 0
        host_ms = 0.0
        for warm in (True, False):
            t0 = time.perf_counter()
            for i in range(FRAMES if not warm else 10):
                h0 = time.perf_counter()
                scene.update(DT, 1)                                   # animation + TLAS rebuild
                frame = pack.with_frame(scene.frame_view())
                host_ms += (time.perf_counter() - h0) * 1e3 if not warm else 0.0
                r.submit(frame, outs[i & 1])
                if i >= 1:
                    s = r.wait()
                    rays += (s["primary_rays"] + s["shadow_rays"] + s["reflection_rays"] + s["refraction_rays"]) if not warm else 0
            s = r.wait()
            rays += (s["primary_rays"] + s["shadow_rays"] + s["reflection_rays"] + s["refraction_rays"]) if not warm else 0
            elapsed = time.perf_counter() - t0
        line = {"workload": "configs[2]: SCENE_DYNAMIC, 6 instances / 5 BLAS, per-frame TLAS rebuild, 1920x1080, 3 bounces, anisotropic, %d frames at dt=1/60" % FRAMES,
                "gpu": {"frames_per_s": FRAMES / elapsed, "ms_per_frame": elapsed * 1e3 / FRAMES, "mrays_per_s": rays / elapsed / 1e6,
                        "host_update_ms_per_frame": host_ms / FRAMES, "api": "Scene::update + b200rt_submit_frame / b200rt_wait_frame"}}
        binary = os.path.join(REPO, "oracle", "_ref", "bin", "ref_l8_b3_aniso_dyn")
        if os.path.exists(binary):
            out = subprocess.run([binary, scene_file, "--frames", "60", "--dt", repr(DT), "--width", str(W), "--height", str(H)], capture_output=True, text=True, check=True).stdout
            info = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
            ms = sorted(info["ms_total"][5:])
            line["reference_cpu"] = {"ms_per_frame_median": ms[len(ms) // 2], "frames_per_s": 1e3 / ms[len(ms) // 2], "threads": info["threads"], "frames": 60}
        print(json.dumps(line))
        del outs
        r.close()


if __name__ == "__main__":
    main()
