#!/usr/bin/env python3
"""A few frames of the bench scene through b200rt_render_frame, for ncu captures of single kernels
(development tool):  python tools/ncu_frame.py && ncu --set full -k regex:fxaa_kernel -s 2 -c 1 -o out python tools/ncu_frame.py"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python")); sys.path.insert(0, REPO)
import bench
from b200rt.device import Renderer
w = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "sponza1080"]
_, scene = bench.build_scene(w)
pack = scene.pack()
r = Renderer(w["width"], w["height"], w["bounces"], w["filter"])
r.upload_static(pack)
for _ in range(6):     # the tile order needs a few frames to settle on the measured costs
    stats = r.render(pack)["stats"]
print(stats["device_ms_render"], stats["device_ms_fxaa"])
r.close()
