#!/usr/bin/env python3
"""How long does one GPU's share of the frame take, and how much of that with part of the GPU idle?
(development tool)    python tools/tail_probe.py [sponza|sibenik]

tile_stride s emulates one rank of an s-GPU run on a single GPU (tiles t with t % s == 0).
B200RT_LIB selects a kernel variant under test."""
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
sys.path.insert(0, REPO)

import bench  # noqa: E402
from b200rt.device import Renderer  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "sponza"
if workload == "sponza":
    _, scene = bench.build_scene()
    w, h, filt = 1920, 1080, "tri"
else:
    from b200rt import standin
    from b200rt.host import HostScene
    scene = HostScene(standin.sibenik_standin(width=3840, height=2160))
    scene.update(0.0, 0)
    w, h, filt = 3840, 2160, "aniso"
pack = scene.pack()
for stride in (1, 2, 4, 8):
    r = Renderer(w, h, 3, filt, tile_stride=stride, tile_phase=0)
    r.upload_static(pack)
    for _ in range(4):
        s = r.render(pack, want_presented=False)["stats"]
    ms, _ = r.time_resident(10, bench.L2_FLUSH_BYTES)
    r.close()
    print(json.dumps({"workload": workload, "lib": os.path.basename(os.environ.get("B200RT_LIB", "libb200rt.so")), "tile_stride": stride,
                      "render_ms": round(ms, 4), "last_frame_tail_ms": round(s["device_ms_tail"], 4)}), flush=True)
