#!/usr/bin/env python3
"""Per-level timing of the deep-pixel queue (development tool; needs build/libb200rt_trace.so = -DRT_QUEUE_TRACE).
    B200RT_LIB=cpu-raytracer_b200/build/libb200rt_trace.so python tools/queue_trace.py sponza|dielectrics [stride] [queue]"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python")); sys.path.insert(0, os.path.join(REPO, "tools")); sys.path.insert(0, REPO)
import variant_bench
from b200rt.device import Renderer
name = sys.argv[1] if len(sys.argv) > 1 else "sponza"
stride = int(sys.argv[2]) if len(sys.argv) > 2 else 1
queue = sys.argv[3] if len(sys.argv) > 3 else "all"
scene, (w, h, bounces, filt) = variant_bench.build_scene(name)
pack = scene.pack()
r = Renderer(w, h, bounces, filt, tile_stride=stride, tile_phase=0, deep_queue=queue)
r.upload_static(pack)
for i in range(6):
    out = r.render(pack, want_presented=False)
print(name, "stride", stride, "queue", queue, "render ms", out["stats"]["device_ms_render"], "tail", out["stats"]["device_ms_tail"], file=sys.stderr)
r.last_frame_stats()
r.close()
