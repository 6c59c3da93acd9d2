#!/usr/bin/env python3
"""One small golden frame with the deep-pixel queue on (development tool; run under compute-sanitizer)."""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python")); sys.path.insert(0, os.path.join(REPO, "tests"))
import numpy as np
import conftest
from b200rt.device import Renderer
g = conftest.Golden(sys.argv[1] if len(sys.argv) > 1 else "dielectrics_b8_ewa")
m = g.meta
for queue in ("glass", "all"):
    r = Renderer(m["width"], m["height"], m["bounces"], m["filter"], deep_queue=queue)
    r.upload_static(g.pack)
    out = r.render(g.pack)
    r.close()
    s = out["stats"]
    print(queue, "frame buffer equals the reference's:", bool(np.array_equal(out["frame_buffer"], g.ref_frame_buffer)), "differing px", int((out["frame_buffer"] != g.ref_frame_buffer).sum()),
          "queue", s["queue_nodes"], s["queue_bounce_items"], s["queue_shadow_items"], s["queue_fallbacks"], "ms", round(s["device_ms_render"], 3), flush=True)
