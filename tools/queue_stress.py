#!/usr/bin/env python3
"""Renders many frames through the deep-pixel queue and checks every one of them (development tool / GPU call).

    python tools/queue_stress.py [frames per case]

The queue's protocol is concurrent; a rare stall or a lost item shows only in repetition. For each case the scene is
animated (SCENE_DYNAMIC) or the camera moved a little every frame, the frame rendered with deep_queue glass and all, and
compared with the depth-first frame of the same inputs: frame buffer and ray counters must be identical. Each case runs
under its own timeout in a child process, so a stall is reported, not waited for. One JSON line per case.
"""
import json
import os
import subprocess
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(scene_name, queue, frames):
    sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
    sys.path.insert(0, os.path.join(REPO, "tools"))
    import numpy as np
    import variant_bench
    from b200rt.device import Renderer
    scene, (w, h, bounces, filt) = variant_bench.build_scene(scene_name)
    plain = Renderer(w, h, bounces, filt, deep_queue="off")
    queued = Renderer(w, h, bounces, filt, deep_queue=queue)
    pack = scene.pack()
    plain.upload_static(pack); queued.upload_static(pack)
    bad, stalled, t0, queue_nodes = 0, 0, time.time(), 0
    for i in range(frames):
        scene.update(1.0 / 60.0, 1 if scene_name == "dynamic" else 0)
        frame_pack = pack.with_frame(scene.frame_view())
        a = plain.render(frame_pack, want_presented=False)
        b = queued.render(frame_pack, want_presented=False)
        queue_nodes += b["stats"]["queue_nodes"]
        same = np.array_equal(a["frame_buffer"], b["frame_buffer"]) and all(a["stats"][k] == b["stats"][k] for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays"))
        bad += 0 if same else 1
        if i > 0 and b["stats"]["queue_nodes"] == 0:
            stalled += 1                      # the watchdog fired earlier: the context went depth-first
    print(json.dumps({"scene": scene_name, "queue": queue, "frames": frames, "frames_that_differ": bad, "frames_without_queue": stalled,
                      "queue_nodes_total": queue_nodes, "seconds": round(time.time() - t0, 1)}), flush=True)


def main():
    if os.environ.get("STRESS_CHILD"):
        return child(os.environ["STRESS_SCENE"], os.environ["STRESS_QUEUE"], int(os.environ["STRESS_FRAMES"]))
    frames = sys.argv[1] if len(sys.argv) > 1 else "300"
    for scene in ("dynamic", "dielectrics", "sponza"):
        for queue in ("glass", "all"):
            env = dict(os.environ, STRESS_CHILD="1", STRESS_SCENE=scene, STRESS_QUEUE=queue, STRESS_FRAMES=frames)
            try:
                p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=float(os.environ.get("STRESS_TIMEOUT", "240")))
                line = [l for l in p.stdout.splitlines() if l.startswith("{")]
                print(line[-1] if line else json.dumps({"scene": scene, "queue": queue, "error": (p.stderr or p.stdout)[-600:]}), flush=True)
            except subprocess.TimeoutExpired:
                print(json.dumps({"scene": scene, "queue": queue, "error": "timeout: a stall the watchdog did not end"}), flush=True)


if __name__ == "__main__":
    main()
