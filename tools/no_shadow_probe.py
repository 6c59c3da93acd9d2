#!/usr/bin/env python3
"""What do the shadow rays cost inside the render kernel? (development tool)

Times the bench frame and the other full-size scenes twice: as they are, and with the lights removed
(the closest-hit rays, texture filtering and the recursion are unchanged: which rays are spawned does
not depend on the lighting; only the direct-lighting terms go). The difference is what tracing the
shadow rays in line costs, tail included. One JSON line per scene.
"""
import json, os, sys, tempfile
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python")); sys.path.insert(0, REPO)
import bench
from b200rt.device import Renderer
from b200rt.pack import ScenePack
from b200rt import scenes, standin
from b200rt.host import HostScene


def build(name):
    if name == "sponza":
        _, scene = bench.build_scene(); return scene, 1920, 1080, 3, "tri"
    if name == "sibenik":
        s = HostScene(standin.sibenik_standin(width=3840, height=2160)); s.update(0.0, 0); return s, 3840, 2160, 3, "aniso"
    data = os.path.join(REPO, "oracle", "_ref", "Data")
    if name == "dielectrics":
        s = HostScene(standin.dielectrics_hall(data, width=1920, height=1080)); s.update(0.0, 0); return s, 1920, 1080, 8, "ewa"
    builder, bounces, filt, frames = {"dynamic": (scenes.dynamic, 3, "aniso", 45)}[name]
    s = HostScene(builder(data, os.path.join(tempfile.mkdtemp(), name + ".scene"), 1920, 1080))
    for _ in range(frames + 1):
        s.update(1.0 / 60.0 if frames else 0.0, 1 if frames else 0)
    return s, 1920, 1080, bounces, filt


for name in (sys.argv[1:] or ["sponza", "sibenik", "dynamic", "dielectrics"]):
    scene, w, h, bounces, filt = build(name)
    pack = scene.pack()
    arrays = dict(pack.a)
    arrays["f_lights"] = arrays["f_lights"][:0]
    line = {"scene": name}
    for label, p in (("as_is", pack), ("no_lights", ScenePack(arrays))):
        for stride in (1, 8):
            r = Renderer(w, h, bounces, filt, tile_stride=stride, tile_phase=0)
            r.upload_static(p)
            for _ in range(3):
                s = r.render(p, want_presented=False)["stats"]
            ms, _ = r.time_resident(10, bench.L2_FLUSH_BYTES)
            s = r.render(p, want_presented=False)["stats"]
            r.close()
            line["%s_stride%d" % (label, stride)] = {"render_ms": round(ms, 4), "tail_ms": round(s["device_ms_tail"], 4), "shadow_rays": s["shadow_rays"],
                                                     "other_rays": s["primary_rays"] + s["reflection_rays"] + s["refraction_rays"]}
    print(json.dumps(line), flush=True)
