#!/usr/bin/env python3
"""Times kernel variants / run-time settings of libb200rt.so side by side (development tool).

    python tools/variant_bench.py SPEC [SPEC ...]
    SPEC = label[,lib=path/to/libb200rt_x.so][,queue=off|glass|all|auto][,arith=exact|fast][,inline=N][,flush=0|1][,looks=N][,shift=N][,batch=N][,keep=0|1]

Each SPEC runs in its own process (B200RT_LIB / B200RT_DEEP_QUEUE / B200RT_ARITHMETIC / B200RT_QUEUE_INLINE_BELOW
select library and settings). First the golden fixtures are rendered and compared with the real reference's stored
frames (exact arithmetic: ids equal, frame buffer within the test bars; fast: ids-%, PSNR and max channel error are
reported), then the scenes named in VARIANT_SCENES (default sponza,dielectrics; also sibenik, dynamic) are timed with
resident inputs (CUDA events inside the library, L2 flushed between launches), the whole frame and — VARIANT_SHARE=8 —
one GPU's share of an 8-GPU run (tile_stride 8). One JSON line per SPEC.
"""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_scene(name):
    sys.path.insert(0, REPO)
    import bench
    from b200rt import scenes, standin
    from b200rt.host import HostScene
    data = os.path.join(REPO, "oracle", "_ref", "Data")
    if name == "sponza":
        scene = HostScene(standin.sponza_standin(width=1920, height=1080)); frames, spec = 0, (1920, 1080, 3, "tri")
    elif name == "sibenik":
        scene = HostScene(standin.sibenik_standin(width=3840, height=2160)); frames, spec = 0, (3840, 2160, 3, "aniso")
    elif name == "dielectrics":
        scene = HostScene(standin.dielectrics_hall(data, width=1920, height=1080)); frames, spec = 0, (1920, 1080, 8, "ewa")
    elif name == "dynamic":
        import tempfile
        scene = HostScene(scenes.dynamic(data, os.path.join(tempfile.mkdtemp(), "dynamic.scene"), 1920, 1080)); frames, spec = 45, (1920, 1080, 3, "aniso")
    else:
        raise SystemExit("unknown scene " + name)
    for _ in range(frames + 1):
        scene.update(1.0 / 60.0 if frames else 0.0, 1 if frames else 0)
    return scene, spec


def child():
    sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
    sys.path.insert(0, os.path.join(REPO, "tests"))
    sys.path.insert(0, REPO)
    import numpy as np
    import bench
    import conftest
    from b200rt.device import Renderer

    res = {"spec": os.environ["VARIANT_LABEL"], "lib": os.path.basename(os.environ.get("B200RT_LIB", "libb200rt.so")),
           "queue": os.environ.get("B200RT_DEEP_QUEUE", "default"), "arith": os.environ.get("B200RT_ARITHMETIC", "exact"), "inline": os.environ.get("B200RT_QUEUE_INLINE_BELOW", "0")}
    exact = res["arith"] == "exact"
    bad, worst = [], {"ids": 1.0, "psnr": 99.0, "max_err": 0}
    for name in conftest.golden_names():
        g = conftest.Golden(name)
        m = g.meta
        r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
        r.upload_static(g.pack)
        out = r.render(g.pack, want_ids=True)
        r.close()
        a, b = conftest.unpack_rgb(out["frame_buffer"]), conftest.unpack_rgb(g.ref_frame_buffer)
        diff = np.abs(a - b)
        ids = float((out["ids"] == g.ref_ids).all(-1).mean())
        worst = {"ids": min(worst["ids"], ids), "psnr": min(worst["psnr"], conftest.psnr(a, b)), "max_err": max(worst["max_err"], int(diff.max()))}
        rays_ok = [int(out["stats"][k]) for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays")] == [int(x) for x in g.ref_rays]
        if exact and (ids < 1.0 or (diff.max(-1) > 0).mean() > 2e-3 or diff.max() > 8 or not rays_ok):
            bad.append([name, ids, int((diff.max(-1) > 0).sum()), int(diff.max()), rays_ok])
        if not exact and (ids < 0.999 or conftest.psnr(a, b) < 50.0):
            bad.append([name, ids, conftest.psnr(a, b), int(diff.max())])
    res["parity_failures"] = bad
    res["goldens_worst"] = worst

    share = int(os.environ.get("VARIANT_SHARE", "0"))
    steps = int(os.environ.get("VARIANT_STEPS", "10"))
    for scene_name in [n for n in os.environ.get("VARIANT_SCENES", "sponza,dielectrics").split(",") if n]:
        scene, (w, h, bounces, filt) = build_scene(scene_name)
        pack = scene.pack()
        entry = {}
        for stride in [1] + ([share] if share > 1 else []):
            r = Renderer(w, h, bounces, filt, tile_stride=stride, tile_phase=0)
            r.upload_static(pack)
            for _ in range(3):
                first = r.render(pack, want_presented=False)     # the tile order settles on the measured costs
            s = first["stats"]
            rays = s["primary_rays"] + s["shadow_rays"] + s["reflection_rays"] + s["refraction_rays"]
            flush = bench.L2_FLUSH_BYTES if os.environ.get("VARIANT_FLUSH", "1") != "0" else 0    # VARIANT_FLUSH=0: warm L2, as a frame loop leaves it
            r.time_resident(3, flush)
            ms_render, ms_fxaa = r.time_resident(steps, flush)
            r.close()
            key = "" if stride == 1 else "_share%d" % stride
            entry["ms_render" + key] = round(ms_render, 4)
            entry["tail_ms" + key] = round(s["device_ms_tail"], 4)
            if stride == 1:
                entry.update({"ms_fxaa": round(ms_fxaa, 4), "mrays_s": round(rays / (ms_render + ms_fxaa) / 1e3, 1), "rays": rays,
                              "checksum": int(first["frame_buffer"].astype(np.uint64).sum()),
                              "queue_items": [s["queue_nodes"], s["queue_bounce_items"], s["queue_shadow_items"], s["queue_fallbacks"]]})
        res[scene_name] = entry
    print(json.dumps(res), flush=True)


def main():
    if os.environ.get("VARIANT_CHILD"):
        return child()
    for spec in sys.argv[1:] or ["shipped"]:
        parts = spec.split(",")
        env = dict(os.environ, VARIANT_CHILD="1", VARIANT_LABEL=parts[0])
        for item in parts[1:]:
            key, value = item.split("=", 1)
            if key == "lib":
                env["B200RT_LIB"] = os.path.abspath(value)
            elif key == "queue":
                env["B200RT_DEEP_QUEUE"] = value
            elif key == "arith":
                env["B200RT_ARITHMETIC"] = value
            elif key == "inline":
                env["B200RT_QUEUE_INLINE_BELOW"] = value
            elif key == "flush":
                env["VARIANT_FLUSH"] = value
            elif key == "looks":
                env["B200RT_QUEUE_IDLE_LOOKS"] = value
            elif key == "shift":
                env["B200RT_QUEUE_IDLE_SHIFT"] = value
            elif key == "batch":
                env["B200RT_QUEUE_ITEM_BATCH"] = value
            elif key == "keep":
                env["B200RT_QUEUE_KEEP_REFRACTION"] = value
            else:
                raise SystemExit("bad SPEC item " + item)
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=float(os.environ.get("VARIANT_TIMEOUT", "180")))
            line = [l for l in p.stdout.splitlines() if l.startswith("{")]
            print(line[-1] if line else json.dumps({"spec": parts[0], "error": (p.stderr or p.stdout)[-800:]}), flush=True)
        except subprocess.TimeoutExpired:
            print(json.dumps({"spec": parts[0], "error": "timeout"}), flush=True)


if __name__ == "__main__":
    main()
