#!/usr/bin/env python3
"""Times kernel variants of libb200rt.so side by side on the bench workload (development tool).

    python tools/variant_bench.py build/libb200rt_a.so build/libb200rt_b.so ...

Each variant runs in its own process (B200RT_LIB selects the library): first the golden fixtures
must come out identical to the real reference (ids exact, frame buffer within the test bars),
then the Sponza stand-in frame is timed with resident inputs (CUDA events inside the library,
L2 flushed between launches). One JSON line per variant.
"""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
    sys.path.insert(0, os.path.join(REPO, "tests"))
    sys.path.insert(0, REPO)
    import numpy as np
    import bench
    import conftest
    from b200rt.device import Renderer

    res = {"lib": os.path.basename(os.environ.get("B200RT_LIB", "libb200rt.so"))}
    bad = []
    for name in conftest.golden_names():
        g = conftest.Golden(name)
        m = g.meta
        r = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
        r.upload_static(g.pack)
        out = r.render(g.pack, want_ids=True)
        r.close()
        ids_ok = bool(np.array_equal(out["ids"], g.ref_ids))
        diff = np.abs(conftest.unpack_rgb(out["frame_buffer"]) - conftest.unpack_rgb(g.ref_frame_buffer))
        if not ids_ok or (diff.max(-1) > 0).mean() > 2e-3 or diff.max() > 8:
            bad.append([name, ids_ok, int((diff.max(-1) > 0).sum()), int(diff.max())])
    res["parity_failures"] = bad

    which = os.environ.get("VARIANT_SCENES", "sponza").split(",")
    for scene_name in which:
        if scene_name == "sponza":
            _, scene = bench.build_scene()
            w, h, bounces, filt = bench.WIDTH, bench.HEIGHT, bench.BOUNCES, bench.FILTER
        elif scene_name == "sibenik":
            from b200rt import standin
            from b200rt.host import HostScene
            scene = HostScene(standin.sibenik_standin(width=3840, height=2160))
            scene.update(0.0, 0)
            w, h, bounces, filt = 3840, 2160, 3, "tri"
        else:   # dynamic | dielectrics: the reference's shipped meshes (oracle/_ref/Data travels with the repo)
            import tempfile
            from b200rt import scenes
            from b200rt.host import HostScene
            data = os.path.join(REPO, "oracle", "_ref", "Data")
            builder, bounces, filt, frames = {"dynamic": (scenes.dynamic, 3, "aniso", 45), "dielectrics": (scenes.dielectrics_small, 8, "ewa", 0)}[scene_name]
            w, h = 1920, 1080
            scene = HostScene(builder(data, os.path.join(tempfile.mkdtemp(), scene_name + ".scene"), w, h))
            for _ in range(frames + 1):
                scene.update(1.0 / 60.0 if frames else 0.0, 1 if frames else 0)
        pack = scene.pack()
        r = Renderer(w, h, bounces, filt)
        r.upload_static(pack)
        first = r.render(pack, want_presented=False)
        s = first["stats"]
        rays = s["primary_rays"] + s["shadow_rays"] + s["reflection_rays"] + s["refraction_rays"]
        r.time_resident(3, bench.L2_FLUSH_BYTES)
        ms_render, ms_fxaa = r.time_resident(int(os.environ.get("VARIANT_STEPS", "20")), bench.L2_FLUSH_BYTES)
        r.close()
        res[scene_name] = {"ms_render": round(ms_render, 4), "ms_fxaa": round(ms_fxaa, 4), "mrays_s": round(rays / (ms_render + ms_fxaa) / 1e3, 1),
                           "checksum": int(first["frame_buffer"].astype(np.uint64).sum())}
    print(json.dumps(res), flush=True)


def main():
    if os.environ.get("VARIANT_CHILD"):
        return child()
    libs = sys.argv[1:] or [os.path.join(REPO, "cpu-raytracer_b200", "libb200rt.so")]
    for lib in libs:
        env = dict(os.environ, VARIANT_CHILD="1", B200RT_LIB=os.path.abspath(lib))
        p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True)
        line = [l for l in p.stdout.splitlines() if l.startswith("{")]
        print(line[-1] if line else json.dumps({"lib": lib, "error": (p.stderr or p.stdout)[-600:]}), flush=True)


if __name__ == "__main__":
    main()
