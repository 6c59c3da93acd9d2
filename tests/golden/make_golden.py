#!/usr/bin/env python3
"""Generates the golden fixtures under tests/golden/ from the REAL reference renderer.

Run in the build container only (needs /root/reference and oracle/_ref, see
oracle/ref_recipe/build_ref.py and stage_assets.py):

    python tests/golden/make_golden.py

For every case it
  1. writes a scene description (cpu-raytracer_b200/python/b200rt/scenes.py),
  2. runs the id-instrumented, -ffp-contract=off build of the reference in --per-ray mode
     (one ray per packet through the reference's own Raytracer::bounce) and keeps its frame
     buffer, primary-hit ids and hit records, plus the per-frame instance matrices / TLAS it built,
  3. loads the same scene through the host library, advances it the same number of frames and
     captures the flat arrays (ScenePack) the device path and the oracle consume,
  4. stores everything in one compressed .npz.
The fixtures travel with the repository, so the GPU box needs neither the reference nor its assets.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))

from b200rt import scenes  # noqa: E402
from b200rt.host import HostScene  # noqa: E402

DATA = os.path.join(REPO, "oracle", "_ref", "Data")
BIN = os.path.join(REPO, "oracle", "_ref", "bin")

W, H = 480, 270

# name, scene builder, reference variant, bounces, filter, animation, frames to advance (dt = 1/60)
CASES = [
    ("monkey_b1_aniso", scenes.monkey_plane, "l8_b1_aniso_ids", 1, "aniso", 0, 0),
    ("dynamic_b3_aniso_f0", scenes.dynamic, "l8_b3_aniso_dyn_ids", 3, "aniso", 1, 0),
    ("dynamic_b3_aniso_f45", scenes.dynamic, "l8_b3_aniso_dyn_ids", 3, "aniso", 1, 45),
    ("dynamic_b3_tri_f0", scenes.dynamic, "l8_b3_tri_ids", 3, "tri", 0, 0),
    ("dielectrics_b8_ewa", scenes.dielectrics_small, "l8_b8_ewa_ids", 8, "ewa", 0, 0),
    # the other settings of the reference's hot-path switches (Config.h: BVH_TRAVERSAL_STRATEGY, TEXTURE_SAMPLE_MODE);
    # "switch_" fixtures are read by tests/test_config_switches.py only
    ("switch_naive_dynamic_b3_f45", scenes.dynamic, "l8_b3_aniso_naive_dyn_ids", 3, "aniso", 1, 45),
    ("switch_nearest_monkey_b1", scenes.monkey_plane, "l8_b1_aniso_nearest_ids", 1, "aniso", 0, 0),
    ("switch_bilinear_monkey_b1", scenes.monkey_plane, "l8_b1_aniso_bilinear_ids", 1, "aniso", 0, 0),
]
DT = 1.0 / 60.0


def run_reference(variant, scene_file, frames, dt, tmp):
    prefix = os.path.join(tmp, "ref")
    cmd = [os.path.join(BIN, "ref_" + variant), scene_file, "--out", prefix, "--per-ray",
           "--frames", str(frames + 1), "--dt", repr(dt), "--dump", str(frames), "--width", str(W), "--height", str(H)]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    info = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    tag = "%s.f%04d" % (prefix, frames)
    res = {
        "ref_frame_buffer": np.fromfile(tag + ".u32", dtype=np.uint32).reshape(H, W),
        "ref_ids": np.fromfile(tag + ".ids", dtype=np.int32).reshape(H, W, 2),
        "ref_hit": np.fromfile(tag + ".hit", dtype=np.float32).reshape(H, W, 4),
        "ref_frame_inputs": np.fromfile(tag + ".frame", dtype=np.uint8),
        "ref_rays": np.array([info["rays"][k] for k in ("primary", "shadow", "reflection", "refraction")], dtype=np.int64),
    }
    return res


def main():
    only = set(sys.argv[1:])
    with tempfile.TemporaryDirectory() as tmp:
        for name, builder, variant, bounces, filt, animation, frames in CASES:
            if only and name not in only:
                continue
            scene_file = builder(DATA, os.path.join(tmp, name + ".scene"), W, H)
            dt = DT if animation else 0.0
            ref = run_reference(variant, scene_file, frames, dt, tmp)

            scene = HostScene(scene_file)
            for _ in range(frames + 1):            # the reference calls scene.update(dt) before EVERY frame (Main.cpp:54)
                scene.update(dt, animation)
            pack = scene.pack()

            arrays = dict(pack.a)
            arrays.update(ref)
            traversal = "naive" if "_naive" in variant else "ordered"
            texture_mode = "nearest" if "_nearest" in variant else "bilinear" if "_bilinear" in variant else "mipmap"
            arrays["meta"] = np.frombuffer(json.dumps(dict(name=name, variant=variant, bounces=bounces, filter=filt,
                                                           traversal=traversal, texture_mode=texture_mode,
                                                           animation=animation, frames=frames, dt=dt, width=W, height=H)).encode(), dtype=np.uint8)
            path = os.path.join(HERE, name + ".npz")
            np.savez_compressed(path, **arrays)
            print("%-24s %7.1f KiB  rays %s" % (name, os.path.getsize(path) / 1024.0, ref["ref_rays"].tolist()))


if __name__ == "__main__":
    main()
