"""GPU, BASELINE.json's five configs at their FULL sizes against the REAL reference run on the box.

oracle/_ref (the reference compiled headless from its own sources by oracle/ref_recipe, plus its
staged assets) travels to the GPU box. For every config the id-instrumented, -ffp-contract=off
build of the reference renders the frame in --per-ray mode (one ray per packet through its own
Raytracer::bounce) and the CUDA path renders the same scene file through the host library and the
C ABI: primary-hit ids, hit distances and material ids must be the same bits; the pre-FXAA frame
buffer may differ only where the last ulp of a transcendental decides a rounding; the ray counters
must agree. What was observed is appended to gpurun_out/full_size_reference.jsonl. Skipped when
oracle/_ref did not travel (the committed goldens under tests/golden/ are the same comparison at
480x270).
"""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import REPO, psnr, unpack_rgb

pytestmark = pytest.mark.gpu

DATA = os.path.join(REPO, "oracle", "_ref", "Data")
BIN = os.path.join(REPO, "oracle", "_ref", "bin")
DT = 1.0 / 60.0


def _scene_file(kind, tmp, width, height):
    from b200rt import scenes, standin
    if kind == "sponza":
        return standin.sponza_standin(width=width, height=height)
    if kind == "sibenik":
        return standin.sibenik_standin(width=width, height=height)
    builder = {"monkey": scenes.monkey_plane, "dynamic": scenes.dynamic, "dielectrics": scenes.dielectrics_small}[kind]
    return builder(DATA, os.path.join(tmp, kind + ".scene"), width, height)


# config, scene, width, height, bounces, filter, reference build, animated, frames advanced before the compared one
CASES = [
    ("configs[0]", "monkey", 1280, 720, 1, "aniso", "l8_b1_aniso_ids", 0, 0),
    ("configs[1]", "sponza", 1920, 1080, 3, "tri", "l8_b3_tri_ids", 0, 0),
    ("configs[2]", "dynamic", 1920, 1080, 3, "aniso", "l8_b3_aniso_dyn_ids", 1, 20),
    ("configs[3]", "dielectrics", 1920, 1080, 8, "ewa", "l8_b8_ewa_ids", 0, 0),
    ("configs[4]", "sibenik", 3840, 2160, 3, "aniso", "l8_b3_aniso_ids", 0, 0),
]


def _reference(variant, scene_file, width, height, frames, dt, tmp):
    prefix = os.path.join(tmp, "ref")
    cmd = [os.path.join(BIN, "ref_" + variant), scene_file, "--out", prefix, "--per-ray", "--frames", str(frames + 1),
           "--dt", repr(dt), "--dump", str(frames), "--width", str(width), "--height", str(height)]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    info = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    tag = "%s.f%04d" % (prefix, frames)
    return dict(frame_buffer=np.fromfile(tag + ".u32", dtype=np.uint32).reshape(height, width),
                ids=np.fromfile(tag + ".ids", dtype=np.int32).reshape(height, width, 2),
                hit=np.fromfile(tag + ".hit", dtype=np.float32).reshape(height, width, 4),
                rays=info["rays"])


def _record(config, kind, width, height, bounces, filt, out, ref):
    """Appends what was observed to gpurun_out/full_size_reference.jsonl (copied into profiles/)."""
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    line = {"config": config, "scene": kind, "size": [width, height], "bounces": bounces, "filter": filt,
            "pixels": width * height,
            "primary_hit_ids_differ": int((out["ids"] != ref["ids"]).any(-1).sum()),
            "hit_distance_bits_differ": int((out["hit"][..., 0].view(np.uint32) != ref["hit"][..., 0].view(np.uint32)).sum()),
            "frame_buffer_pixels_differ": int((out["frame_buffer"] != ref["frame_buffer"]).sum()),
            "max_channel_difference": int(np.abs(a - b).max()), "psnr_db": psnr(a, b),
            "rays": {k: [int(out["stats"][k + "_rays"]), int(ref["rays"][k])] for k in ("primary", "shadow", "reflection", "refraction")}}
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    with open(os.path.join(REPO, "gpurun_out", "full_size_reference.jsonl"), "a") as f:
        f.write(json.dumps(line) + "\n")


@pytest.fixture(scope="module", params=CASES, ids=[c[0] + "-" + c[1] for c in CASES])
def case(request, tmp_path_factory):
    config, kind, width, height, bounces, filt, variant, animated, frames = request.param
    if not os.path.exists(os.path.join(BIN, "ref_" + variant)) or not os.path.exists(os.path.join(DATA, "Monkey.obj")):
        pytest.skip("oracle/_ref (reference binary + assets) is not on this box")
    from b200rt.device import Renderer
    from b200rt.host import HostScene
    tmp = str(tmp_path_factory.mktemp(kind))
    scene_file = _scene_file(kind, tmp, width, height)

    scene = HostScene(scene_file)                   # builds (or loads) the BLAS caches the reference then reads
    dt = DT if animated else 0.0
    for _ in range(frames + 1):                     # Main.cpp:54: scene.update(dt) before EVERY frame
        scene.update(dt, animated)
    pack = scene.pack()
    r = Renderer(width, height, bounces, filt)
    r.upload_static(pack)
    out = r.render(pack, want_ids=True, want_presented=False)
    r.close()

    ref = _reference(variant, scene_file, width, height, frames, dt, tmp)
    _record(config, kind, width, height, bounces, filt, out, ref)
    return dict(out=out, ref=ref, size=(width, height))


def test_primary_hits_are_the_references(case):
    out, ref = case["out"], case["ref"]
    same = (out["ids"] == ref["ids"]).all(-1)
    assert same.mean() >= 0.999                                          # north_star's bar
    assert same.all(), f"{(~same).sum()} pixels with a different primary hit"
    # distance and material id carry no transcendental: the same bits
    assert np.array_equal(out["hit"][..., 0].view(np.uint32), ref["hit"][..., 0].view(np.uint32))
    assert np.array_equal(out["hit"][..., 1], ref["hit"][..., 1])
    # u, v: exact for triangles and planes; spheres go through atan2f / acosf (CUDA vs libmvec)
    np.testing.assert_allclose(out["hit"][..., 2:], ref["hit"][..., 2:], rtol=0, atol=2e-6)


def test_frame_buffer_is_the_references(case):
    out, ref = case["out"], case["ref"]
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    diff = np.abs(a - b)
    assert psnr(a, b) >= 50.0                                            # north_star's bar
    # observed: identical but for a handful of pixels one level off, where the last ulp of a
    # transcendental (expf / acosf / log2f: CUDA vs libmvec) decides a rounding
    assert (diff.max(-1) > 0).mean() <= 2e-3, f"{(diff.max(-1) > 0).sum()} pixels differ"
    assert diff.max() <= 8, f"max channel error {diff.max()}"


def test_ray_counters_are_the_references(case):
    s, r = case["out"]["stats"], case["ref"]["rays"]
    width, height = case["size"]
    assert s["primary_rays"] == width * height == r["primary"]
    for mine, theirs in (("shadow_rays", "shadow"), ("reflection_rays", "reflection"), ("refraction_rays", "refraction")):
        assert abs(s[mine] - r[theirs]) <= max(2, r[theirs] // 10000), mine
