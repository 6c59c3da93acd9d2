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

Beside that pinned comparison, per config:
  * the reference AS SHIPPED — 4x2 packets (Raytracer.cpp:28-42), multiply-adds contracted (the g++ twin of the MSVC
    /fp:fast release build) — against the same GPU frame, with north_star's bars (PSNR >= 50 dB) and the max channel
    error recorded ("stock" rows);
  * the deep-pixel queue (deep_queue glass / all / auto): the same frame buffer and counters, bit for bit;
  * the FAST arithmetic unit against the pinned reference frame: >= 99.9 % primary-hit ids, >= 50 dB.
configs[3] is the scene SURVEY 8(d) states: Diamond.obj, Magnifier.obj and the glass spheres of Scene.cpp:8-19 INSIDE
the (stand-in) Sponza hall, 8 bounces, EWA — the scene `bench.py --workload dielectrics` times.
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
    if kind == "dielectrics_hall":
        return standin.dielectrics_hall(DATA, width=width, height=height)
    builder = {"monkey": scenes.monkey_plane, "dynamic": scenes.dynamic}[kind]
    return builder(DATA, os.path.join(tmp, kind + ".scene"), width, height)


# config, scene, width, height, bounces, filter, reference build, animated, frames advanced before the compared one
CASES = [
    ("configs[0]", "monkey", 1280, 720, 1, "aniso", "l8_b1_aniso_ids", 0, 0),
    ("configs[1]", "sponza", 1920, 1080, 3, "tri", "l8_b3_tri_ids", 0, 0),
    ("configs[2]", "dynamic", 1920, 1080, 3, "aniso", "l8_b3_aniso_dyn_ids", 1, 20),
    ("configs[3]", "dielectrics_hall", 1920, 1080, 8, "ewa", "l8_b8_ewa_ids", 0, 0),
    ("configs[4]", "sibenik", 3840, 2160, 3, "aniso", "l8_b3_aniso_ids", 0, 0),
]


def _reference(variant, scene_file, width, height, frames, dt, tmp, per_ray=True):
    prefix = os.path.join(tmp, "ref_" + variant)
    cmd = [os.path.join(BIN, "ref_" + variant), scene_file, "--out", prefix, "--frames", str(frames + 1),
           "--dt", repr(dt), "--dump", str(frames), "--width", str(width), "--height", str(height)] + (["--per-ray"] if per_ray else [])
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    info = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    tag = "%s.f%04d" % (prefix, frames)
    result = dict(frame_buffer=np.fromfile(tag + ".u32", dtype=np.uint32).reshape(height, width), rays=info["rays"])
    if variant.endswith("_ids"):
        result.update(ids=np.fromfile(tag + ".ids", dtype=np.int32).reshape(height, width, 2),
                      hit=np.fromfile(tag + ".hit", dtype=np.float32).reshape(height, width, 4))
    return result


def _append(line):
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    with open(os.path.join(REPO, "gpurun_out", "full_size_reference.jsonl"), "a") as f:
        f.write(json.dumps(line) + "\n")


def _record(config, kind, width, height, bounces, filt, out, ref):
    """Appends what was observed to gpurun_out/full_size_reference.jsonl (copied into profiles/)."""
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    line = {"row": "pinned (per-ray, -ffp-contract=off)", "config": config, "scene": kind, "size": [width, height], "bounces": bounces, "filter": filt,
            "pixels": width * height,
            "primary_hit_ids_differ": int((out["ids"] != ref["ids"]).any(-1).sum()),
            "hit_distance_bits_differ": int((out["hit"][..., 0].view(np.uint32) != ref["hit"][..., 0].view(np.uint32)).sum()),
            "frame_buffer_pixels_differ": int((out["frame_buffer"] != ref["frame_buffer"]).sum()),
            "max_channel_difference": int(np.abs(a - b).max()), "psnr_db": psnr(a, b),
            "rays": {k: [int(out["stats"][k + "_rays"]), int(ref["rays"][k])] for k in ("primary", "shadow", "reflection", "refraction")}}
    _append(line)


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
    return dict(out=out, ref=ref, size=(width, height), config=config, kind=kind, pack=pack, bounces=bounces, filter=filt,
                variant=variant, scene_file=scene_file, frames=frames, dt=dt, tmp=tmp)


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


def test_stock_build_of_the_reference_is_within_north_stars_bar(case):
    """The reference as shipped (packets, contracted multiply-adds) on the same scene file and <obj>.bvh caches."""
    stock_variant = case["variant"][:-len("_ids")]
    if not os.path.exists(os.path.join(BIN, "ref_" + stock_variant)):
        pytest.skip("no stock build of this configuration in oracle/_ref")
    width, height = case["size"]
    stock = _reference(stock_variant, case["scene_file"], width, height, case["frames"], case["dt"], case["tmp"], per_ray=False)
    a, b = unpack_rgb(case["out"]["frame_buffer"]), unpack_rgb(stock["frame_buffer"])
    differ = (a != b).any(-1)
    _append({"row": "stock (4x2 packets, contracted)", "config": case["config"], "scene": case["kind"], "size": [width, height],
             "frame_buffer_pixels_differ": int(differ.sum()), "fraction_of_pixels_equal": float(1.0 - differ.mean()),
             "max_channel_difference": int(np.abs(a - b).max()), "psnr_db": psnr(a, b),
             "rays": {k: [int(case["out"]["stats"][k + "_rays"]), int(stock["rays"][k])] for k in ("primary", "shadow", "reflection", "refraction")}})
    assert psnr(a, b) >= 50.0                                            # north_star's bar
    assert differ.mean() <= 0.01


@pytest.mark.parametrize("queue", ["glass", "all", "auto"])
def test_deep_pixel_queue_gives_the_same_bits(case, queue):
    from b200rt.device import Renderer
    width, height = case["size"]
    r = Renderer(width, height, case["bounces"], case["filter"], deep_queue=queue)
    r.upload_static(case["pack"])
    frames = [r.render(case["pack"], want_presented=False) for _ in range(9 if queue == "auto" else 1)]   # AUTO: through its comparison frames
    r.close()
    for out in frames:
        assert np.array_equal(out["frame_buffer"], case["out"]["frame_buffer"])
        for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays"):
            assert out["stats"][k] == case["out"]["stats"][k]
    s = frames[-1]["stats"]
    _append({"row": "deep_queue=" + queue, "config": case["config"], "scene": case["kind"], "identical": True,
             "queue_nodes": int(s["queue_nodes"]), "queue_bounce_items": int(s["queue_bounce_items"]), "queue_shadow_items": int(s["queue_shadow_items"]),
             "queue_fallbacks": int(s["queue_fallbacks"]), "queued_frames": [bool(o["stats"]["queue_nodes"] > 0) for o in frames]})


def test_fast_arithmetic_is_within_north_stars_bar(case):
    from b200rt.device import Renderer
    width, height = case["size"]
    r = Renderer(width, height, case["bounces"], case["filter"], arithmetic="fast")
    r.upload_static(case["pack"])
    out = r.render(case["pack"], want_ids=True, want_presented=False)
    r.close()
    ref = case["ref"]
    a, b = unpack_rgb(out["frame_buffer"]), unpack_rgb(ref["frame_buffer"])
    ids = float((out["ids"] == ref["ids"]).all(-1).mean())
    _append({"row": "arithmetic=fast vs pinned", "config": case["config"], "scene": case["kind"], "ids_equal": ids, "psnr_db": psnr(a, b),
             "max_channel_difference": int(np.abs(a - b).max()), "frame_buffer_pixels_differ": int((a != b).any(-1).sum())})
    assert ids >= 0.999 and psnr(a, b) >= 50.0                          # north_star's bars
