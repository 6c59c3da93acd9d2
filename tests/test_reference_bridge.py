"""The reference itself, with Main.cpp:56-59 replaced by libb200rt.so — the drop-in, run.

oracle/ref_recipe/bridge/B200Bridge.h is the binding INTEGRATION.md tells a maintainer of the reference to add. Here it
is compiled against the reference's own headers into the headless reference build (variant "..._bridge"): the
reference's loaders build the Scene, the reference's Scene::update animates it and rebuilds the TLAS, and every frame is
rendered twice from that one Scene object — by the reference's Raytracer on the host cores and, through the bridge and
the C ABI, by the library. The two frames of one process are compared: no scene file parser of ours, no host library of
ours, no packing code of ours is between the reference's data and the kernels.

The CPU test links the binary against the kernels' source compiled for the host (tests/emu); the gpu test against the
real library on the B200.
"""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import REPO, psnr, unpack_rgb

BINARY = os.path.join(REPO, "oracle", "_ref", "bin", "ref_l8_b3_aniso_dyn_ids_bridge")
DATA = os.path.join(REPO, "oracle", "_ref", "Data")
EMU = os.path.join(REPO, "tests", "emu")

pytestmark = pytest.mark.skipif(not os.path.exists(BINARY), reason="oracle/_ref has no bridge build (python oracle/ref_recipe/build_ref.py --default)")


def run_bridge(tmp_path, width, height, frames, dump, library_dir=None, per_ray=True):
    from b200rt import scenes
    scene = scenes.dynamic(DATA, str(tmp_path / "dynamic.scene"), width, height)
    prefix = str(tmp_path / "out")
    env = dict(os.environ)
    if library_dir:
        env["LD_LIBRARY_PATH"] = library_dir + os.pathsep + env.get("LD_LIBRARY_PATH", "")     # searched before the binary's RUNPATH
    command = [BINARY, scene, "--out", prefix, "--frames", str(frames), "--dt", repr(1.0 / 60.0), "--dump", ",".join(str(d) for d in dump), "--bridge"]
    if per_ray:
        command.append("--per-ray")
    p = subprocess.run(command, env=env, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    report = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    result = []
    for d in dump:
        tag = "%s.f%04d" % (prefix, d)
        result.append({"cpu": np.fromfile(tag + ".u32", dtype=np.uint32).reshape(height, width),
                       "gpu": np.fromfile(tag + ".gpu.u32", dtype=np.uint32).reshape(height, width),
                       "gpu_presented": np.fromfile(tag + ".gpu.rgb8", dtype=np.uint8).reshape(height, width, 3)})
    return report, result


def check(report, frames):
    import oracle_ffi
    for f in frames:
        assert np.array_equal(f["gpu"], f["cpu"])                                               # Window::plot words, bit for bit
        assert np.array_equal(f["gpu_presented"], oracle_ffi.present(f["gpu"], True))
    assert report["gpu_rays"] == report["rays"]                                                 # PerformanceStats of the last frame


@pytest.mark.timeout(900)
def test_reference_renders_through_the_bridge_on_the_emulated_kernels(tmp_path):
    subprocess.run(["make", "-C", EMU, "-s"], check=True)
    stand_in = tmp_path / "lib"
    stand_in.mkdir()
    os.symlink(os.path.join(EMU, "libb200rt_emu.so"), str(stand_in / "libb200rt.so"))
    report, frames = run_bridge(tmp_path, 160, 96, frames=3, dump=[0, 2], library_dir=str(stand_in))
    check(report, frames)


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_reference_renders_through_the_bridge_on_the_gpu(tmp_path):
    """Config 3 (the animated scene) at 640x360, 46 frames of Scene::update; frames 0, 20 and 45 compared."""
    report, frames = run_bridge(tmp_path, 640, 360, frames=46, dump=[0, 20, 45])
    check(report, frames)
    assert len(report["ms_bridge"]) == 46


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_bridge_against_the_references_packet_mode(tmp_path):
    """The shipped 4x2 packets (Raytracer.cpp:28-42) instead of one ray per packet: same image (test_stock_reference)."""
    report, frames = run_bridge(tmp_path, 640, 360, frames=2, dump=[1], per_ray=False)
    f = frames[0]
    a, b = unpack_rgb(f["gpu"]), unpack_rgb(f["cpu"])
    assert psnr(a, b) >= 50.0 and (a != b).any(-1).mean() < 0.002
