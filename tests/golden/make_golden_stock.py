#!/usr/bin/env python3
"""Adds what the reference produces in its STOCK configuration to the fixtures: tests/golden/stock_frames.npz.

The fixtures of make_golden.py come from the reference's --per-ray mode (one ray per packet, no FMA contraction):
the semantics one CUDA thread has to reproduce bit for bit. The reference as shipped traces packets of
SIMD_LANE_SIZE rays through a shared stack, and its release build lets the compiler contract multiply-adds. For
every case of make_golden.py this script stores

  <case>/stock_frame_buffer    the timed build (ref_l8_*, -mfma contraction, 8-ray packets): frame buffer
  <case>/packet_sha256         the id-instrumented build (-ffp-contract=off) in packet mode: SHA-256 of its frame
                               buffer, of its primary-hit ids and of its primary-hit distances (3 x 32 bytes: the
                               arrays themselves equal the per-ray ones, which the digests let a test assert)

so that tests/test_stock_reference.py can state how far the per-ray fixtures (hence the device path, which equals
them bit for bit) are from the stock reference. Run in the build container only (needs oracle/_ref):

    python tests/golden/make_golden_stock.py
"""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import make_golden as mg  # noqa: E402

W, H = mg.W, mg.H


def run(variant, scene_file, frames, dt, tmp):
    prefix = os.path.join(tmp, "stock")
    subprocess.run([os.path.join(mg.BIN, "ref_" + variant), scene_file, "--out", prefix, "--frames", str(frames + 1), "--dt", repr(dt),
                    "--dump", str(frames), "--width", str(W), "--height", str(H)], check=True, capture_output=True, text=True)
    tag = "%s.f%04d" % (prefix, frames)
    res = {"frame_buffer": np.fromfile(tag + ".u32", dtype=np.uint32).reshape(H, W)}
    if os.path.exists(tag + ".ids"):
        res["ids"] = np.fromfile(tag + ".ids", dtype=np.int32).reshape(H, W, 2)
        res["t"] = np.fromfile(tag + ".hit", dtype=np.float32).reshape(H, W, 4)[..., 0].copy()
    for f in os.listdir(tmp):
        if f.startswith("stock."):
            os.remove(os.path.join(tmp, f))
    return res


def main():
    arrays = {}
    # Primary-hit ids index the leaves of the tree in use, and the release build's SBVH builder (contracted
    # multiply-adds in the SAH) does not produce the same tree as the -ffp-contract=off build the fixtures were made
    # with: start from no caches and let the id-instrumented build write them, as make_golden.py does.
    for root, _, files in os.walk(mg.DATA):
        for f in files:
            if f.endswith((".bvh", ".b200bvh")):
                os.remove(os.path.join(root, f))
    with tempfile.TemporaryDirectory() as tmp:
        for name, builder, variant, bounces, filt, animation, frames in mg.CASES:
            if name.startswith("switch_"):
                continue
            scene_file = builder(mg.DATA, os.path.join(tmp, name + ".scene"), W, H)
            dt = mg.DT if animation else 0.0
            packet = run(variant, scene_file, frames, dt, tmp)               # first: writes the <obj>.bvh caches
            stock = run(variant.replace("_ids", ""), scene_file, frames, dt, tmp)
            arrays[name + "/stock_frame_buffer"] = stock["frame_buffer"]
            arrays[name + "/packet_sha256"] = np.frombuffer(b"".join(hashlib.sha256(packet[k].tobytes()).digest()
                                                                     for k in ("frame_buffer", "ids", "t")), dtype=np.uint8)
            print(name)
    path = os.path.join(HERE, "stock_frames.npz")
    np.savez_compressed(path, **arrays)
    print("%s %.1f KiB" % (path, os.path.getsize(path) / 1024.0))


if __name__ == "__main__":
    main()
