#!/usr/bin/env python3
"""TEST INFRASTRUCTURE — stage the reference's shipped ASSETS (not sources) for the oracle runs.

The reference writes its `<obj>.bvh` caches next to the OBJ files (BottomLevelBVH.cpp:26,53) and
/root/reference is read-only, so the meshes / textures the parity scenes use are copied into the
git-ignored oracle/_ref/Data/. A procedural sky probe replaces the missing Sky_Probes/*.float
(SURVEY.md §8(d)). Nothing here is needed on the GPU box: the committed fixtures under
tests/golden/ carry the flattened scene arrays.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("B200RT_REFERENCE", "/root/reference")
DATA = os.path.join(os.path.dirname(HERE), "_ref", "Data")

FILES = ["Monkey.obj", "Monkey.mtl", "Diamond.obj", "Diamond.mtl", "icosphere.obj", "icosphere.mtl", "Rock.obj", "Rock.mtl",
         "Torus.obj", "Torus.mtl", "Magnifier.obj", "Magnifier.mtl", "Concave.obj", "Concave.mtl", "floor.png", "bricks.png",
         "Cube.obj", "Cube.mtl", "Plane.obj", "Plane.mtl"]


def main():
    if not os.path.isdir(REF):
        print(f"reference not present at {REF}; nothing staged")
        return 0
    os.makedirs(DATA, exist_ok=True)
    for f in FILES:
        dst = os.path.join(DATA, f)
        if not os.path.exists(dst):
            shutil.copyfile(os.path.join(REF, "Data", f), dst)
    # Linux is case sensitive: the shipped scene asks for "Floor.png" (Scene.cpp:23, Cube.mtl)
    if not os.path.exists(os.path.join(DATA, "Floor.png")):
        shutil.copyfile(os.path.join(DATA, "floor.png"), os.path.join(DATA, "Floor.png"))
    sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
    from b200rt.scenes import write_sky
    sky = os.path.join(DATA, "sky_256.float")
    if not os.path.exists(sky):
        write_sky(sky, 256)
    print("assets staged in", DATA)
    return 0


if __name__ == "__main__":
    sys.exit(main())
