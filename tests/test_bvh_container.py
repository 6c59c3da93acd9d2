"""CPU: the versioned `<obj>.b200bvh` container around the reference's headerless `<obj>.bvh` dump
(SURVEY 8(f) row 4): round trip, every kind of damage detected, loader preference and fallback."""
import ctypes as C
import struct
import zlib

import numpy as np
import pytest

from b200rt import _ffi, scenes, standin
from b200rt.host import HostScene


def _lib():
    lib = _ffi.host_lib()
    lib.b200rt_bvh_to_container.argtypes = [C.c_char_p, C.c_char_p]
    lib.b200rt_bvh_container_to_legacy.argtypes = [C.c_char_p, C.c_char_p]
    return lib


@pytest.fixture()
def mesh_scene(tmp_path):
    standin.lens(str(tmp_path / "lens.obj"), n=12)
    scenes.write_sky(str(tmp_path / "sky.float"), 16)
    s = scenes.SceneWriter(32, 32)
    s.camera((14.0, 3.0, 0.0), (0.0, -0.707107, 0.0, 0.707107))
    s.sky(str(tmp_path / "sky.float"))
    s.mesh(str(tmp_path / "lens.obj"), (5.0, 3.0, 0.0))
    s.point_light((50.0, 50.0, 50.0), (6.0, 12.0, 6.0))
    return tmp_path, s.write(str(tmp_path / "one.scene"))


def test_loader_writes_both_forms_and_they_agree(mesh_scene):
    d, path = mesh_scene
    first = HostScene(path).pack()
    legacy, container = (d / "lens.obj.bvh").read_bytes(), (d / "lens.obj.b200bvh").read_bytes()
    magic, endian, version, tris, nodes, refs, crc = struct.unpack("<8sIIIIII", container[:32])
    assert magic == b"B200BVH\0" and endian == 0x01020304 and version == 1
    assert container[32:] == legacy                         # same payload: the reference can read what we cache
    assert crc == zlib.crc32(legacy)
    assert len(legacy) == 12 + tris * 100 + nodes * 32 + refs * 4
    (d / "lens.obj.bvh").unlink()                          # container alone is enough
    again = HostScene(path).pack()
    for key in ("s_nodes", "s_triangles_hot", "s_triangles_cold"):
        assert np.array_equal(first.a[key], again.a[key]), key


def test_conversion_round_trip(mesh_scene, tmp_path):
    d, path = mesh_scene
    HostScene(path).close()
    lib = _lib()
    legacy = d / "lens.obj.bvh"
    assert lib.b200rt_bvh_to_container(str(legacy).encode(), str(tmp_path / "a.b200bvh").encode()) == 0
    assert lib.b200rt_bvh_container_to_legacy(str(tmp_path / "a.b200bvh").encode(), str(tmp_path / "a.bvh").encode()) == 0
    a, b = (tmp_path / "a.bvh").read_bytes(), legacy.read_bytes()
    assert len(a) == len(b)
    # identical except node 1, which the loader clears (unreachable, garbage in reference-written files)
    tris = struct.unpack("<i", b[:4])[0]
    node1 = 4 + tris * 100 + 4 + 32
    assert a[:node1] == b[:node1] and a[node1 + 32:] == b[node1 + 32:]


@pytest.mark.parametrize("damage", ["magic", "endian", "version", "flip", "truncate", "append", "count"])
def test_damage_is_detected_and_the_loader_falls_back(mesh_scene, damage):
    d, path = mesh_scene
    good = HostScene(path).pack()
    container = d / "lens.obj.b200bvh"
    blob = bytearray(container.read_bytes())
    if damage == "magic":
        blob[0] ^= 0xFF
    elif damage == "endian":
        blob[8:12] = struct.pack(">I", 0x01020304)
    elif damage == "version":
        blob[12:16] = struct.pack("<I", 99)
    elif damage == "flip":
        blob[len(blob) // 2] ^= 0x01
    elif damage == "truncate":
        blob = blob[:-5]
    elif damage == "append":
        blob += b"\0\0\0\0"
    elif damage == "count":
        blob[20:24] = struct.pack("<I", struct.unpack("<I", blob[20:24])[0] + 1)
    container.write_bytes(bytes(blob))
    lib = _lib()
    assert lib.b200rt_bvh_container_to_legacy(str(container).encode(), str(d / "x.bvh").encode()) != 0
    assert lib.b200rt_host_last_error()
    # the scene loader ignores the damaged container and reads the reference-format dump beside it
    again = HostScene(path).pack()
    assert np.array_equal(good.a["s_nodes"], again.a["s_nodes"])
