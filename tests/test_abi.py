"""CPU: the C ABI. Both shared libraries load, export every function the headers under include/
declare, the ctypes mirrors have the C compiler's struct sizes, and the device library FAILS LOUDLY
when there is no GPU (there is no CPU fallback behind b200rt_create). No compute calls here.
"""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

from b200rt import _ffi
from conftest import REPO

INCLUDE = os.path.join(REPO, "include")


def _declared_functions(header):
    text = open(os.path.join(INCLUDE, header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rt_[a-z0-9_]+)\s*\(", text)))


def test_device_library_exports_every_declared_symbol():
    names = _declared_functions("b200rt.h")
    assert {"b200rt_create", "b200rt_upload_static", "b200rt_render_frame", "b200rt_destroy", "b200rt_last_error"} <= set(names)
    lib = _ffi.device_lib()
    for name in names:
        assert hasattr(lib, name), name


def test_host_library_exports_every_declared_symbol():
    names = _declared_functions("b200rt_host.h")
    assert "b200rt_scene_load" in names and "b200rt_scene_update" in names
    lib = _ffi.host_lib()
    for name in names:
        assert hasattr(lib, name), name


def test_headers_are_plain_c(tmp_path):
    """No C++ / torch types cross the boundary: the headers compile as C11 and give the sizes the
    ctypes mirrors assume."""
    src = tmp_path / "sizes.c"
    records = ["bvh_node", "triangle_hot", "triangle_cold", "blas", "material", "texture", "instance", "sphere", "plane", "light",
               "camera", "static_scene", "frame", "config", "stats", "outputs"]
    body = "".join('printf("%s %%zu\\n", sizeof(b200rt_%s));\n' % (r, r) for r in records)
    src.write_text('#include <stdio.h>\n#include "b200rt_host.h"\nint main(void){\n%sreturn 0;}\n' % body)
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-I", INCLUDE, str(src), "-o", str(exe)], check=True)
    sizes = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for name, expect in _ffi.SIZEOF.items():
        assert int(sizes[name]) == expect, name
    assert int(sizes["camera"]) == C.sizeof(_ffi.Camera)
    assert int(sizes["static_scene"]) == C.sizeof(_ffi.StaticScene)
    assert int(sizes["frame"]) == C.sizeof(_ffi.Frame)
    assert int(sizes["config"]) == C.sizeof(_ffi.Config)
    assert int(sizes["stats"]) == C.sizeof(_ffi.Stats)
    assert int(sizes["outputs"]) == C.sizeof(_ffi.Outputs)


def test_reference_struct_layouts():
    """the records that are byte-for-byte the reference's."""
    assert _ffi.SIZEOF["bvh_node"] == 32          # BVHNode.h:10-16
    assert _ffi.SIZEOF["triangle_hot"] == 36      # BottomLevelBVH.h:6-10
    assert _ffi.SIZEOF["triangle_cold"] == 64     # BottomLevelBVH.h:12-22


def test_create_rejects_bad_configs_with_a_message():
    lib = _ffi.device_lib()
    h = C.c_void_p()
    for kwargs in (dict(width=0, height=8), dict(width=8, height=8, bounces=9), dict(width=8, height=8, tile_stride=2, tile_phase=2)):
        args = dict(width=8, height=8, bounces=1)
        args.update(kwargs)
        cfg = _ffi.make_config(**args)
        assert lib.b200rt_create(C.byref(cfg), C.byref(h)) != 0
        assert not h.value
        assert lib.b200rt_last_error(None)


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_no_gpu_means_an_error_not_a_fallback(golden):
    from b200rt.device import DeviceError, Renderer
    with pytest.raises(DeviceError, match="CUDA"):
        Renderer(golden.meta["width"], golden.meta["height"], golden.meta["bounces"], golden.meta["filter"])


def test_product_package_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under cpu-raytracer_b200/ may import, link or name it."""
    pkg = os.path.join(REPO, "cpu-raytracer_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(root, f), errors="ignore").read()
                assert "oracle_ffi" not in text and "liboracle" not in text and "whitted_oracle" not in text, os.path.join(root, f)
