"""TEST INFRASTRUCTURE — ctypes binding of the CPU oracle (oracle/liboracle.so).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs. The product package (cpu-raytracer_b200/) never imports this module.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "cpu-raytracer_b200", "python"))

from b200rt import _ffi  # noqa: E402  (struct layouts of include/b200rt.h only)

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            raise OSError(f"{path} missing: run __graft_entry__.build()")
        _lib = C.CDLL(path)
        _lib.oracle_render.argtypes = [C.POINTER(_ffi.StaticScene), C.POINTER(_ffi.Frame), C.POINTER(_ffi.Config), C.POINTER(_ffi.Outputs), C.c_int]
        _lib.oracle_present.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        _lib.oracle_present.restype = None
        _lib.oracle_gamma_table.argtypes = [C.c_void_p]
        _lib.oracle_gamma_table.restype = None
    return _lib


def render(pack, bounces, mipmap_filter="aniso", width=None, height=None, enable_fxaa=True, threads=None,
           want_ids=True, want_float=False, tile_stride=1, tile_phase=0, max_anisotropy=8.0,
           traversal="ordered", texture_mode="mipmap"):
    """Render one frame of a ScenePack with the CPU oracle. Returns a dict of numpy arrays + stats."""
    w = width or pack.width
    h = height or pack.height
    cfg = _ffi.make_config(w, h, bounces, mipmap_filter, max_anisotropy, enable_fxaa, 0, tile_stride, tile_phase, False, traversal, texture_mode)
    static, frame = pack.static_view(), pack.frame_view()

    res = {
        "frame_buffer": np.zeros((h, w), dtype=np.uint32),
        "presented": np.zeros((h, w, 3), dtype=np.uint8),
    }
    if want_ids:
        res["ids"] = np.full((h, w, 2), -1, dtype=np.int32)
        res["hit"] = np.zeros((h, w, 4), dtype=np.float32)
    if want_float:
        res["colour"] = np.zeros((h, w, 3), dtype=np.float32)
    stats = _ffi.Stats()
    out = _ffi.Outputs(res["frame_buffer"].ctypes.data, res["presented"].ctypes.data,
                       res["colour"].ctypes.data if want_float else None,
                       res["ids"].ctypes.data if want_ids else None,
                       res["hit"].ctypes.data if want_ids else None, C.pointer(stats))
    rc = lib().oracle_render(C.byref(static), C.byref(frame), C.byref(cfg), C.byref(out), threads or (os.cpu_count() or 1))
    if rc != 0:
        raise RuntimeError("oracle_render failed")
    res["stats"] = stats.as_dict()
    return res


def present(frame_buffer, enable_fxaa=True):
    fb = np.ascontiguousarray(frame_buffer, dtype=np.uint32)
    h, w = fb.shape
    out = np.zeros((h, w, 3), dtype=np.uint8)
    lib().oracle_present(fb.ctypes.data, w, h, 1 if enable_fxaa else 0, out.ctypes.data)
    return out


def gamma_table():
    t = np.zeros(256, dtype=np.float32)
    lib().oracle_gamma_table(t.ctypes.data)
    return t
