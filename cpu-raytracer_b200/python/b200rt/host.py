"""HostScene: Python face of libb200rt_host.so — the reference's Scene surface on the host.

    scene = HostScene("config1.scene")
    scene.update(dt=0.0)                 # Scene::update: transforms, camera, per-frame TLAS rebuild
    pack = scene.pack()                  # flat arrays for the device path
"""
import ctypes as C

import numpy as np

from . import _ffi
from .pack import ScenePack


class HostError(RuntimeError):
    pass


class HostScene:
    def __init__(self, scene_file, width=None, height=None):
        self._lib = _ffi.host_lib()
        self._h = C.c_void_p()
        if self._lib.b200rt_scene_load(str(scene_file).encode(), C.byref(self._h)) != 0:
            raise HostError(self._lib.b200rt_host_last_error().decode())
        if width and height:
            self.resize(width, height)

    def close(self):
        if self._h:
            self._lib.b200rt_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc < 0:
            raise HostError(self._lib.b200rt_host_last_error().decode())
        return rc

    def resize(self, width, height):
        self._check(self._lib.b200rt_scene_resize(self._h, width, height))

    @property
    def resolution(self):
        w, h = C.c_int(), C.c_int()
        self._check(self._lib.b200rt_scene_resolution(self._h, C.byref(w), C.byref(h)))
        return w.value, h.value

    def set_camera(self, position=None, rotation=None):
        p = (C.c_float * 3)(*position) if position is not None else None
        r = (C.c_float * 4)(*rotation) if rotation is not None else None
        self._check(self._lib.b200rt_scene_set_camera(self._h, p, r))

    def update(self, dt=0.0, animation=0):
        self._check(self._lib.b200rt_scene_update(self._h, float(dt), int(animation)))

    def static_view(self):
        s = _ffi.StaticScene()
        self._check(self._lib.b200rt_scene_static_view(self._h, C.byref(s)))
        return s

    def frame_view(self):
        f = _ffi.Frame()
        self._check(self._lib.b200rt_scene_frame_view(self._h, C.byref(f)))
        return f

    def pack(self):
        w, h = self.resolution
        return ScenePack.capture(self.static_view(), self.frame_view(), w, h)

    def blas_original_indices(self, blas):
        n = self._check(self._lib.b200rt_scene_blas_original_indices(self._h, blas, None, 0))
        out = np.zeros(n, dtype=np.int32)
        self._check(self._lib.b200rt_scene_blas_original_indices(self._h, blas, out.ctypes.data, n))
        return out


def tile_count(width, height):
    tx, ty = C.c_int(), C.c_int()
    n = _ffi.host_lib().b200rt_tile_count(width, height, C.byref(tx), C.byref(ty))
    return n, tx.value, ty.value
