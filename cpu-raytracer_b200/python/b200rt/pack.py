"""ScenePack: the flat arrays behind b200rt_static_scene / b200rt_frame as numpy buffers.

A pack can be captured from the host library (`HostScene.pack()`), stored as a compressed .npz
(the golden fixtures under tests/golden/ are packs) and turned back into the ctypes views that
b200rt_upload_static / b200rt_render_frame — and the CPU oracle — take.
"""
import ctypes as C

import numpy as np

from . import _ffi

_STATIC_ARRAYS = [  # (field, count field, record bytes)
    ("nodes", "node_count", _ffi.SIZEOF["bvh_node"]),
    ("triangles_hot", "triangle_count", _ffi.SIZEOF["triangle_hot"]),
    ("triangles_cold", "triangle_count", _ffi.SIZEOF["triangle_cold"]),
    ("blas", "blas_count", _ffi.SIZEOF["blas"]),
    ("materials", "material_count", _ffi.SIZEOF["material"]),
    ("textures", "texture_count", _ffi.SIZEOF["texture"]),
    ("texels", "texel_count", 12),
]
_FRAME_ARRAYS = [
    ("instances", "instance_count", _ffi.SIZEOF["instance"]),
    ("tlas_nodes", "tlas_node_count", _ffi.SIZEOF["bvh_node"]),
    ("tlas_indices", "instance_count", 4),
    ("spheres", "sphere_count", _ffi.SIZEOF["sphere"]),
    ("planes", "plane_count", _ffi.SIZEOF["plane"]),
    ("lights", "light_count", _ffi.SIZEOF["light"]),
]


def _copy(ptr, nbytes):
    if not ptr or nbytes == 0:
        return np.zeros(0, dtype=np.uint8)
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(nbytes,)).copy()


class ScenePack:
    def __init__(self, arrays):
        self.a = {k: np.ascontiguousarray(v) for k, v in arrays.items()}

    # ---- capture -------------------------------------------------------------------------
    @classmethod
    def capture(cls, static_view, frame_view, width, height):
        a = {"resolution": np.array([width, height], dtype=np.int32)}
        for field, count_field, rec in _STATIC_ARRAYS:
            a["s_" + field] = _copy(getattr(static_view, field), int(getattr(static_view, count_field)) * rec)
        sky_texels = static_view.sky_size * static_view.sky_size + 1
        a["s_sky"] = _copy(static_view.sky, sky_texels * 12)
        a["s_sky_size"] = np.array([static_view.sky_size], dtype=np.int32)
        if static_view.texels_rgb8 and static_view.gamma_lut:   # level-0 8-bit images + table: input of the GPU texture ingest
            level0 = sum(int(t[2]) * int(t[3]) for t in a["s_textures"].view("<i4").reshape(-1, 22)) if a["s_textures"].size else 0
            a["s_texels_rgb8"] = _copy(static_view.texels_rgb8, level0 * 3)
            a["s_gamma_lut"] = _copy(static_view.gamma_lut, 256 * 4)
        a.update(cls._capture_frame(frame_view))
        return cls(a)

    @staticmethod
    def _capture_frame(frame_view):
        a = {}
        for field, count_field, rec in _FRAME_ARRAYS:
            a["f_" + field] = _copy(getattr(frame_view, field), int(getattr(frame_view, count_field)) * rec)
        cam = frame_view.camera
        a["f_camera"] = np.array(list(cam.position) + list(cam.top_left_corner) + list(cam.x_axis) + list(cam.y_axis), dtype=np.float32)
        a["f_ambient"] = np.array(list(frame_view.ambient), dtype=np.float32)
        return a

    def with_frame(self, frame_view):
        a = dict(self.a)
        a.update(self._capture_frame(frame_view))
        return ScenePack(a)

    # ---- storage -------------------------------------------------------------------------
    def save(self, path):
        np.savez_compressed(path, **self.a)

    @classmethod
    def load(cls, path):
        with np.load(path) as z:
            return cls({k: z[k] for k in z.files})

    # ---- views ---------------------------------------------------------------------------
    @property
    def width(self):
        return int(self.a["resolution"][0])

    @property
    def height(self):
        return int(self.a["resolution"][1])

    def _ptr(self, key):
        arr = self.a[key]
        return arr.ctypes.data if arr.size else None

    def static_view(self, gpu_texture_ingest=False):
        """gpu_texture_ingest: hand the device the 8-bit level-0 images instead of the float pyramids."""
        s = _ffi.StaticScene()
        for field, count_field, rec in _STATIC_ARRAYS:
            setattr(s, field, self._ptr("s_" + field))
            setattr(s, count_field, self.a["s_" + field].size // rec)
        s.sky = self._ptr("s_sky")
        s.sky_size = int(self.a["s_sky_size"][0])
        if gpu_texture_ingest:
            if "s_texels_rgb8" not in self.a:
                raise ValueError("this pack carries no 8-bit texture images")
            s.texels = None
            s.texels_rgb8 = self._ptr("s_texels_rgb8")
            s.gamma_lut = self._ptr("s_gamma_lut")
        return s

    def frame_view(self):
        f = _ffi.Frame()
        for field, count_field, rec in _FRAME_ARRAYS:
            setattr(f, field, self._ptr("f_" + field))
            if field != "tlas_indices":
                setattr(f, count_field, self.a["f_" + field].size // rec)
        cam = self.a["f_camera"]
        for i in range(3):
            f.camera.position[i] = cam[i]
            f.camera.top_left_corner[i] = cam[3 + i]
            f.camera.x_axis[i] = cam[6 + i]
            f.camera.y_axis[i] = cam[9 + i]
            f.ambient[i] = self.a["f_ambient"][i]
        return f

    # ---- typed accessors (tests, statistics) ---------------------------------------------------
    def triangle_count(self):
        return self.a["s_triangles_hot"].size // _ffi.SIZEOF["triangle_hot"]

    def node_count(self):
        return self.a["s_nodes"].size // _ffi.SIZEOF["bvh_node"]

    def nodes(self):
        dt = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("left_or_first", "<i4"), ("count", "<i4")])
        return self.a["s_nodes"].view(dt)

    def blas(self):
        return self.a["s_blas"].view("<i4").reshape(-1, 5)

    def static_bytes(self):
        return sum(v.nbytes for k, v in self.a.items() if k.startswith("s_"))

    def frame_bytes(self):
        return sum(v.nbytes for k, v in self.a.items() if k.startswith("f_"))
