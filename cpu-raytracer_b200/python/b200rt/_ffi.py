"""ctypes mirrors of include/b200rt.h and include/b200rt_host.h (plain C ABI, no torch types)."""
import ctypes as C
import os

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # cpu-raytracer_b200/
REPO_ROOT = os.path.dirname(PKG_ROOT)

MAX_MIP_LEVELS = 16
FILTER_TRILINEAR, FILTER_ANISOTROPIC, FILTER_EWA = 0, 1, 2
FILTERS = {"tri": FILTER_TRILINEAR, "aniso": FILTER_ANISOTROPIC, "ewa": FILTER_EWA}
TRAVERSALS = {"ordered": 0, "naive": 1}                       # BVH_TRAVERSAL_STRATEGY (Config.h)
TEXTURE_MODES = {"mipmap": 0, "bilinear": 1, "nearest": 2}    # TEXTURE_SAMPLE_MODE (Config.h)

# sizeof of the POD records (checked against the C side by tests/test_abi.py)
SIZEOF = dict(bvh_node=32, triangle_hot=36, triangle_cold=64, blas=20, material=48, texture=88,
              instance=144, sphere=32, plane=48, light=48)


class StaticScene(C.Structure):
    _fields_ = [
        ("nodes", C.c_void_p), ("node_count", C.c_int32),
        ("triangles_hot", C.c_void_p), ("triangles_cold", C.c_void_p), ("triangle_count", C.c_int32),
        ("blas", C.c_void_p), ("blas_count", C.c_int32),
        ("materials", C.c_void_p), ("material_count", C.c_int32),
        ("textures", C.c_void_p), ("texture_count", C.c_int32),
        ("texels", C.c_void_p), ("texel_count", C.c_int64),
        ("sky", C.c_void_p), ("sky_size", C.c_int32),
        ("texels_rgb8", C.c_void_p), ("gamma_lut", C.c_void_p),
    ]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("top_left_corner", C.c_float * 3), ("x_axis", C.c_float * 3), ("y_axis", C.c_float * 3)]


class Frame(C.Structure):
    _fields_ = [
        ("camera", Camera), ("ambient", C.c_float * 3),
        ("instances", C.c_void_p), ("instance_count", C.c_int32),
        ("tlas_nodes", C.c_void_p), ("tlas_node_count", C.c_int32),
        ("tlas_indices", C.c_void_p),
        ("spheres", C.c_void_p), ("sphere_count", C.c_int32),
        ("planes", C.c_void_p), ("plane_count", C.c_int32),
        ("lights", C.c_void_p), ("light_count", C.c_int32),
    ]


class Config(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("bounces", C.c_int32), ("mipmap_filter", C.c_int32),
        ("max_anisotropy", C.c_float), ("enable_fxaa", C.c_int32), ("device", C.c_int32),
        ("tile_stride", C.c_int32), ("tile_phase", C.c_int32), ("count_traversal", C.c_int32),
        ("traversal_naive", C.c_int32), ("texture_sample_mode", C.c_int32),
        ("arithmetic", C.c_int32), ("deep_queue", C.c_int32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("reflection_rays", C.c_uint64), ("refraction_rays", C.c_uint64),
        ("nodes_visited", C.c_uint64), ("triangles_tested", C.c_uint64), ("texels_fetched", C.c_uint64),
        ("hits_accepted", C.c_uint64), ("instances_entered", C.c_uint64),
        ("device_ms_render", C.c_float), ("device_ms_fxaa", C.c_float), ("device_ms_total", C.c_float), ("device_ms_tail", C.c_float),
        ("queue_nodes", C.c_uint32), ("queue_bounce_items", C.c_uint32), ("queue_shadow_items", C.c_uint32), ("queue_fallbacks", C.c_uint32),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


class PeerHandle(C.Structure):          # b200rt_peer_handle (a cudaIpcMemHandle_t)
    _fields_ = [("bytes", C.c_ubyte * 64)]


BARRIER_BYTES = 64                      # B200RT_BARRIER_BYTES


class Outputs(C.Structure):
    _fields_ = [
        ("frame_buffer", C.c_void_p), ("presented_rgb8", C.c_void_p), ("colour_float", C.c_void_p),
        ("primary_ids", C.c_void_p), ("primary_hit", C.c_void_p), ("stats", C.POINTER(Stats)),
    ]


ARITHMETIC = {"exact": 0, "fast": 1}
DEEP_QUEUE = {"off": 0, "glass": 1, "all": 2, "auto": 3}
DEFAULT_DEEP_QUEUE = "auto"   # the context times both walks on its own first frames and keeps the faster (b200rt.h)


def make_config(width, height, bounces, mipmap_filter="aniso", max_anisotropy=8.0, enable_fxaa=True, device=0,
                tile_stride=1, tile_phase=0, count_traversal=False, traversal="ordered", texture_mode="mipmap",
                arithmetic=None, deep_queue=None):
    f = FILTERS[mipmap_filter] if isinstance(mipmap_filter, str) else int(mipmap_filter)
    arithmetic = arithmetic or os.environ.get("B200RT_ARITHMETIC", "exact")     # the environment overrides the defaults (tools/, variant timing)
    deep_queue = deep_queue or os.environ.get("B200RT_DEEP_QUEUE", DEFAULT_DEEP_QUEUE)
    return Config(width, height, bounces, f, max_anisotropy, 1 if enable_fxaa else 0, device, tile_stride, tile_phase, int(count_traversal),
                  TRAVERSALS[traversal], TEXTURE_MODES[texture_mode], ARITHMETIC[arithmetic], DEEP_QUEUE[deep_queue])


def _load(path):
    if not os.path.exists(path):
        raise OSError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` first")
    return C.CDLL(path, mode=C.RTLD_GLOBAL)


_host = None
_device = None


def host_lib():
    """libb200rt_host.so — host scene library (no CUDA)."""
    global _host
    if _host is None:
        lib = _load(os.path.join(PKG_ROOT, "libb200rt_host.so"))
        lib.b200rt_host_last_error.restype = C.c_char_p
        lib.b200rt_scene_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        lib.b200rt_scene_free.argtypes = [C.c_void_p]
        lib.b200rt_scene_free.restype = None
        lib.b200rt_scene_resize.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.b200rt_scene_resolution.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.b200rt_scene_set_camera.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.b200rt_scene_update.argtypes = [C.c_void_p, C.c_float, C.c_int]
        lib.b200rt_scene_static_view.argtypes = [C.c_void_p, C.POINTER(StaticScene)]
        lib.b200rt_scene_frame_view.argtypes = [C.c_void_p, C.POINTER(Frame)]
        lib.b200rt_scene_blas_original_indices.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.b200rt_tile_count.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _host = lib
    return _host


def device_lib():
    """libb200rt.so — the CUDA render path. There is no CPU fallback: a missing library is an error."""
    global _device
    if _device is None:
        lib = _load(os.environ.get("B200RT_LIB") or os.path.join(PKG_ROOT, "libb200rt.so"))  # B200RT_LIB: a kernel variant under test (tools/)
        lib.b200rt_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
        lib.b200rt_destroy.argtypes = [C.c_void_p]
        lib.b200rt_destroy.restype = None
        lib.b200rt_last_error.argtypes = [C.c_void_p]
        lib.b200rt_last_error.restype = C.c_char_p
        lib.b200rt_upload_static.argtypes = [C.c_void_p, C.POINTER(StaticScene)]
        lib.b200rt_render_frame.argtypes = [C.c_void_p, C.POINTER(Frame), C.POINTER(Outputs)]
        lib.b200rt_render_frame_device.argtypes = [C.c_void_p, C.POINTER(Frame), C.c_uint64, C.POINTER(Stats)]
        lib.b200rt_render_resident_noalloc.argtypes = [C.c_void_p, C.c_uint64]
        lib.b200rt_last_frame_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        lib.b200rt_peer_alloc.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(PeerHandle)]
        lib.b200rt_peer_free.argtypes = [C.c_void_p, C.c_uint64]
        lib.b200rt_peer_open.argtypes = [C.c_void_p, C.POINTER(PeerHandle), C.POINTER(C.c_uint64)]
        lib.b200rt_peer_close.argtypes = [C.c_void_p, C.c_uint64]
        lib.b200rt_peer_read.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
        lib.b200rt_barrier_setup.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int32, C.c_int32]
        lib.b200rt_barrier.argtypes = [C.c_void_p]
        lib.b200rt_present_device.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_float)]
        lib.b200rt_set_band_targets.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int32, C.c_int32, C.c_int32]
        lib.b200rt_present_rows_device.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(C.c_float)]
        lib.b200rt_time_resident.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.b200rt_download_texels.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        lib.b200rt_submit_frame.argtypes = [C.c_void_p, C.POINTER(Frame), C.POINTER(Outputs)]
        lib.b200rt_wait_frame.argtypes = [C.c_void_p, C.POINTER(Stats)]
        lib.b200rt_host_alloc.argtypes = [C.c_uint64, C.POINTER(C.c_void_p)]
        lib.b200rt_host_free.argtypes = [C.c_void_p]
        lib.b200rt_host_free.restype = None
        lib.b200rt_set_stream.argtypes = [C.c_void_p, C.c_uint64]
        lib.b200rt_reset_stream.argtypes = [C.c_void_p]
        lib.b200rt_stream.argtypes = [C.c_void_p]
        lib.b200rt_stream.restype = C.c_uint64
        _device = lib
    return _device
