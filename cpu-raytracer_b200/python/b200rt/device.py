"""Renderer: Python face of libb200rt.so (include/b200rt.h) — the CUDA render path.

    r = Renderer(width, height, bounces=3, mipmap_filter="tri")
    r.upload_static(pack)                   # BLAS, materials, textures, sky -> HBM (once)
    out = r.render(pack)                    # per-frame data up, trace + FXAA, images down

There is deliberately no fallback: if the CUDA library or a GPU is missing this raises.
"""
import ctypes as C

import numpy as np

from . import _ffi


class DeviceError(RuntimeError):
    pass


class _PinnedBlock:
    """Page-locked host memory (b200rt_host_alloc) that lives as long as anything refers to it: the numpy arrays of
    alloc_outputs() keep their block alive through their base object, so closing the renderer cannot pull it away."""

    def __init__(self, lib, nbytes):
        self._lib, self.ptr = lib, C.c_void_p()
        if lib.b200rt_host_alloc(nbytes, C.byref(self.ptr)) != 0:
            raise DeviceError(lib.b200rt_last_error(None).decode())

    def __del__(self):
        if self.ptr:
            self._lib.b200rt_host_free(self.ptr)
            self.ptr = C.c_void_p()


class Renderer:
    def __init__(self, width, height, bounces, mipmap_filter="aniso", max_anisotropy=8.0, enable_fxaa=True, device=0,
                 tile_stride=1, tile_phase=0, count_traversal=False, traversal="ordered", texture_mode="mipmap",
                 arithmetic=None, deep_queue=None):
        self._lib = _ffi.device_lib()
        self.config = _ffi.make_config(width, height, bounces, mipmap_filter, max_anisotropy, enable_fxaa, device,
                                       tile_stride, tile_phase, count_traversal, traversal, texture_mode, arithmetic, deep_queue)
        self._h = C.c_void_p()
        if self._lib.b200rt_create(C.byref(self.config), C.byref(self._h)) != 0:
            raise DeviceError(self._lib.b200rt_last_error(None).decode())
        self.width, self.height = width, height
        self._keep = None

    def close(self):
        if self._h:
            self._lib.b200rt_destroy(self._h)
            self._h = C.c_void_p()       # (page-locked arrays of alloc_outputs() own their memory: _PinnedBlock)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise DeviceError(self._lib.b200rt_last_error(self._h).decode())

    def upload_static(self, pack, gpu_texture_ingest=False):
        view = pack.static_view(gpu_texture_ingest)
        self._check(self._lib.b200rt_upload_static(self._h, C.byref(view)))

    def download_texels(self, first, count):
        out = np.zeros((count, 3), dtype=np.float32)
        self._check(self._lib.b200rt_download_texels(self._h, out.ctypes.data, first, count))
        return out

    def alloc_outputs(self):
        """Page-locked output images (b200rt_host_alloc): render(..., out=...) then DMAs straight into them."""
        h, w = self.height, self.width
        res = {}
        for name, shape, dtype in (("frame_buffer", (h, w), np.uint32), ("presented", (h, w, 3), np.uint8)):
            nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
            block = _PinnedBlock(self._lib, nbytes)
            buf = (C.c_uint8 * nbytes).from_address(block.ptr.value)
            buf._block = block             # numpy array -> ctypes buffer -> block: freed when the last view goes
            res[name] = np.frombuffer(buf, dtype=dtype).reshape(shape)
        return res

    def render(self, pack, want_ids=False, want_float=False, want_presented=True, want_frame_buffer=True, out=None):
        """One frame through b200rt_render_frame (host buffers in, host buffers out)."""
        h, w = self.height, self.width
        res = out if out is not None else {}
        if want_frame_buffer and "frame_buffer" not in res:
            res["frame_buffer"] = np.zeros((h, w), dtype=np.uint32)
        if want_presented and "presented" not in res:
            res["presented"] = np.zeros((h, w, 3), dtype=np.uint8)
        if want_ids:
            res["ids"] = np.full((h, w, 2), -1, dtype=np.int32)
            res["hit"] = np.zeros((h, w, 4), dtype=np.float32)
        if want_float:
            res["colour"] = np.zeros((h, w, 3), dtype=np.float32)
        stats = _ffi.Stats()
        outputs = _ffi.Outputs(
            res["frame_buffer"].ctypes.data if want_frame_buffer else None,
            res["presented"].ctypes.data if want_presented else None,
            res["colour"].ctypes.data if want_float else None,
            res["ids"].ctypes.data if want_ids else None,
            res["hit"].ctypes.data if want_ids else None,
            C.pointer(stats))
        frame = pack.frame_view()
        self._check(self._lib.b200rt_render_frame(self._h, C.byref(frame), C.byref(outputs)))
        res["stats"] = stats.as_dict()
        return res

    def submit(self, pack, out):
        """Pipelined frame (b200rt_submit_frame): `out` from alloc_outputs(); returns at once. Pair with wait()."""
        outputs = _ffi.Outputs(out["frame_buffer"].ctypes.data if "frame_buffer" in out else None,
                               out["presented"].ctypes.data if "presented" in out else None, None, None, None, None)
        frame = pack.frame_view()
        self._check(self._lib.b200rt_submit_frame(self._h, C.byref(frame), C.byref(outputs)))

    def wait(self):
        """Blocks until the oldest submitted frame is in its output images; returns its stats."""
        stats = _ffi.Stats()
        self._check(self._lib.b200rt_wait_frame(self._h, C.byref(stats)))
        return stats.as_dict()

    def render_device(self, pack, frame_buffer_ptr, wait=True):
        """Render this context's tiles into a device buffer (uint32 [H][W]) owned by the caller.
        wait=False only enqueues on the context's stream (no statistics)."""
        frame = pack.frame_view()
        if not wait:
            self._check(self._lib.b200rt_render_frame_device(self._h, C.byref(frame), C.c_uint64(frame_buffer_ptr), None))
            return None
        stats = _ffi.Stats()
        self._check(self._lib.b200rt_render_frame_device(self._h, C.byref(frame), C.c_uint64(frame_buffer_ptr), C.byref(stats)))
        return stats.as_dict()

    def render_resident_noalloc(self, frame_buffer_ptr):
        """The resident frame once more into a caller-owned device image, through the entry point that allocates nothing
        (b200rt_render_resident_noalloc: scratch in a static device array). Only enqueues on the context's stream."""
        self._check(self._lib.b200rt_render_resident_noalloc(self._h, C.c_uint64(frame_buffer_ptr)))

    def last_frame_stats(self):
        """Counters of the frame rendered last (waits for the stream); device_ms_render = the kernel's own globaltimer span."""
        stats = _ffi.Stats()
        self._check(self._lib.b200rt_last_frame_stats(self._h, C.byref(stats)))
        return stats.as_dict()

    def present_device(self, frame_buffer_ptr, presented_ptr, wait=True):
        if not wait:
            self._check(self._lib.b200rt_present_device(self._h, C.c_uint64(frame_buffer_ptr), C.c_uint64(presented_ptr), None))
            return None
        ms = C.c_float()
        self._check(self._lib.b200rt_present_device(self._h, C.c_uint64(frame_buffer_ptr), C.c_uint64(presented_ptr), C.byref(ms)))
        return ms.value

    def set_band_targets(self, targets, band_rows, halo_rows):
        """Multi-GPU exchange fused into the pixel store (b200rt_set_band_targets): targets[b] = device address, in
        this process, of the uint32 [H][W] image of the GPU that owns band b. An empty list: single target again."""
        array = (C.c_uint64 * max(len(targets), 1))(*targets)
        self._check(self._lib.b200rt_set_band_targets(self._h, array, len(targets), band_rows, halo_rows))

    def present_rows_device(self, frame_buffer_ptr, presented_ptr, row_begin, row_end, wait=True):
        """FXAA + gamma over rows [row_begin, row_end) of a full-size image (this GPU's band)."""
        if not wait:
            self._check(self._lib.b200rt_present_rows_device(self._h, C.c_uint64(frame_buffer_ptr), C.c_uint64(presented_ptr), row_begin, row_end, None))
            return None
        ms = C.c_float()
        self._check(self._lib.b200rt_present_rows_device(self._h, C.c_uint64(frame_buffer_ptr), C.c_uint64(presented_ptr), row_begin, row_end, C.byref(ms)))
        return ms.value

    def time_resident(self, repeat=1, flush_l2_bytes=0):
        """Average device time (ms) of render and FXAA kernels over `repeat` launches on resident inputs."""
        a, b = C.c_float(), C.c_float()
        self._check(self._lib.b200rt_time_resident(self._h, repeat, C.c_uint64(flush_l2_bytes), C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_stream(self, stream):
        """Issue this context's work on a caller-owned CUDA stream (integer handle; 0 is CUDA's default stream)."""
        self._check(self._lib.b200rt_set_stream(self._h, C.c_uint64(stream)))

    def reset_stream(self):
        """Back to the context's own stream."""
        self._check(self._lib.b200rt_reset_stream(self._h))

    @property
    def stream(self):
        return self._lib.b200rt_stream(self._h)
