"""Tile partition across ranks and the one collective of the path.

The reference hands 32x32 tiles to worker threads off one atomic counter (WorkerThread.cpp:53-65).
Across GPUs the hand-out is static: tile t (row-major, Window.cpp:11-12) belongs to rank
t % world — an interleave that balances sky rows against geometry rows — and inside a GPU the
device-side counter of the render kernel takes over. Scene data is replicated; the only exchange
is the pre-FXAA frame buffer travelling to rank 0, which then runs the FXAA pass over the full
image (its taps reach across tile borders and wrap at the frame border).

Every rank renders into a zero-initialised uint32 [H][W] buffer and touches only its own tiles, so
a SUM reduction to rank 0 *is* the gather (x + 0 = x, bit for bit), with no packing kernels on
either side. Works with any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).

PyTorch is imported only by the pieces that wrap a torch collective or hand out a tensor view; the band exchange
itself (PeerImages: peer memory and the device barrier of include/b200rt.h) runs through the C ABI alone.
"""
TILE = 32  # Window.h:32-33


def tile_grid(width, height):
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


def tiles_of_rank(width, height, world, rank):
    """Row-major ids of the tiles rank `rank` renders."""
    tx, ty = tile_grid(width, height)
    return list(range(rank, tx * ty, world))


def tile_rect(width, height, tile):
    """(x, y, w, h) of a tile, clipped at the frame edge like WorkerThread.cpp:57-61."""
    tx, _ = tile_grid(width, height)
    x, y = (tile % tx) * TILE, (tile // tx) * TILE
    return x, y, min(TILE, width - x), min(TILE, height - y)


def owner_mask(width, height, world, rank):
    """bool [H][W]: pixels owned by `rank`."""
    import torch
    tx, ty = tile_grid(width, height)
    ids = (torch.arange(ty).unsqueeze(1) * tx + torch.arange(tx).unsqueeze(0)) % world == rank
    return ids.repeat_interleave(TILE, 0).repeat_interleave(TILE, 1)[:height, :width]


class FrameGather:
    """Moves the disjoint per-rank tiles of one frame to rank 0 with a single reduction."""

    def __init__(self, group=None, dst=0):
        self.group, self.dst = group, dst

    def __call__(self, frame_buffer):
        """frame_buffer: int32 tensor [H][W], zero outside this rank's tiles. In place on dst."""
        import torch.distributed as dist
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.reduce(frame_buffer, dst=self.dst, op=dist.ReduceOp.SUM, group=self.group)
        return frame_buffer


# ------------------------------------------------------------------------------ bands ---
# The exchange bench.py uses when peer memory is available. Gathering the frame on rank 0 leaves
# rank 0 with everything that follows tracing: the FXAA pass over the whole image and both images'
# trip to the host over ONE PCIe link, while the other GPUs wait. Instead the frame is cut into N
# horizontal bands; every rank's render kernel stores each pixel straight into the image of the
# rank that owns the pixel's band (b200rt_set_band_targets: NVLink peer stores, 1/N of the frame
# per link, plus HALO rows either side of a band so that the FXAA taps find their neighbours), and
# every rank then filters its own band and sends it to the host over its own PCIe link.

HALO = 4  # rows fragment_fxaa.glsl reaches above / below a pixel: |dir| <= 8 px (:46), taps at +-1/2 of it (:52-57), diagonals < 1 px


def fxaa_reach(width, height):
    """Rows the shader reaches above / below a pixel: 4 along the edge direction, and the diagonal taps'
    vertical offset, which is H/W pixels (the shader adds inv_screen_size.x to both coordinates,
    fragment_fxaa.glsl:25-28): more than 4 only for a frame over 4.5 times as high as wide."""
    return max(HALO, int(0.5 + height / width))


def band_partition(height, world, multiple=4, width=None):
    """(band_rows, [(row_begin, row_end)] per rank): equal bands, a multiple of the FXAA kernel's 4-row
    blocks; the last band takes the remainder. None when the frame is too short to give every rank a
    band of at least HALO rows, or so narrow that the shader reaches further than HALO rows (then:
    gather on rank 0)."""
    if width is not None and fxaa_reach(width, height) > HALO:
        return None
    rows = -(-height // world)
    rows = -(-rows // multiple) * multiple
    if rows < HALO or (world - 1) * rows >= height or height - (world - 1) * rows < HALO:
        return None
    return rows, [(r * rows, min(height, (r + 1) * rows)) for r in range(world)]


def band_rows_needed(height, begin, end, halo=HALO):
    """Row indices (cyclic: the taps wrap at the frame border) a rank must hold to filter rows [begin, end)."""
    return [y % height for y in range(begin - halo, end + halo)]


def band_destinations(height, world, y, halo=HALO):
    """Ranks that receive the pixels of row y — what store_pixel_banded (csrc/kernels.cuh) does."""
    rows, bands = band_partition(height, world)
    b = min(y // rows, world - 1)
    begin, end = bands[b]
    dst = [b]
    if y - begin < halo:
        dst.append((b - 1) % world)
    if end - y <= halo:
        dst.append((b + 1) % world)
    return sorted(set(dst))


class BandExchangeGeneric:
    """The band exchange over any torch.distributed backend (gloo in the CPU tests): every rank ends up
    with the rows of its band + halo, taken from whichever rank traced them. One all_reduce(SUM) of
    images that are zero outside the sender's tiles — same trick as FrameGather; the GPU path moves
    only the rows each rank needs (PeerImages)."""

    def __init__(self, group=None):
        self.group = group

    def __call__(self, frame_buffer):
        import torch.distributed as dist
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(frame_buffer, op=dist.ReduceOp.SUM, group=self.group)
        return frame_buffer


class PeerImages:
    """Device buffers of every rank that every rank of the node has mapped (include/b200rt.h: b200rt_peer_alloc /
    b200rt_peer_open — cudaIpc handles, NVLink peer access), plus the ranks' flag blocks for the device-side barrier
    (b200rt_barrier_setup / b200rt_barrier). No framework is involved on the data path: the 64-byte handles are the
    only thing the processes exchange, once, at set-up — through `exchange`, a callable that takes this rank's bytes
    and returns every rank's (bench.py passes an all_gather over its process group, tests/test_peer_exchange.py uses
    files; a C++ host would use a pipe).

    sizes[k] = bytes of buffer k (every rank allocates the same list); ptrs[k][r] = rank r's buffer k as mapped in this
    process — for an image, the list b200rt_set_band_targets wants."""

    def __init__(self, renderer, world, rank, exchange, sizes):
        import ctypes as C
        from . import _ffi
        self.renderer, self.world, self.rank, self.sizes = renderer, world, rank, list(sizes)
        self._lib, self._h = renderer._lib, renderer._h
        mine, handles = [], []
        for size in self.sizes + [_ffi.BARRIER_BYTES]:    # the last allocation is the barrier's flag block
            ptr, handle = C.c_uint64(), _ffi.PeerHandle()
            renderer._check(self._lib.b200rt_peer_alloc(self._h, size, C.byref(ptr), C.byref(handle)))
            mine.append(ptr.value)
            handles.append(bytes(handle.bytes))
        self._mine = mine
        everyone = exchange(b"".join(handles))            # [world] x (len(sizes) + 1) x 64 bytes
        self._opened = []
        table = []
        for r in range(world):
            row = []
            for k in range(len(mine)):
                if r == rank:
                    row.append(mine[k])
                    continue
                handle = _ffi.PeerHandle.from_buffer_copy(everyone[r][64 * k:64 * (k + 1)])
                ptr = C.c_uint64()
                renderer._check(self._lib.b200rt_peer_open(self._h, C.byref(handle), C.byref(ptr)))
                self._opened.append(ptr.value)
                row.append(ptr.value)
            table.append(row)
        self.ptrs = [[table[r][k] for r in range(world)] for k in range(len(self.sizes))]
        flags = (C.c_uint64 * world)(*[table[r][len(self.sizes)] for r in range(world)])
        renderer._check(self._lib.b200rt_barrier_setup(self._h, flags, world, rank))

    def barrier(self):
        """Device-side barrier of all ranks, enqueued on the renderer's stream (one tiny kernel, no host involvement)."""
        self.renderer._check(self._lib.b200rt_barrier(self._h))

    def read(self, k, dtype, shape):
        """This rank's buffer k copied to the host (b200rt_peer_read: waits for the renderer's stream)."""
        import numpy as np
        out = np.zeros(shape, dtype=dtype)
        assert out.nbytes <= self.sizes[k]
        self.renderer._check(self._lib.b200rt_peer_read(self._h, self._mine[k], out.ctypes.data, out.nbytes))
        return out

    def tensor(self, k, shape, itemsize):
        """This rank's buffer k as a torch tensor over the same device memory (int32 / uint8): a view, not a copy."""
        import torch

        class _Span:
            pass
        span = _Span()
        span.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<i4" if itemsize == 4 else "|u1", "data": (self._mine[k], False), "version": 2}
        return torch.as_tensor(span, device="cuda")

    def close(self):
        for p in self._opened:
            self._lib.b200rt_peer_close(self._h, p)
        for p in self._mine:
            self._lib.b200rt_peer_free(self._h, p)
        self._opened, self._mine = [], []


class PrivateHostImage:
    """Same face as SharedHostImage, but page-locked memory of this process only: where /dev/shm is too
    small for the frame, every rank keeps the band it delivered (a distributed frame)."""

    def __init__(self, shape, dtype):
        import numpy as np
        import torch
        self.tensor = torch.empty(tuple(shape), dtype=torch.from_numpy(np.zeros(1, dtype=dtype)).dtype).pin_memory()
        self.array = self.tensor.numpy()
        self.registered = True

    def register(self):
        return True

    def close(self):
        pass


def shared_memory_free_bytes(path="/dev/shm"):
    import os
    try:
        st = os.statvfs(path)
        return st.f_bavail * st.f_frsize
    except OSError:
        return 0


class SharedHostImage:
    """A host image all ranks of the node write their bands into: POSIX shared memory, page-locked in
    every process (cudaHostRegister), so each GPU's DMA engine delivers its band over its own PCIe
    link and the consumer (rank 0) reads one contiguous frame."""

    def __init__(self, name, shape, dtype, create):
        import numpy as np
        import torch
        from multiprocessing import shared_memory
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.shm = shared_memory.SharedMemory(name=name, create=create, size=nbytes)
        self.owner = create
        if not create:   # Python < 3.13 registers attached segments for clean-up too: only the creator unlinks
            try:
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.array = np.ndarray(shape, dtype=dtype, buffer=self.shm.buf)
        self.tensor = torch.from_numpy(self.array)
        self.nbytes = nbytes
        self.registered = False

    def register(self):
        """Page-locks the mapping in this process. Call after every rank has written its own rows once:
        registration faults all pages in, and a page lives on the NUMA node of whoever touches it first."""
        import torch
        rc = torch.cuda.cudart().cudaHostRegister(self.tensor.data_ptr(), self.nbytes, 0)
        self.registered = int(rc) == 0
        return self.registered

    def close(self):
        if self.registered:
            import torch
            torch.cuda.cudart().cudaHostUnregister(self.tensor.data_ptr())
            self.registered = False
        del self.tensor, self.array
        try:
            self.shm.close()
        except BufferError:     # a view of the image is still alive somewhere: the mapping goes with the process
            pass
        if self.owner:
            try:
                self.shm.unlink()
            except FileNotFoundError:
                pass
