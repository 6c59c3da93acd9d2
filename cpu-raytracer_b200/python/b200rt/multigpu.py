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
"""
import torch
import torch.distributed as dist

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
    tx, ty = tile_grid(width, height)
    ids = (torch.arange(ty).unsqueeze(1) * tx + torch.arange(tx).unsqueeze(0)) % world == rank
    return ids.repeat_interleave(TILE, 0).repeat_interleave(TILE, 1)[:height, :width]


class FrameGather:
    """Moves the disjoint per-rank tiles of one frame to rank 0 with a single reduction."""

    def __init__(self, group=None, dst=0):
        self.group, self.dst = group, dst

    def __call__(self, frame_buffer):
        """frame_buffer: int32 tensor [H][W], zero outside this rank's tiles. In place on dst."""
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
    only the rows each rank needs (BandFrames)."""

    def __init__(self, group=None):
        self.group = group

    def __call__(self, frame_buffer):
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(frame_buffer, op=dist.ReduceOp.SUM, group=self.group)
        return frame_buffer


class BandFrames:
    """One full-size pre-FXAA image per rank in CUDA symmetric memory, each mapped into every rank.

    ptrs[r] = rank r's image as seen from this process: the list b200rt_set_band_targets wants. Only
    the rows of a rank's band (+ halo) are ever written in its image."""

    def __init__(self, height, width, world, rank, group=None, dtype=torch.int32, channels=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        shape = (height, width) if channels is None else (height, width, channels)
        self.local = symm_mem.empty(shape, dtype=dtype, device=torch.device("cuda", torch.cuda.current_device()))
        self.local.zero_()
        self.handle = symm_mem.rendezvous(self.local, self.group)
        self.ptrs = [int(self.handle.buffer_ptrs[r]) for r in range(world)]
        self.rank = rank

    def barrier(self):
        """Device-side barrier of all ranks, enqueued on the current stream."""
        self.handle.barrier()


class PrivateHostImage:
    """Same face as SharedHostImage, but page-locked memory of this process only: where /dev/shm is too
    small for the frame, every rank keeps the band it delivered (a distributed frame)."""

    def __init__(self, shape, dtype):
        import numpy as np
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
        rc = torch.cuda.cudart().cudaHostRegister(self.tensor.data_ptr(), self.nbytes, 0)
        self.registered = int(rc) == 0
        return self.registered

    def close(self):
        if self.registered:
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
