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


class PeerFrame:
    """Rank 0's frame buffer mapped into every rank's address space (NVLink peer memory).

    The render kernel of every rank stores its pixels straight into rank 0's image — the gather is
    fused into the kernel's final store (4 bytes per pixel over NVLink, 8.3 MB per 1080p frame in
    total), so there is no zero-fill, no send buffer and no reduction; one device-side barrier tells
    rank 0 that all tiles have landed. Needs CUDA symmetric memory (torch.distributed._symmetric_memory);
    FrameGather is the transport-agnostic form of the same exchange.
    """

    def __init__(self, height, width, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.local = symm_mem.empty((height, width), dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
        self.local.zero_()
        self.handle = symm_mem.rendezvous(self.local, self.group)
        self.root_ptr = int(self.handle.buffer_ptrs[0])   # rank 0's buffer, as seen from this rank

    def barrier(self):
        """Device-side barrier of all ranks, enqueued on the current stream."""
        self.handle.barrier()
