"""CPU, world_size 2 over gloo: the N>1 path of SURVEY 8(e) — static tile interleave + one
reduction that gathers the frame on rank 0 (b200rt/multigpu.py, the code bench.py runs over NCCL).
The per-rank renderer here is the CPU oracle standing in for a GPU; the partition, the gather and
the reassembled frame are what is under test.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200rt import multigpu
from conftest import GOLDEN, Golden


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, result_path):
    import oracle_ffi
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = Golden(name)
        m = g.meta
        part = oracle_ffi.render(g.pack, m["bounces"], m["filter"], want_ids=False, threads=2, tile_stride=world, tile_phase=rank)
        fb = torch.from_numpy(part["frame_buffer"].view(np.int32).copy())
        mask = multigpu.owner_mask(m["width"], m["height"], world, rank)
        assert not fb[~mask].any(), "a rank wrote outside its tiles"
        multigpu.FrameGather()(fb)
        rays = torch.tensor([part["stats"]["primary_rays"]], dtype=torch.int64)
        dist.all_reduce(rays)
        if rank == 0:
            np.save(result_path, fb.numpy().view(np.uint32))
            assert int(rays) == m["width"] * m["height"]
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["monkey_b1_aniso", "dynamic_b3_tri_f0"])
def test_two_ranks_reassemble_the_reference_frame(tmp_path, name):
    assert os.path.exists(os.path.join(GOLDEN, name + ".npz"))
    result = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(2, _free_port(), name, result), nprocs=2, join=True)
    assert np.array_equal(np.load(result), Golden(name).ref_frame_buffer)


def test_partition_is_a_disjoint_cover():
    for (w, h) in ((1280, 720), (1920, 1080), (3840, 2160), (33, 65)):
        tx, ty = multigpu.tile_grid(w, h)
        for world in (1, 2, 4, 8):
            seen = sorted(t for r in range(world) for t in multigpu.tiles_of_rank(w, h, world, r))
            assert seen == list(range(tx * ty))
            total = sum(int(multigpu.owner_mask(w, h, world, r).sum()) for r in range(world))
            assert total == w * h
            counts = [len(multigpu.tiles_of_rank(w, h, world, r)) for r in range(world)]
            assert max(counts) - min(counts) <= 1
    assert multigpu.tile_rect(1920, 1080, 2039) == (1888, 1056, 32, 24)   # last row is 24 px high
