"""The multi-GPU band exchange (b200rt/multigpu.py, include/b200rt.h: b200rt_set_band_targets).

CPU: the partition, who receives which row, that HALO rows are all the FXAA pass needs (checked
against the oracle's restatement of fragment_fxaa.glsl, wrap-around included), and the whole
exchange at world_size 2 over gloo with the oracle standing in for the GPUs.
GPU: N contexts on one device play the N ranks — the render kernel's banded store and the row-range
FXAA must reassemble the single-GPU frame bit for bit and touch nothing outside band + halo.
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


def test_bands_cover_the_frame_and_rows_reach_who_needs_them():
    for height in (720, 1080, 2160, 65, 16):
        for world in (2, 4, 8):
            part = multigpu.band_partition(height, world)
            if part is None:
                assert height // world < 2 * multigpu.HALO + 4
                continue
            rows, bands = part
            assert bands[0][0] == 0 and bands[-1][1] == height and rows % 4 == 0
            assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
            assert all(end - begin >= multigpu.HALO for begin, end in bands)
            received = [set() for _ in range(world)]
            for y in range(height):
                for r in multigpu.band_destinations(height, world, y):
                    received[r].add(y)
            for r, (begin, end) in enumerate(bands):
                assert received[r] == set(multigpu.band_rows_needed(height, begin, end)), (height, world, r)


@pytest.mark.parametrize("name", ["dynamic_b3_aniso_f45", "monkey_b1_aniso"])
def test_halo_rows_are_all_the_fxaa_pass_needs(name):
    import oracle_ffi
    g = Golden(name)
    fb = g.ref_frame_buffer
    height = fb.shape[0]
    whole = oracle_ffi.present(fb, enable_fxaa=True)
    rng = np.random.default_rng(7)
    for world in (2, 4, 8):
        part = multigpu.band_partition(height, world)
        if part is None:
            continue
        for begin, end in part[1]:
            held = np.zeros(height, dtype=bool)
            held[multigpu.band_rows_needed(height, begin, end)] = True
            mine = fb.copy()
            mine[~held] = rng.integers(0, 1 << 24, size=(int((~held).sum()), fb.shape[1]), dtype=np.uint32)   # what this rank never received
            assert np.array_equal(oracle_ffi.present(mine, enable_fxaa=True)[begin:end], whole[begin:end]), (world, begin)


def test_tall_narrow_frames_are_gathered_instead():
    """The diagonal taps' vertical offset is H/W pixels: beyond 4.5 the shader reaches past HALO rows."""
    assert multigpu.fxaa_reach(1920, 1080) == multigpu.HALO and multigpu.fxaa_reach(1080, 1920) == multigpu.HALO
    assert multigpu.fxaa_reach(100, 449) == multigpu.HALO and multigpu.fxaa_reach(100, 450) == 5
    assert multigpu.band_partition(1080, 2, width=1920) is not None
    assert multigpu.band_partition(4500, 2, width=1000) is None


def test_one_row_less_of_halo_is_not_enough():
    """Pins HALO to the shader's reach: with HALO - 1 rows some band of a real frame comes out different."""
    import oracle_ffi
    fb = Golden("dynamic_b3_aniso_f45").ref_frame_buffer
    height = fb.shape[0]
    whole = oracle_ffi.present(fb, enable_fxaa=True)
    different = 0
    for begin, end in multigpu.band_partition(height, 8)[1]:
        held = np.zeros(height, dtype=bool)
        held[multigpu.band_rows_needed(height, begin, end, halo=multigpu.HALO - 1)] = True
        mine = fb.copy()
        mine[~held] ^= 0x00FFFFFF
        different += int((oracle_ffi.present(mine, enable_fxaa=True)[begin:end] != whole[begin:end]).any())
    assert different > 0


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
        height = m["height"]
        part = oracle_ffi.render(g.pack, m["bounces"], m["filter"], want_ids=False, threads=2, tile_stride=world, tile_phase=rank)
        fb = torch.from_numpy(part["frame_buffer"].view(np.int32).copy())
        multigpu.BandExchangeGeneric()(fb)
        begin, end = multigpu.band_partition(height, world)[1][rank]
        held = np.zeros(height, dtype=bool)
        held[multigpu.band_rows_needed(height, begin, end)] = True
        mine = fb.numpy().view(np.uint32).copy()
        mine[~held] = 0xDEADBEEF                                     # a rank only ever holds its band + halo
        band = torch.from_numpy(oracle_ffi.present(mine, enable_fxaa=True)[begin:end].copy())
        rows = multigpu.band_partition(height, world)[0]
        padded = torch.zeros(rows, m["width"], 3, dtype=torch.uint8)
        padded[: end - begin] = band
        pieces = [torch.zeros_like(padded) for _ in range(world)] if rank == 0 else None
        dist.gather(padded, pieces, dst=0)
        if rank == 0:
            np.save(result_path, torch.cat(pieces)[:height].numpy())
    finally:
        dist.destroy_process_group()


def test_two_ranks_filter_their_bands_into_the_reference_image(tmp_path):
    import oracle_ffi
    name = "dynamic_b3_tri_f0"
    assert os.path.exists(os.path.join(GOLDEN, name + ".npz"))
    result = str(tmp_path / "presented.npy")
    mp.spawn(_worker, args=(2, _free_port(), name, result), nprocs=2, join=True)
    assert np.array_equal(np.load(result), oracle_ffi.present(Golden(name).ref_frame_buffer, enable_fxaa=True))


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_banded_store_and_row_fxaa_reassemble_the_frame(golden, world):
    from b200rt.device import DeviceError, Renderer
    m = golden.meta
    width, height = m["width"], m["height"]
    part = multigpu.band_partition(height, world)
    if part is None:
        pytest.skip("frame too short for %d bands" % world)
    rows, bands = part
    whole = Renderer(width, height, m["bounces"], m["filter"])
    whole.upload_static(golden.pack)
    expect = whole.render(golden.pack)
    whole.close()

    sentinel = -1                                                     # 0xFFFFFFFF: no pixel has it (top byte is always 0)
    images = [torch.full((height, width), sentinel, dtype=torch.int32, device="cuda") for _ in range(world)]
    presented = torch.zeros(height, width, 3, dtype=torch.uint8, device="cuda")
    ranks = []
    for r in range(world):
        ctx = Renderer(width, height, m["bounces"], m["filter"], tile_stride=world, tile_phase=r)
        ctx.upload_static(golden.pack)
        ctx.set_band_targets([t.data_ptr() for t in images], rows, multigpu.HALO)
        ctx.render_device(golden.pack, images[r].data_ptr())
        ranks.append(ctx)
    for r, (begin, end) in enumerate(bands):
        ranks[r].present_rows_device(images[r].data_ptr(), presented.data_ptr(), begin, end)
    torch.cuda.synchronize()

    for r, (begin, end) in enumerate(bands):
        got = images[r].cpu().numpy().view(np.uint32)
        held = np.zeros(height, dtype=bool)
        held[multigpu.band_rows_needed(height, begin, end)] = True
        assert np.array_equal(got[held], expect["frame_buffer"][held]), r
        assert (got[~held] == 0xFFFFFFFF).all(), "rank %d received rows outside band + halo" % r
    assert np.array_equal(presented.cpu().numpy(), expect["presented"])

    # back to a single target; refused configurations
    ranks[0].set_band_targets([], 0, 0)
    single = torch.zeros(height, width, dtype=torch.int32, device="cuda")
    ranks[0].render_device(golden.pack, single.data_ptr())
    mask = multigpu.owner_mask(width, height, world, 0).numpy()
    assert np.array_equal(single.cpu().numpy().view(np.uint32)[mask], expect["frame_buffer"][mask])
    with pytest.raises(DeviceError, match="halo"):
        ranks[0].set_band_targets([t.data_ptr() for t in images], rows, 3)
    with pytest.raises(DeviceError, match="cover"):
        ranks[0].set_band_targets([t.data_ptr() for t in images], height, multigpu.HALO)   # band 0 would be the whole frame
    for ctx in ranks:
        ctx.close()
