"""GPU, two PROCESSES, ctypes only: the band exchange through the C ABI's own peer memory and device barrier
(include/b200rt.h: b200rt_peer_alloc / _open, b200rt_barrier) — no PyTorch, no NCCL anywhere in the workers.

What a C++ host (the reference is one) would do to replace the worker threads' shared frame_buffer (Window.h:56-65,
WorkerThread.cpp:53-65) with GPUs in separate processes: each process renders its interleaved tiles, the render kernel
stores every pixel into the image of the process that owns the pixel's band, a device barrier says "all pixels have
landed", each process runs the FXAA pass over its own band into rank 0's presented image, second barrier. Rank 0 must
hold the single-GPU frame, bit for bit. The two processes share GPU 0 when the box has one GPU (contexts of different
processes time-slice, so the barrier kernels do meet; slow but correct), else they take GPUs 0 and 1.
"""
import multiprocessing as mp
import os
import sys
import time

import numpy as np
import pytest

from conftest import REPO, Golden

pytestmark = pytest.mark.gpu


def _exchange_through_files(directory, tag, rank, world):
    def exchange(mine):
        path = os.path.join(directory, "%s_%d.bin" % (tag, rank))
        with open(path + ".part", "wb") as f:
            f.write(mine)
        os.replace(path + ".part", path)
        everyone = []
        for r in range(world):
            other = os.path.join(directory, "%s_%d.bin" % (tag, r))
            deadline = time.time() + 60
            while not os.path.exists(other):
                if time.time() > deadline:
                    raise RuntimeError("rank %d never published its handles" % r)
                time.sleep(0.01)
            with open(other, "rb") as f:
                everyone.append(f.read())
        return everyone
    return exchange


def _worker(rank, world, name, directory, devices):
    sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
    sys.path.insert(0, os.path.join(REPO, "tests"))
    assert "torch" not in sys.modules
    from b200rt.device import Renderer
    from b200rt import multigpu            # imports torch lazily only where a tensor view is asked for
    g = Golden(name)
    m = g.meta
    w, h = m["width"], m["height"]
    r = Renderer(w, h, m["bounces"], m["filter"], device=devices[rank], tile_stride=world, tile_phase=rank)
    r.upload_static(g.pack)
    peers = multigpu.PeerImages(r, world, rank, _exchange_through_files(directory, name, rank, world), [h * w * 4, h * w * 3])
    band_rows, ranges = multigpu.band_partition(h, world, width=w)
    y0, y1 = ranges[rank]
    r.set_band_targets(peers.ptrs[0], band_rows, multigpu.HALO)
    for _ in range(2):                                               # twice: the barrier's epochs must keep working
        r.render_device(g.pack, peers.ptrs[0][rank], wait=False)     # pixels go to the band owners over peer memory
        peers.barrier()                                              # all pixels of the frame have landed
        r.present_rows_device(peers.ptrs[0][rank], peers.ptrs[1][0], y0, y1, wait=False)   # my band, filtered into rank 0's presented image
        peers.barrier()
    held = multigpu.band_rows_needed(h, y0, y1)
    mine = peers.read(0, np.uint32, (h, w))
    np.save(os.path.join(directory, "band_%d.npy" % rank), np.stack([np.array(held), np.zeros(len(held), dtype=np.int64)]))
    np.save(os.path.join(directory, "rows_%d.npy" % rank), mine[held])
    if rank == 0:
        np.save(os.path.join(directory, "presented.npy"), peers.read(1, np.uint8, (h, w, 3)))
    peers.barrier()
    r.last_frame_stats()                                             # drains the stream before the mappings go away
    peers.close()
    r.close()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("name", ["dynamic_b3_aniso_f45"])
def test_two_processes_exchange_bands_through_the_c_abi(tmp_path, name):
    import ctypes as C
    from b200rt import _ffi
    from b200rt.device import Renderer
    count = C.c_int()
    cudart = C.CDLL("libcudart.so.12") if False else None   # the device count comes from the library itself: create on ordinal 1 and see
    devices = [0, 0]
    try:
        probe = Renderer(8, 8, 0, "tri", device=1)
        probe.close()
        devices = [0, 1]
    except Exception:
        pass
    g = Golden(name)
    m = g.meta
    whole = Renderer(m["width"], m["height"], m["bounces"], m["filter"])
    whole.upload_static(g.pack)
    expect = whole.render(g.pack)
    whole.close()

    ctx = mp.get_context("spawn")
    workers = [ctx.Process(target=_worker, args=(rank, 2, name, str(tmp_path), devices)) for rank in range(2)]
    for p in workers:
        p.start()
    for p in workers:
        p.join(240)
    for p in workers:
        if p.is_alive():
            p.kill()
            pytest.fail("a worker hung")
        assert p.exitcode == 0
    for rank in range(2):
        held = np.load(str(tmp_path / ("band_%d.npy" % rank)))[0]
        assert np.array_equal(np.load(str(tmp_path / ("rows_%d.npy" % rank))), expect["frame_buffer"][held])
    assert np.array_equal(np.load(str(tmp_path / "presented.npy")), expect["presented"])
