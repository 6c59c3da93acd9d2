#!/usr/bin/env python3
"""bench.py — the headline benchmark of the render hot path (BASELINE.json: Mrays/s and frames/s,
Sponza 1920x1080, trilinear mip-maps, point + directional light with shadows, 3 bounces).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one frame of the hot path: trace every tile (primary rays, shadow rays, reflection /
refraction recursion, texture filtering) + the FXAA pass. Prints ONE JSON line on rank 0.

  value   whole-job Mrays/s with every input resident in HBM (kernels only, CUDA events on the
          launching stream, L2 flushed between steps), rays = active rays of the frame
  e2e     the same metric through the public per-frame calls (b200rt_submit_frame / b200rt_wait_frame; on N
          GPUs b200rt_render_frame_device + b200rt_present_rows_device per rank): host Scene::update,
          pinned-host -> device copy of the frame data, kernels, device -> host read of both images (on N
          GPUs every rank delivers its row band over its own PCIe link into one frame in shared host memory)
  roofline  render kernel: algorithmic bytes of the frame (DESIGN.md "Algorithmic bytes") / its
          CUDA-event duration, against the measured HBM copy peak (MEASURED_PEAKS.json)
  cpu_baseline  the reference's own AVX renderer (oracle/_ref, SIMD_LANE_SIZE 8, all host threads)
          on a bounded sample of the same workload

Sponza's OBJ is absent from the reference checkout (.MISSING_LARGE_BLOBS) and there is no network:
the geometry is the deterministic procedural stand-in of cpu-raytracer_b200/python/b200rt/standin.py
(seed 1234, ~262 k triangles), `data` says so.

--impl reference: times the reference's CPU implementation of the same path on the host cores
(rank 0 only), same metric / unit / config.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))

WIDTH, HEIGHT, BOUNCES, FILTER = 1920, 1080, 3, "tri"
REF_VARIANT = "l8_b3_tri"
METRIC = "Mrays/s, Sponza(stand-in) 1920x1080, trilinear mip-maps, point+directional light with shadows, 3 bounces"
WORKLOAD = "configs[1]: Sponza 1920x1080 trilinear, point + directional lights with shadows, 3 bounces (stand-in geometry, seed 1234)"
SCENE = "sponza"
DATA_NOTE = "synthetic (procedural Sponza stand-in, seed 1234; sponza.obj is missing from the reference checkout)"
L2_FLUSH_BYTES = 256 << 20

# --workload sibenik4k: BASELINE.json configs[4] (the multi-GPU scaling case of north_star), not the headline line
OTHER_WORKLOADS = {
    "sibenik4k": dict(WIDTH=3840, HEIGHT=2160, BOUNCES=3, FILTER="aniso", REF_VARIANT="l8_b3_aniso", SCENE="sibenik",
                      DATA_NOTE="synthetic (procedural Sibenik stand-in, seed 4321; sibenik.obj is missing from the reference checkout)",
                      METRIC="Mrays/s, Sibenik(stand-in) 3840x2160, anisotropic mip-maps, spot+point light with shadows, 3 bounces",
                      WORKLOAD="configs[4]: Sibenik 3840x2160, spot + point shadows, 3 bounces, tiles partitioned across the GPUs (stand-in geometry, seed 4321)"),
}


def workload_config(n_gpus):
    return {
        "workload": WORKLOAD,
        "resolution": [WIDTH, HEIGHT], "bounces": BOUNCES, "mipmap_filter": {"tri": "trilinear", "aniso": "anisotropic", "ewa": "ewa"}[FILTER], "fxaa": True,
        "tiles": "32x32, tile t -> GPU t %% %d, device-side work counter per GPU" % n_gpus,
        "l2": "flushed between timed steps (256 MiB memset, outside the timed region)",
        "ray_count": "active rays (primary + shadow + reflection + refraction), reference's PerformanceStats per ray",
    }


def load_peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    QUERY = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.device, self.samples, self.proc = device, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()

    def summary(self):
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples if len(s) >= 6 for i in range(4) if s[2 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def build_scene():
    from b200rt import standin
    from b200rt.host import HostScene
    scene_file = (standin.sponza_standin if SCENE == "sponza" else standin.sibenik_standin)(width=WIDTH, height=HEIGHT)
    scene = HostScene(scene_file)
    scene.update(0.0, 0)
    return scene_file, scene


def algorithmic_bytes(stats, pixels):
    """bytes the reference's algorithm moves for this frame (render kernel only)."""
    return (32 * stats["nodes_visited"] + 36 * stats["triangles_tested"] + (64 + 64) * stats["hits_accepted"]
            + 64 * stats["instances_entered"] + 12 * stats["texels_fetched"] + 4 * pixels)


# ------------------------------------------------------------------------- reference / CPU arm ---
def reference_frames(scene_file, frames, threads=None):
    """Runs the reference's AVX renderer (oracle/_ref) for `frames` frames; returns ms per frame list or None."""
    binary = os.path.join(REPO, "oracle", "_ref", "bin", "ref_" + REF_VARIANT)
    if not os.path.exists(binary):
        return None
    env = dict(os.environ)
    if threads:
        env["REF_THREADS"] = str(threads)
    out = subprocess.run([binary, scene_file, "--frames", str(frames), "--width", str(WIDTH), "--height", str(HEIGHT)],
                         capture_output=True, text=True, env=env, check=True).stdout
    info = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    return info


def cpu_baseline(scene_file, pack, rays, frames):
    """The reference's CPU path on this box's host cores, bounded sample (`frames` full frames after 1 warm-up)."""
    cores = os.cpu_count() or 1
    info = None
    try:
        info = reference_frames(scene_file, frames + 1)
    except (subprocess.CalledProcessError, OSError, ValueError, IndexError):
        info = None
    if info is not None:
        ms = statistics.median(info["ms_total"][1:])
        return {"value": rays / ms / 1e3, "unit": "Mrays/s", "cores": info["threads"], "kind": "reference",
                "sample": "%d full 1920x1080 frames of the same scene after 1 warm-up, median; reference AVX path (SIMD_LANE_SIZE 8) built by oracle/ref_recipe, %d threads" % (frames, info["threads"]),
                "ms_per_frame": ms, "frames_per_s": 1e3 / ms,
                "rays_reference_convention": sum(info["rays"].values())}
    # the reference binary did not travel: time the C port instead (only place bench.py may run oracle/)
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import oracle_ffi
    times = []
    for _ in range(max(2, frames // 2)):
        t = time.perf_counter()
        oracle_ffi.render(pack, BOUNCES, FILTER, want_ids=False, threads=cores)
        times.append((time.perf_counter() - t) * 1e3)
    ms = statistics.median(times[1:]) if len(times) > 1 else times[0]
    return {"value": rays / ms / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": "%d full frames, C oracle port (scalar per ray), %d threads" % (len(times), cores), "ms_per_frame": ms, "frames_per_s": 1e3 / ms}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scene_file, scene = build_scene()
    pack = scene.pack()
    # active rays of the frame, from the C oracle's counters (deterministic, = the GPU path's counters)
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import oracle_ffi
    stats = oracle_ffi.render(pack, BOUNCES, FILTER, want_ids=False)["stats"]
    rays = stats["primary_rays"] + stats["shadow_rays"] + stats["reflection_rays"] + stats["refraction_rays"]

    frames = args.warmup + args.steps
    info = None
    try:
        info = reference_frames(scene_file, frames)
    except (subprocess.CalledProcessError, OSError, ValueError, IndexError):
        info = None
    if info is not None:
        ms = info["ms_total"][args.warmup:]
        kind, cores = "reference", info["threads"]
        sample = "each step = one full 1920x1080 frame, reference AVX path (SIMD_LANE_SIZE 8), %d threads" % cores
    else:
        cores = os.cpu_count() or 1
        ms = []
        for i in range(frames):
            t = time.perf_counter()
            oracle_ffi.render(pack, BOUNCES, FILTER, want_ids=False, threads=cores)
            if i >= args.warmup:
                ms.append((time.perf_counter() - t) * 1e3)
        kind = "port"
        sample = "each step = one full 1920x1080 frame, C oracle port, %d threads" % cores
    total_ms = sum(ms)
    value = rays * len(ms) / total_ms / 1e3
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": len(ms), "warmup": args.warmup,
        "ms_per_step": total_ms / len(ms), "frames_per_s": 1e3 * len(ms) / total_ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": DATA_NOTE, "config": workload_config(args.gpus),
        "rays_per_frame": rays,
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=RESULT, flush=True)
    return 0


# ------------------------------------------------------------------------------------ GPU arm ---
def run_gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from b200rt.device import Renderer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the render path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if world > 1 and rank != 0:
        dist.barrier()                                     # rank 0 generates the stand-in assets and the BVH caches first
    scene_file, scene = build_scene()
    if world > 1 and rank == 0:
        dist.barrier()
    pack = scene.pack()
    pixels = WIDTH * HEIGHT

    # one instrumented frame (outside every timed region): ray counts + traversal counters of the WHOLE frame
    counting = Renderer(WIDTH, HEIGHT, BOUNCES, FILTER, device=local_rank, count_traversal=True)
    counting.upload_static(pack)
    full_stats = counting.render(pack, want_presented=False)["stats"]
    counting.close()
    rays = full_stats["primary_rays"] + full_stats["shadow_rays"] + full_stats["reflection_rays"] + full_stats["refraction_rays"]

    renderer = Renderer(WIDTH, HEIGHT, BOUNCES, FILTER, device=local_rank, tile_stride=world, tile_phase=rank)
    renderer.upload_static(pack)
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    clocks = ClockSampler(local_rank)

    if world == 1:
        # ---- resident: kernels only, CUDA events on the launching stream (inside the library) ----------
        first = renderer.render(pack)                      # uploads the frame once; also the e2e warm-up
        for _ in range(args.warmup):
            renderer.time_resident(1, L2_FLUSH_BYTES)
        barrier()
        with clocks:
            ms_render, ms_fxaa = renderer.time_resident(args.steps, L2_FLUSH_BYTES)
        barrier()
        ms_step = ms_render + ms_fxaa
        warm_render, warm_fxaa = renderer.time_resident(args.steps, 0)

        # ---- e2e: the frame loop a user writes — Scene::update on the host, then the frame through the public
        # per-frame call with HOST buffers in and out. Pipelined form (b200rt_submit_frame / b200rt_wait_frame):
        # frame n's images are copied out while frame n+1 is prepared and traced, as the reference's own
        # asynchronous GL present overlaps its next frame. Every step uploads its frame block and downloads
        # both of its images; the clock runs from the first submit to the arrival of the last image. --------
        outs = [renderer.alloc_outputs(), renderer.alloc_outputs()]   # page-locked host images, two frames in flight

        def frame_loop(n):
            for i in range(n):
                scene.update(0.0, 0)                       # host per-frame prep incl. TLAS rebuild (Scene.cpp:139-171)
                renderer.submit(pack.with_frame(scene.frame_view()), outs[i & 1])
                if i >= 1:
                    renderer.wait()                        # frame i-1 has arrived in outs[(i-1) & 1]
            renderer.wait()

        frame_loop(args.warmup)
        barrier()
        t0 = time.perf_counter()
        frame_loop(args.steps)
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps

        # the same loop, one frame at a time (b200rt_render_frame: nothing overlaps), for reference
        t0 = time.perf_counter()
        for _ in range(args.steps):
            scene.update(0.0, 0)
            renderer.render(pack.with_frame(scene.frame_view()), out=outs[0])
        barrier()
        e2e_sync_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        if not (np.array_equal(outs[0]["frame_buffer"], first["frame_buffer"]) and np.array_equal(outs[1]["presented"], first["presented"])):
            raise SystemExit("pipelined frames differ from the synchronous frame")
        h2d = pack.frame_bytes()
        d2h = pixels * 4 + pixels * 3
        launches = 4 * args.steps      # per frame: reset_counters_kernel, render_kernel, fxaa_kernel, tile_schedule_kernel
        kernel_ms = ms_render
        e2e_api = "Scene::update + b200rt_submit_frame/b200rt_wait_frame (two frames in flight)"
        band_d2h_ms = None
    else:
        # ---- N GPUs: each rank traces its interleaved tiles, one collective moves the frame to rank 0,
        # rank 0 runs FXAA. Timed on the device with events on the stream everything is issued on. ------
        stream = torch.cuda.Stream()                       # ONE stream for kernels, barriers / collectives and copies of this rank
        torch.cuda.set_stream(stream)
        renderer.set_stream(stream.cuda_stream)
        fb = torch.zeros(HEIGHT, WIDTH, dtype=torch.int32, device="cuda")
        presented = torch.empty(HEIGHT, WIDTH, 3, dtype=torch.uint8, device="cuda")
        host_fb = torch.empty(HEIGHT, WIDTH, dtype=torch.int32).pin_memory()
        host_rgb = torch.empty(HEIGHT, WIDTH, 3, dtype=torch.uint8).pin_memory()

        from b200rt.multigpu import HALO, BandFrames, FrameGather, PrivateHostImage, SharedHostImage, band_partition, shared_memory_free_bytes
        gather = FrameGather()
        bands = None
        partition = band_partition(HEIGHT, world, width=WIDTH)
        if os.environ.get("B200RT_GATHER", "bands") == "bands" and partition is not None:
            try:
                # two sets alternate, so no rank has to be told that the previous frame has been consumed:
                # frame n+2 is written into the images of frame n only after every rank has passed the
                # barriers of frame n+1, which it reaches after its own FXAA of frame n (stream order)
                frames = [BandFrames(HEIGHT, WIDTH, world, rank) for _ in range(2)]
                shown = [BandFrames(HEIGHT, WIDTH, world, rank, dtype=torch.uint8, channels=3) for _ in range(2)]   # rank 0's copy is the presented image
                bands = partition
            except Exception as e:                         # no symmetric memory on this box: NCCL reduction instead
                if rank == 0:
                    print("peer memory unavailable (%s): gathering with an NCCL reduce" % (str(e).splitlines() or ["?"])[0], file=sys.stderr)
        if bands is not None:
            band_rows, ranges = bands
            y0, y1 = ranges[rank]

        phase_events = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        phase_ms = [0.0, 0.0, 0.0]
        frame_number = [0]

        def step(phases=False):
            """One frame, everything resident: trace this rank's tiles, exchange, FXAA; the presented image ends on rank 0."""
            if bands is not None:
                b = frame_number[0] & 1
                frame_number[0] += 1
                renderer.set_band_targets(frames[b].ptrs, band_rows, HALO)
                if phases: phase_events[0].record(stream)
                renderer.render_device(pack, frames[b].ptrs[rank], wait=False)   # every pixel is stored into the image of its band's owner over NVLink
                if phases: phase_events[1].record(stream)
                frames[b].barrier()                        # all pixels of the frame have landed
                if phases: phase_events[2].record(stream)
                renderer.present_rows_device(frames[b].ptrs[rank], shown[b].ptrs[0], y0, y1, wait=False)   # this rank's band, filtered into rank 0's presented image
                shown[b].barrier()                         # rank 0 holds the whole presented image
                if phases: phase_events[3].record(stream)
            else:
                fb.zero_()                                 # every rank: only its own tiles may be non-zero going into the SUM
                renderer.render_device(pack, fb.data_ptr(), wait=False)
                gather(fb)                                 # tiles are disjoint and the rest is zero: SUM == gather (b200rt/multigpu.py)
                if rank == 0:
                    renderer.present_device(fb.data_ptr(), presented.data_ptr(), wait=False)

        # the exchanged frame must be the single-GPU frame, bit for bit (checked once, untimed): every rank
        # checks the band (+ halo) it received, rank 0 the presented image
        step()
        torch.cuda.synchronize()
        whole = Renderer(WIDTH, HEIGHT, BOUNCES, FILTER, device=local_rank)
        whole.upload_static(pack)
        expect = whole.render(pack)
        whole.close()
        if bands is not None:
            from b200rt.multigpu import band_rows_needed
            held = band_rows_needed(HEIGHT, y0, y1)
            ok = np.array_equal(frames[0].local.cpu().numpy().view(np.uint32)[held], expect["frame_buffer"][held])
            if rank == 0:
                ok = ok and np.array_equal(shown[0].local.cpu().numpy(), expect["presented"])
        else:
            ok = rank != 0 or (np.array_equal(fb.cpu().numpy().view(np.uint32), expect["frame_buffer"]) and np.array_equal(presented.cpu().numpy(), expect["presented"]))
        if not ok:
            print("rank %d: the exchanged frame differs from the single-GPU frame" % rank, file=sys.stderr)
        verdict = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(verdict, op=dist.ReduceOp.MIN)
        gather_verified = bool(verdict.item())

        for _ in range(args.warmup):
            step()
        barrier()
        with clocks:
            start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            total = 0.0
            for _ in range(args.steps):
                flush.fill_(1)                              # evict L2, outside the timed region
                barrier()
                if bands is not None:
                    # the ranks' streams line up on a device-side barrier (untimed): the clock of every rank starts at the same
                    # moment, with the step already queued behind it, instead of whenever its host got out of the host barrier
                    frames[frame_number[0] & 1].barrier()
                start.record(stream)
                step(phases=bands is not None)
                end.record(stream)
                torch.cuda.synchronize()
                total += start.elapsed_time(end)
                if bands is not None:
                    for k in range(3):
                        phase_ms[k] += phase_events[k].elapsed_time(phase_events[k + 1]) / args.steps
        barrier()
        t = torch.tensor([total / args.steps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_step = float(t.item())
        ms_render, ms_fxaa, warm_render, warm_fxaa = ms_step, 0.0, None, None
        kernel_ms = None

        # ---- end to end: the frame loop with host buffers, pipelined like the single-GPU loop. Every rank
        # filters its band and sends it (both images) to the host over its OWN PCIe link, into one frame in
        # shared page-locked memory; frame n leaves on a second stream while frame n+1 is traced. ----------
        if bands is not None:
            tag = "b200rt_%s_" % os.environ.get("MASTER_PORT", "0")
            # one frame for the whole job in POSIX shared memory — if /dev/shm has room for two of them (a container's
            # default is 64 MiB); else every rank keeps the band it delivered in page-locked memory of its own
            room = torch.tensor([shared_memory_free_bytes()], dtype=torch.int64, device="cuda")
            dist.all_reduce(room, op=dist.ReduceOp.MIN)
            one_frame = int(room.item()) > 2 * pixels * 7 + (64 << 20) and os.environ.get("B200RT_SHARED_FRAME", "1") != "0"
            def shared(name, shape, dtype, rows=None):
                if not one_frame:
                    return PrivateHostImage(shape, dtype)
                if rank == 0:
                    try:
                        image = SharedHostImage(tag + name, shape, dtype, create=True)
                    except FileExistsError:                  # left behind by a run that died
                        from multiprocessing import shared_memory
                        stale = shared_memory.SharedMemory(name=tag + name); stale.close(); stale.unlink()
                        image = SharedHostImage(tag + name, shape, dtype, create=True)
                    dist.barrier()
                else:
                    dist.barrier()
                    image = SharedHostImage(tag + name, shape, dtype, create=False)
                if rows is not None:
                    image.array[rows[0]:rows[1]] = 0         # first touch: this rank's rows live on this rank's NUMA node
                elif rank == 0:
                    image.array[:] = 0
                dist.barrier()
                image.register()
                return image
            host_fb = [shared("fb%d" % b, (HEIGHT, WIDTH), np.int32, (y0, y1)) for b in range(2)]
            host_rgb = [shared("rgb%d" % b, (HEIGHT, WIDTH, 3), np.uint8, (y0, y1)) for b in range(2)]
            arrived = shared("arrived", (2, world), np.int64)      # arrived[b][r] = number of the last frame whose band rank r has delivered into images b
            arrived.array[:] = 0
            dist.barrier()
            page_locked = all(h.registered and h.tensor.is_pinned() for h in host_fb + host_rgb + [arrived])
            presented_local = [torch.empty(HEIGHT, WIDTH, 3, dtype=torch.uint8, device="cuda") for _ in range(2)]
            sequence = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(2)]
            copy_stream = torch.cuda.Stream()
            fxaa_done = [torch.cuda.Event(), torch.cuda.Event()]
            copied = [torch.cuda.Event(), torch.cuda.Event()]
            delivered = [0]                                # frames of this loop and the ones before it

            def e2e_loop(n):
                first = delivered[0]
                for i in range(first, first + n):
                    b = i & 1
                    scene.update(0.0, 0)
                    frame_pack = pack.with_frame(scene.frame_view())
                    renderer.set_band_targets(frames[b].ptrs, band_rows, HALO)
                    renderer.render_device(frame_pack, frames[b].ptrs[rank], wait=False)
                    if i > first:
                        stream.wait_event(copied[b ^ 1])     # my band of frame i-1 has left images b^1: beyond this barrier the others may trace frame i+1 into them
                    frames[b].barrier()
                    renderer.present_rows_device(frames[b].ptrs[rank], presented_local[b].data_ptr(), y0, y1, wait=False)
                    fxaa_done[b].record(stream)
                    copy_stream.wait_event(fxaa_done[b])
                    with torch.cuda.stream(copy_stream):
                        host_fb[b].tensor[y0:y1].copy_(frames[b].local[y0:y1], non_blocking=True)
                        host_rgb[b].tensor[y0:y1].copy_(presented_local[b][y0:y1], non_blocking=True)
                        sequence[b].fill_(i + 1)
                        arrived.tensor[b, rank:rank + 1].copy_(sequence[b], non_blocking=True)
                        copied[b].record(copy_stream)
                    if i > first:
                        copied[b ^ 1].synchronize()          # frame i-1: this rank's band is in host memory ...
                        if rank == 0 and one_frame:          # ... and the consumer waits for everybody's
                            while int(arrived.array[b ^ 1].min()) < i:
                                pass
                last = (first + n - 1) & 1
                copied[last].synchronize()
                if rank == 0 and one_frame:
                    while int(arrived.array[last].min()) < first + n:
                        pass
                delivered[0] = first + n
                torch.cuda.synchronize()

            e2e_loop(max(2, args.warmup))
            barrier()
            t0 = time.perf_counter()
            e2e_loop(args.steps)
            barrier()
            elapsed = (time.perf_counter() - t0) * 1e3 / args.steps
            ok = True       # both host frames of the pipeline must be the single-GPU frame: every rank checks the band it delivered, rank 0 the whole frame
            for b in range(2):
                rows = slice(0, HEIGHT) if (rank == 0 and one_frame) else slice(y0, y1)
                ok = ok and np.array_equal(host_fb[b].array.view(np.uint32)[rows], expect["frame_buffer"][rows]) and np.array_equal(host_rgb[b].array[rows], expect["presented"][rows])
            verdict = torch.tensor([1 if ok else 0], device="cuda")
            dist.all_reduce(verdict, op=dist.ReduceOp.MIN)
            gather_verified = bool(gather_verified and verdict.item())
            barrier()
            # how long the band's trip to the host takes on its own, all ranks copying at once (diagnostic, untimed part of the run)
            copy_events = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
            with torch.cuda.stream(copy_stream):
                copy_events[0].record(copy_stream)
                for _ in range(10):
                    host_fb[0].tensor[y0:y1].copy_(frames[0].local[y0:y1], non_blocking=True)
                    host_rgb[0].tensor[y0:y1].copy_(presented_local[0][y0:y1], non_blocking=True)
                copy_events[1].record(copy_stream)
            torch.cuda.synchronize()
            t = torch.tensor([copy_events[0].elapsed_time(copy_events[1]) / 10], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            band_d2h_ms = float(t.item())
            barrier()
            for h in host_fb + host_rgb + [arrived]:
                h.close()
            e2e_api = ("Scene::update + b200rt_render_frame_device per rank (pixels stored into the band owners' images over NVLink) + b200rt_present_rows_device per rank + "
                       "D2H of every rank's band over its own PCIe link into %s %s host memory; frame n+1 traced while frame n leaves"
                       % ("one frame in shared" if one_frame else "its own (no shared frame: /dev/shm too small, or B200RT_SHARED_FRAME=0)", "page-locked" if page_locked else "PAGEABLE (cudaHostRegister failed)"))
        else:
            def step_e2e():
                scene.update(0.0, 0)
                fb.zero_()
                renderer.render_device(pack.with_frame(scene.frame_view()), fb.data_ptr(), wait=False)
                gather(fb)
                if rank == 0:
                    renderer.present_device(fb.data_ptr(), presented.data_ptr(), wait=False)
                    host_fb.copy_(fb, non_blocking=True)
                    host_rgb.copy_(presented, non_blocking=True)
                    torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                step_e2e()
            barrier()
            elapsed = (time.perf_counter() - t0) * 1e3 / args.steps
            e2e_api = "Scene::update + b200rt_render_frame_device per rank + NCCL reduce + b200rt_present_device + D2H on rank 0"
        t = torch.tensor([elapsed], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
        h2d = pack.frame_bytes() * world                  # every rank uploads the frame block
        d2h = pixels * 4 + pixels * 3 + (8 * world if bands is not None else 0)   # whole job: the bands add up to both images (+ the arrival stamps)
        launches = 4 * args.steps                          # per rank and frame: reset_counters, render, tile_schedule, fxaa (the NCCL path runs fxaa on rank 0 only)
        e2e_sync_ms = None
        if bands is None:
            band_d2h_ms = None

    if rank == 0:
        peak, peak_source = load_peaks()
        alg_bytes = algorithmic_bytes(full_stats, pixels)
        roofline = None
        if kernel_ms:
            achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
            ncu = {}
            try:
                with open(os.path.join(REPO, "profiles", "ncu_render_kernel.json")) as f:
                    ncu = json.load(f)
            except (OSError, ValueError):
                pass
            if ncu.get("workload", "sponza1080") != args.workload:
                ncu = {}   # the capture describes another frame: no traffic / instruction figures for this one
            sm_hz = (clocks.summary()["sm_mhz"] or 1965.0) * 1e6
            issue = None
            if ncu.get("thread_instructions_per_launch"):
                # second yardstick: lane-instructions the launch executes vs what 148 SMs x 4 schedulers x 32 lanes can issue
                issue_peak = 148 * 4 * 32 * sm_hz
                issue = {"thread_instructions_per_launch": ncu["thread_instructions_per_launch"], "achieved_per_s": ncu["thread_instructions_per_launch"] / (kernel_ms * 1e-3),
                         "peak_per_s": issue_peak, "frac": ncu["thread_instructions_per_launch"] / (kernel_ms * 1e-3) / issue_peak,
                         "threads_per_instruction": ncu.get("threads_per_instruction"), "issue_active_pct": ncu.get("issue_active_pct")}
            roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu.get("dram_bytes_per_launch"),
                        "kernel": "render_kernel<false>", "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes,
                        "algorithmic_bytes_per_ray": alg_bytes / rays, "peak_source": peak_source, "traffic_source": ncu.get("source"),
                        "issue_roofline": issue,
                        "note": "the algorithmic stream (BVH nodes + triangles, 2 KB per ray) is served by L1/L2 (working set 39 MB < 126 MB L2): DRAM traffic per launch is `traffic`, ~1 % of the algorithmic bytes, so frac > 1 against the HBM peak means 'cache-resident', not 'beyond the roofline'; the binding limit is instruction issue (issue_roofline)"}
        baseline = cpu_baseline(scene_file, pack, rays, frames=6) if world == 1 else None
        line = {
            "metric": METRIC, "value": rays / ms_step / 1e3, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "frames_per_s": 1e3 / ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": DATA_NOTE,
            "config": workload_config(world),
            "rays_per_frame": rays, "rays": {k: full_stats[k] for k in ("primary_rays", "shadow_rays", "reflection_rays", "refraction_rays")},
            "ms_render_kernel": ms_render, "ms_fxaa_kernel": ms_fxaa,
            "warm_l2": None if warm_render is None else {"ms_render_kernel": warm_render, "ms_fxaa_kernel": warm_fxaa, "value": rays / (warm_render + warm_fxaa) / 1e3},
            "e2e": {"value": rays / e2e_ms / 1e3, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "frames_per_s": 1e3 / e2e_ms,
                    "api": e2e_api, "unpipelined_ms_per_step": e2e_sync_ms, "band_d2h_ms_alone": band_d2h_ms},
            "gpu_launches": launches,
            "gather_verified": gather_verified if world > 1 else None,
            "rank0_phases_ms": None if world == 1 or bands is None else {"upload_and_render": phase_ms[0], "barrier_all_pixels_landed": phase_ms[1], "fxaa_band_into_rank0_and_barrier": phase_ms[2]},
            "gather": None if world == 1 else ("%d bands of %d rows (+%d halo): render kernel stores each pixel into its band owner's image over NVLink peer memory, device barrier, every rank filters its band"
                                              % (world, bands[0], HALO) if bands is not None else "NCCL reduce(SUM) of disjoint tiles"),
            "clocks": clocks.summary(),
            "roofline": roofline,
            "cpu_baseline": baseline,
        }
        print(json.dumps(line), file=RESULT, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


RESULT = sys.stdout


def keep_stdout_for_the_result():
    """The contract is ONE JSON line on stdout; libraries write there too (NCCL prints its version on
    the first collective). File descriptor 1 is pointed at stderr and the line goes to a private copy."""
    global RESULT
    sys.stdout.flush()
    RESULT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def main():
    keep_stdout_for_the_result()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sponza1080", choices=["sponza1080"] + sorted(OTHER_WORKLOADS))
    args = ap.parse_args()
    if args.workload != "sponza1080":
        globals().update(OTHER_WORKLOADS[args.workload])
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
