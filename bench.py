#!/usr/bin/env python
"""bench.py -- BASELINE.json metric on BASELINE config C2 (configs[1]): primary-ray closest-hit,
1920x1080, 1 spp, procedural 1 002 528-triangle height-field, user program programs/primary.cl
through RD::TraceRays. One "step" = one TraceRays frame (W*H primary rays).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU): the scene is replicated, the frame is split into
contiguous runs of 8x4 screen tiles (one run per rank), and the RGBA8 framebuffer is gathered on
rank 0 with NCCL (the only collective). Scaling is "strong" (the frame is fixed).
Prints ONE JSON line on rank 0. See DESIGN.md "Measurement" for every field's definition.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

WIDTH, HEIGHT, GRID_N = 1920, 1080, 708
WORKLOAD = "C2: primary-ray closest-hit, 1920x1080, 1 spp, height-field n=708 (1002528 triangles), programs/primary.cl"
NODE_BYTES, TRI_BYTES, RAY_BYTES, HIT_BYTES = 80, 48, 32, 16
# dram__bytes_read.sum + dram__bytes_write.sum of one raygen launch of this workload, from the ncu --set full capture
# summarised in profiles/ncu_r1c_primary.txt (93.44 MB + 29.40 MB); re-measured whenever the kernel changes
NCU_DRAM_BYTES_PER_LAUNCH = 124.57e6   # profiles/ncu_r1g_primary.txt: dram read 93.56 MB + write 31.01 MB


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def sample_rows(ref, host, n, cores):
    """Bounded sample of the frame for the CPU legs: image rows around the middle of the frame, sized
    from a 60-row probe so that one pass costs about 1.5 s of wall time (at most the whole frame)."""
    first, count = (HEIGHT // 2 - 30) * WIDTH, 60 * WIDTH
    t0 = time.perf_counter()
    ref.raygen("primary", host, n, first, count, cores)
    dt = max(1e-4, time.perf_counter() - t0)
    rows = int(60 * 1.5 / dt)
    return max(60, min(HEIGHT, rows - rows % 4))


_REAL_STDOUT = None


def claim_stdout():
    """This process's stdout carries ONE JSON line. Libraries print there too -- the reference's builder ("Max BVH depth
    ...", radiance/src/bvh.cpp), NCCL ("NCCL version ...") -- so file descriptor 1 is pointed at stderr for the whole run
    and the line is written to a private duplicate of the original stdout by emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(line + "\n")
    _REAL_STDOUT.flush()


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU code (oracle/_ref: its bvh.cpp + radiance.cl/pbr.cl, the
    same user program compiled as C++) on the host cores, each step a bounded sample of the C2 frame."""
    if rank != 0:
        return None
    pkg, ref = entry.load_package(), entry.load_oracle()
    if not ref.available():
        return json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libradiance_ref.so not built"})
    scenes = pkg.scenes
    cores = os.cpu_count() or 1
    b = scenes.heightfield_scene(GRID_N, WIDTH, HEIGHT)
    t0 = time.time()
    tl = scenes.build_reference_scene(ref, b)
    build_s = time.time() - t0
    n = WIDTH * HEIGHT
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    rt = scenes.rtprop(0, 1, 1, 0)
    host = [np.array(rt).reshape(1), scratch, image, np.array(b.camera).reshape(1), np.array(b.scene).reshape(1),
            b.meshInfoData, b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, None, 0, tl.blob]
    rows = sample_rows(ref, host, n, cores)
    first, count = (HEIGHT // 2 - rows // 2) * WIDTH, rows * WIDTH
    for _ in range(args.warmup):
        ref.raygen("primary", host, n, first, count, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.raygen("primary", host, n, first, count, cores)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    v = count / dt / 1e6
    sample = "%d of %d image rows (%d rays) per step, reference BVH build %.1f s outside the timed region" % (rows, HEIGHT, count, build_s)
    return json.dumps({"impl": "reference", "metric": "Mrays/s (primary closest-hit, 1080p)", "value": v, "unit": "Mrays/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": WORKLOAD, "sample": sample},
                      "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample},
                      "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


def cpu_baseline_leg(pkg, ref, bundle):
    """Reference CPU path on a bounded sample (rank 0, N=1 only): see reference_arm for the sample."""
    scenes = pkg.scenes
    cores = os.cpu_count() or 1
    t0 = time.time()
    tl = scenes.build_reference_scene(ref, bundle)
    build_s = time.time() - t0
    n = WIDTH * HEIGHT
    host = [np.array(scenes.rtprop(0, 1, 1, 0)).reshape(1), np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8),
            np.array(bundle.camera).reshape(1), np.array(bundle.scene).reshape(1), bundle.meshInfoData, bundle.vertexData,
            bundle.indexData, bundle.uvData, bundle.normalData, bundle.materials, None, 0, tl.blob]
    rows = sample_rows(ref, host, n, cores)
    first, count = (HEIGHT // 2 - rows // 2) * WIDTH, rows * WIDTH
    ref.raygen("primary", host, n, first, count, cores)   # warm-up
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        ref.raygen("primary", host, n, first, count, cores)
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
    return {"value": count / best / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "reference",
            "sample": "%d of %d image rows (%d rays), best of 3, all host threads; reference SAH build took %.1f s (not timed)"
                      % (rows, HEIGHT, count, build_s)}


def time_frames(torch, plt, rd, stream, flush, w, h, reps=3):
    """Median device time (ms) of rd_trace_rays over `reps` frames after one warm-up, L2 flushed before each."""
    ms = []
    for it in range(reps + 1):
        with torch.cuda.stream(stream):
            flush.fill_(it)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rd.TraceRaysAsync(plt, w, h)
            e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms[1:]))


def extra_config(torch, pkg, plt, stream, flush, name, bundle, program, spp, depth, modes, width=None, height=None):
    """Path-tracing configs (BASELINE configs[2], configs[3]) through RD::TraceRays: rays per frame counted by a
    count_rays build of the same program, then each mode timed. Returns a dict for the JSON line."""
    rd, scenes = pkg.rd, pkg.scenes
    w, h = width or WIDTH, height or HEIGHT
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    build_ms = (time.perf_counter() - t0) * 1e3
    ds = scenes.DeviceScene(plt, bundle, tlas, w, h, scenes.rtprop(0, spp, depth, 0))
    plt.set_option("count_rays", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, program)
    finally:
        plt.set_option("count_rays", 0)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    plt.ray_count()
    rd.TraceRays(plt, 0, 0, 0, w, h)
    rays = plt.ray_count()
    pipe, module = scenes.make_pipeline(plt, program)
    rd.BindPipeline(plt, pipe)
    out = {"workload": name, "triangles": bundle.triangle_count(), "instances": len(bundle.instances), "spp": spp, "depth": depth,
           "rays_per_frame": rays, "build_ms_incl_upload": build_ms, "unit": "Mrays/s"}
    try:
        for mode, wf, srt, paths in modes:
            plt.set_option("wavefront", wf)
            plt.set_option("wf_sort", srt)
            plt.set_option("wf_paths", paths)
            l0 = plt.launch_count()
            ms = time_frames(torch, plt, rd, stream, flush, w, h, reps=3 if w * h * spp < (1 << 28) else 2)
            out[mode] = {"value": rays / ms / 1e3, "ms_per_frame": ms, "launches_per_frame": (plt.launch_count() - l0) // 4}
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_sort", 1)
        plt.set_option("wf_paths", 1 << 25)
    for h in [tlas] + list(blas):
        h.free()
    return out


def multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world, name, bundle, w, h, spp, depth, frames=3):
    """Path tracing across all ranks with dynamic chunk stealing: frame ms (max over ranks, device events, gather
    included) and Mrays/s with the rays of one frame counted on rank 0 by a count_rays build."""
    rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
    tlas, blas = scenes.build_scene(plt, bundle)
    ds = scenes.DeviceScene(plt, bundle, tlas, w, h, scenes.rtprop(0, spp, depth, 0))
    n = w * h
    image_t = torch.zeros(n * 4, dtype=torch.uint8, device="cuda")
    ds.rdImage = rd.WrapBuffer(plt, image_t.data_ptr(), n * 4, owner=image_t)
    ds.descriptors[2] = ds.rdImage
    rays = 0
    if rank == 0:
        plt.set_option("count_rays", 1)
        try:
            pipe, _ = scenes.make_pipeline(plt, "path")
        finally:
            plt.set_option("count_rays", 0)
        rd.BindPipeline(plt, pipe)
        rd.BindDescriptorSet(plt, ds.descriptors)
        plt.ray_count()
        rd.TraceRays(plt, 0, 0, 0, w, h)
        rays = plt.ray_count()
    plt.set_option("stealing", 1)
    mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)

    def timed(render):
        ms = []
        for f in range(frames + 1):
            torch.cuda.synchronize()
            dist.barrier()
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                image_t.zero_()
                render(f)
                mg.combine_sum(image_t, world, dist)
                e1.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms.append(float(t[0]))
        return float(np.median(ms[1:]))

    steal_ms = timed(lambda f: rd.TraceRaysSteal(plt, w, h, 600 + f, 32))
    owned = torch.tensor([len(rd.StealOwned(plt))], device="cuda")
    allc = [torch.zeros_like(owned) for _ in range(world)]
    dist.all_gather(allc, owned)
    # wavefront stages on the interleaved band partition
    plt.set_option("wavefront", 1)
    try:
        def wave(f):
            ds.set_rtprop(0, spp, depth)
            rd.TraceRaysInterleaved(plt, w, h, rank, world, 1)
        wave_ms = timed(wave)
    finally:
        plt.set_option("wavefront", 0)
    for hnd in [tlas] + list(blas):
        hnd.free()
    best = min(steal_ms, wave_ms)
    return {"workload": name, "triangles": bundle.triangle_count(), "spp": spp, "depth": depth, "rays_per_frame": rays,
            "frame_ms": best, "value": rays / best / 1e3 if rays else None, "unit": "Mrays/s",
            "inline_stealing": {"frame_ms": steal_ms, "partition": "dynamic chunk stealing + NCCL sum-reduce",
                                "chunks_per_rank": [int(c) for c in allc]},
            "wavefront_interleaved": {"frame_ms": wave_ms, "partition": "interleaved tile-row bands + NCCL sum-reduce"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--partition", default="interleave", choices=["interleave", "steal", "static"],
                    help="N>1: interleaved tile-row bands + sum-reduce (default), dynamic chunk stealing + sum-reduce, "
                         "or contiguous tile rows + send/recv gather")
    ap.add_argument("--band", type=int, default=1, help="tile rows per band of the interleaved partition")
    ap.add_argument("--chunk", type=int, default=32, help="work-items per stolen chunk (multiple of 32)")
    ap.add_argument("--extra", default="c3", choices=["c3", "all"], help="extra configs measured at N=1: C3, or C3 and C4")
    ap.add_argument("--readback-bands", type=int, default=24,
                    help="N=1 end-to-end leg: bands of the streamed read-back (rd_buffer_stream_to_host); 0 = plain ReadBuffer")
    ap.add_argument("--no-extra", action="store_true", help="skip the C3 / C4 wavefront measurements (N=1 only)")
    args = ap.parse_args()
    claim_stdout()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        line = reference_arm(args, rank, world)
        if line:
            emit(line)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = entry.load_package()
    rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
    plt = rd.Platform.GetPlatform(local)
    steal = world > 1 and args.partition == "steal"
    interleave = world > 1 and args.partition == "interleave"
    if steal:
        plt.set_option("stealing", 1)
    stream = torch.cuda.ExternalStream(plt.stream(), device=local)

    bundle = scenes.heightfield_scene(GRID_N, WIDTH, HEIGHT)
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    build_ms = (time.perf_counter() - t0) * 1e3            # first build of the process: includes module loading
    for h in [tlas] + list(blas):
        h.free()
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    build_warm_ms = (time.perf_counter() - t0) * 1e3
    ds = scenes.DeviceScene(plt, bundle, tlas, WIDTH, HEIGHT, scenes.rtprop(0, 1, 1, 0))
    # the RGBA8 framebuffer lives in a torch tensor so NCCL can gather it in place (zero copy)
    n = WIDTH * HEIGHT
    image_t = torch.zeros(n * 4, dtype=torch.uint8, device="cuda")
    ds.rdImage = rd.WrapBuffer(plt, image_t.data_ptr(), n * 4, owner=image_t)
    ds.descriptors[2] = ds.rdImage
    streamed = world == 1 and args.readback_bands > 0
    if streamed:   # the module also gets the band-reporting raygen; it only runs once a host mirror is registered (e2e leg)
        plt.set_option("stream_image", 1)
    pipe, module = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    first, count, row0, rows = mg.tile_row_range(rank, world, WIDTH, HEIGHT)
    if steal:
        mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")
        first, count = 0, n
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    frame_no = [0]

    def render_and_gather():
        """One frame of this rank's share + the frame's only collective; returns the event between them."""
        e1 = torch.cuda.Event(enable_timing=True)
        if steal:
            image_t.zero_()
            rd.TraceRaysSteal(plt, WIDTH, HEIGHT, frame_no[0], args.chunk)
            frame_no[0] += 1
            e1.record(stream)
            mg.combine_sum(image_t, world, dist)
        elif interleave:
            image_t.zero_()
            rd.TraceRaysInterleaved(plt, WIDTH, HEIGHT, rank, world, args.band)
            e1.record(stream)
            mg.combine_sum(image_t, world, dist)
        else:
            rd.TraceRaysRange(plt, WIDTH, HEIGHT, first, count)
            e1.record(stream)
            mg.gather_rows(image_t, rank, world, WIDTH, HEIGHT, dist)
        return e1

    def step(timed):
        with torch.cuda.stream(stream):
            flush.fill_(rank + 1)                       # evict the scene from L2 between frames
            e0, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record(stream)
            e1 = render_and_gather()
            e2.record(stream)
        return e0, e1, e2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step(False)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = plt.launch_count()
    evs = [step(True) for _ in range(args.steps)]
    barrier()
    launches = plt.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    frame_ms = sum(e0.elapsed_time(e2) for e0, _, e2 in evs) / args.steps
    kern_ms = sum(e0.elapsed_time(e1) for e0, e1, _ in evs) / args.steps
    t = torch.tensor([frame_ms, kern_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    frame_ms, kern_ms = float(t[0]), float(t[1])

    # ---- end to end through the public API with HOST buffers (sample1.cpp render loop :466-490):
    # WriteBuffer(RTProp) from pinned memory, TraceRays (blocking), ReadBuffer(image) into pinned memory
    host_rt = torch.zeros(4, dtype=torch.int32).pin_memory()
    host_img = torch.zeros(count * 4, dtype=torch.uint8).pin_memory()
    host_rt[1], host_rt[2] = 1, 1

    def e2e_step():
        rd.WriteBuffer(plt, ds.rdRTProp, 16, host_rt.data_ptr())
        with torch.cuda.stream(stream):
            render_and_gather()
        if rank == 0:
            rd.ReadBuffer(plt, ds.rdImage, n * 4 if world > 1 else count * 4, host_full.data_ptr(), 0)
        else:
            plt.sync()
    host_full = torch.zeros(n * 4, dtype=torch.uint8).pin_memory()
    if streamed:
        # same two API calls per frame; the bands of the frame travel to host_full while the later bands render
        rd.StreamBufferToHost(plt, ds.rdImage, host_full.data_ptr(), 4, args.readback_bands)
    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if streamed:
        rd.StreamBufferToHost(plt, None, None)
        with torch.cuda.stream(stream):   # the streamed frame must be the frame the plain path delivers
            render_and_gather()
        check = torch.zeros(n * 4, dtype=torch.uint8).pin_memory()
        rd.ReadBuffer(plt, ds.rdImage, n * 4, check.data_ptr(), 0)
        if not torch.equal(check, host_full):
            raise SystemExit("bench: streamed read-back delivered a different frame than TraceRays + ReadBuffer")
    t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t[0])

    multi = {}
    if world > 1 and not args.no_extra:
        # frame ms of the path-traced configs across the ranks (the "frame ms at 1/2/4/8 B200" half of the metric)
        try:
            multi["C3"] = multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world,
                "C3: path.cl, 1080p 16 spp depth 5, blob-field n=2237", scenes.blob_field_scene(2237, WIDTH, HEIGHT), WIDTH, HEIGHT, 16, 5)
            if args.extra == "all":
                multi["C5"] = multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world,
                    "C5: path.cl, 3840x2160 64 spp depth 5, 5 instances of the n=2237 height-field (50 M triangles)",
                    scenes.replicated_scene(2237, 5, 3840, 2160), 3840, 2160, 64, 5, frames=2)
        except Exception as e:
            multi["error"] = repr(e)[:300]
    if rank == 0:
        # algorithmic bytes per ray: counters-on build of the same traversal over the same primary rays
        ra, rb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16)
        ha, hb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 4)
        rd.GeneratePrimaryRays(plt, ds.rdCam, 0, 0, WIDTH, HEIGHT, ra, rb)
        rd.TraceRayList(plt, tlas, ra, rb, ha, hb, n, rd.TRACE_COUNTERS)
        plt.sync()
        c = rd.TraceCounters(plt)
        nodes_per_ray, tris_per_ray = c["nodes"] / c["rays"], c["tris"] / c["rays"]
        bytes_per_ray = nodes_per_ray * NODE_BYTES + tris_per_ray * TRI_BYTES + RAY_BYTES + HIT_BYTES
        peak, peak_kind = measured_peaks()
        rays_per_launch = n // world if (steal or interleave) else count
        achieved = rays_per_launch * bytes_per_ray / (kern_ms * 1e-3) / 1e9
        out = {
            "metric": "Mrays/s (primary closest-hit, 1080p)", "value": n / (frame_ms * 1e-3) / 1e6, "unit": "Mrays/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": frame_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "l2": "256 MiB flush between timed frames", "rays_per_step": n,
                       "partition": ("single GPU" if world == 1 else
                                     "interleaved bands of %d 8x4-tile row(s), round robin over ranks + NCCL sum-reduce to rank 0" % args.band if interleave else
                                     "dynamic stealing of 8x4-tile chunks (IPC counter, system-scope atomics) + NCCL sum-reduce to rank 0" if steal
                                     else "contiguous 8x4-tile rows per GPU, NCCL send/recv gather to rank 0"),
                       "build_ms_incl_upload": build_ms, "build_ms_incl_upload_warm": build_warm_ms,
                       "bvh": blas[0].stats(), "raygen": module.info()},
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": n / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 16,
                    "d2h_bytes_per_step": n * 4 if world > 1 else count * 4, "ms_per_step": e2e_ms,
                    "readback": ("streamed in %d bands while the frame renders (rd_buffer_stream_to_host)" % args.readback_bands)
                    if streamed else "ReadBuffer after the frame"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH if world == 1 else None, "traffic_unit": "bytes per launch (ncu dram read + write)",
                         "algorithmic_bytes_per_launch": rays_per_launch * bytes_per_ray, "peak_kind": peak_kind, "kernel": "raygen (NVRTC, programs/primary.cl + inlined traversal)",
                         "kernel_ms": kern_ms, "bytes_per_ray": bytes_per_ray, "nodes_per_ray": nodes_per_ray,
                         "tris_per_ray": tris_per_ray},
        }
        if multi:
            out["extra"] = multi
        if world == 1 and not args.no_extra:
            # the path-tracing half of the metric ("primary + 4-bounce path"): BASELINE configs[2] (and configs[3]
            # with --extra all), wavefront mode against the inline raygen kernel of the same program
            inflight = 1 << 25   # paths in flight per wavefront pass (all 16 samples of a 1080p frame); 1 = one sample per pass
            modes = (("inline", 0, 1, inflight), ("wavefront_sorted", 1, 1, inflight), ("wavefront_unsorted", 1, 0, inflight),
                     ("wavefront_sorted_one_sample_per_pass", 1, 1, 1))
            out["extra"] = {}
            try:
                out["extra"]["C3"] = extra_config(torch, pkg, plt, stream, flush,
                    "C3: 4-bounce path tracing (depth 5) with any-hit shadow rays, 1080p 16 spp, blob-field n=2237, programs/path.cl",
                    scenes.blob_field_scene(2237, WIDTH, HEIGHT), "path", 16, 5, modes)
                if args.extra == "all":
                    # single-GPU leg of BASELINE configs[4] (4K, 64 spp, 50 M triangles): the denominator of its 2/4/8-GPU frame ms
                    out["extra"]["C5"] = extra_config(torch, pkg, plt, stream, flush,
                        "C5: path.cl, 3840x2160 64 spp depth 5, 5 instances of the n=2237 height-field (50 M triangles)",
                        scenes.replicated_scene(2237, 5, 3840, 2160), "path", 64, 5, (("wavefront_sorted", 1, 1, inflight),), 3840, 2160)
                    out["extra"]["C4"] = extra_config(torch, pkg, plt, stream, flush,
                        "C4: 64 SBT hit groups, 4096 instances, n=2240 height-field, 1080p 16 spp depth 5, programs/groups.cl",
                        scenes.groups_scene(2240, 64, 64, WIDTH, HEIGHT), "groups", 16, 5, modes)
            except Exception as e:   # the headline line must survive a failure here
                out["extra"]["error"] = repr(e)[:300]
        if world == 1 and not args.no_cpu_baseline:
            ref = entry.load_oracle()
            if ref.available():
                out["cpu_baseline"] = cpu_baseline_leg(pkg, ref, bundle)
            else:
                out["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "reference",
                                       "sample": "oracle/_ref not built on this box"}
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
