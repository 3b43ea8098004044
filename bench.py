#!/usr/bin/env python
"""bench.py -- BASELINE.json metric. Headline = BASELINE config C2 (configs[1]): primary-ray closest-hit, 1920x1080, 1 spp,
procedural 1 002 528-triangle height-field, user program programs/primary.cl through RD::TraceRays. One "step" = one
TraceRays frame (W*H primary rays).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

The same JSON line carries, under "extra", every other BASELINE configuration at its own size (each a few frames):
  P10M  north_star's target: primary closest-hit on the 10 M-triangle height-field (own roofline + issue-slot accounting)
  C1    the reference's own samples/shader.cl + sbt.json (800x600, 16 spp, depth 8) when oracle/_ref/samples is present
  C3    4-bounce path tracing, 10 M triangles, 16 spp: inline / wavefront sorted / wavefront unsorted
  C4    64 hit groups, 4096 instances: hit-group sort on vs off
  C5    4K 64 spp on the 50 M-triangle replicated scene (the single-GPU leg of the multi-GPU configuration)
N > 1 is launched by torchrun (one rank per GPU): the scene is replicated, the frame is dealt to the ranks in
interleaved bands of 8x4-tile rows (rd_trace_rays_interleaved) or taken dynamically in work units from a job-wide
counter (rd_trace_rays_units), and the finished framebuffer bands are gathered on rank 0 with NCCL through the C ABI
(rd_frame_gather / rd_frame_gather_units) -- the only collective. Scaling is "strong" (the frame is fixed); extras at
N > 1: C3 and C5 frame ms. Prints ONE JSON line on rank 0. See DESIGN.md "Measurement" for every field's definition.
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
L2_NOTE = "256 MiB flush between timed frames"
NODE_BYTES, TRI_BYTES, RAY_BYTES, HIT_BYTES = 80, 48, 32, 16
# 148 SMs x 4 schedulers x 1.965 GHz: warp instructions the B200 can issue per second (one per scheduler per cycle)
ISSUE_PEAK = 148 * 4 * 1.965e9


def ncu_reference(key):
    """Per-launch DRAM bytes and warp instructions of the dominant kernel from the committed ncu --set full captures
    (profiles/ncu_traffic.json names the capture file and the commit it was taken at); None when absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(key)
    except Exception:
        return None


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
    tl = ref.build_scene(b)
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
                      "config": {"workload": WORKLOAD, "l2": L2_NOTE, "rays_per_step": WIDTH * HEIGHT, "partition": "single GPU"},
                      "details": {"sample": sample, "note": "CPU arm: each step renders the sample rows of the frame; value = rays of the sample / time"},
                      "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample},
                      "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


def opencl_cpu_device_note():
    """the reference's true path is clBuildProgram on a PoCL CPU device; it is timed only where an OpenCL
    platform exists at run time. This box has no ICD, so the row reads 'PoCL not present' (the timed arm is the reference's own
    sources compiled as C++)."""
    import ctypes.util
    if not (ctypes.util.find_library("OpenCL") and os.path.isdir("/etc/OpenCL/vendors") and os.listdir("/etc/OpenCL/vendors")):
        return "PoCL not present"
    return "an OpenCL ICD is installed on this box; the PoCL path is not wired into this bench"


def cpu_baseline_leg(pkg, ref, bundle):
    """Reference CPU path on a bounded sample (rank 0, N=1 only): see reference_arm for the sample."""
    scenes = pkg.scenes
    cores = os.cpu_count() or 1
    t0 = time.time()
    tl = ref.build_scene(bundle)
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
                      % (rows, HEIGHT, count, build_s),
            "opencl_path": opencl_cpu_device_note()}


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
        plt.set_option("wf_paths", 1 << 26)
    for h in [tlas] + list(blas):
        h.free()
    return out


def primary_config(torch, pkg, plt, stream, flush, name, grid_n, key):
    """Primary closest-hit on a height-field of grid_n x grid_n quads at 1080p (P10M: north_star's target configuration):
    Mrays/s, the traversal's own node / triangle fetch counts, the HBM roofline and the issue-slot accounting."""
    rd, scenes = pkg.rd, pkg.scenes
    bundle = scenes.heightfield_scene(grid_n, WIDTH, HEIGHT)
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    build_ms = (time.perf_counter() - t0) * 1e3   # first build of this size in the process: pool growth, first pageable upload
    for hnd in [tlas] + list(blas):
        hnd.free()
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    build_warm_ms = (time.perf_counter() - t0) * 1e3
    ds = scenes.DeviceScene(plt, bundle, tlas, WIDTH, HEIGHT, scenes.rtprop(0, 1, 1, 0))
    pipe, module = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    ms = time_frames(torch, plt, rd, stream, flush, WIDTH, HEIGHT, reps=9)
    n = WIDTH * HEIGHT
    ra, rb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16)
    ha, hb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 4)
    rd.GeneratePrimaryRays(plt, ds.rdCam, 0, 0, WIDTH, HEIGHT, ra, rb)
    rd.TraceRayList(plt, tlas, ra, rb, ha, hb, n, rd.TRACE_COUNTERS)
    plt.sync()
    c = rd.TraceCounters(plt)
    for hnd in (ra, rb, ha, hb):
        hnd.free()
    nodes_per_ray, tris_per_ray = c["nodes"] / c["rays"], c["tris"] / c["rays"]
    bytes_per_ray = nodes_per_ray * NODE_BYTES + tris_per_ray * TRI_BYTES + RAY_BYTES + HIT_BYTES
    peak, peak_kind = measured_peaks()
    achieved = n * bytes_per_ray / (ms * 1e-3) / 1e9
    ncu = ncu_reference(key) or {}
    out = {"workload": name, "triangles": bundle.triangle_count(), "value": n / ms / 1e3, "unit": "Mrays/s", "ms_per_frame": ms,
           "build_ms_incl_upload": build_ms, "build_ms_incl_upload_warm": build_warm_ms, "bvh": blas[0].stats(), "raygen": module.info(),
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": ncu.get("dram_bytes_per_launch"), "traffic_source": ncu.get("source"),
                        "algorithmic_bytes_per_launch": n * bytes_per_ray, "peak_kind": peak_kind,
                        "kernel": "raygen (NVRTC, programs/primary.cl + inlined traversal)", "kernel_ms": ms,
                        "bytes_per_ray": bytes_per_ray, "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray}}
    if ncu.get("warp_inst_per_launch"):
        wi = ncu["warp_inst_per_launch"] / n
        out["issue"] = {"warp_inst_per_ray": wi, "source": ncu.get("source"), "issue_peak_warp_inst_per_s": ISSUE_PEAK,
                        "ceiling_mrays_at_full_issue": ISSUE_PEAK / wi / 1e6, "frac_of_issue_ceiling": (n / ms / 1e3) / (ISSUE_PEAK / wi / 1e6)}
    for hnd in [tlas] + list(blas):
        hnd.free()
    return out


def c1_config(torch, pkg, plt, stream, flush):
    """BASELINE configs[0] on the GPU: the reference's own samples/shader.cl + sbt.json (copied unmodified into the git-ignored
    oracle/_ref/samples/ by oracle/Makefile -- plugin INPUT, read here as a user would read their shader file), SURVEY 8(d)
    C1 scene, 800x600, 16 spp, depth 8."""
    rd, scenes = pkg.rd, pkg.scenes
    cl = os.path.join(ROOT, "oracle", "_ref", "samples", "shader.cl")
    sbt = os.path.join(ROOT, "oracle", "_ref", "samples", "sbt.json")
    if not (os.path.exists(cl) and os.path.exists(sbt)):
        return {"unavailable": "oracle/_ref/samples/shader.cl not present on this box"}
    src, sbt_text = open(cl).read(), open(sbt).read()
    w, h, spp, depth = 800, 600, 16, 8
    b = scenes.cornell_like()
    b.camera["x"], b.camera["y"] = 0.013, 1.037
    tlas, blas = scenes.build_scene(plt, b)
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, spp, depth, 0))
    plt.set_option("count_rays", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, source=src, sbt=sbt_text)
    finally:
        plt.set_option("count_rays", 0)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    plt.ray_count()
    rd.TraceRays(plt, 0, 0, 0, w, h)
    rays = plt.ray_count()
    pipe, module = scenes.make_pipeline(plt, source=src, sbt=sbt_text)
    rd.BindPipeline(plt, pipe)
    ms = time_frames(torch, plt, rd, stream, flush, w, h, reps=5)
    out = {"workload": "C1: samples/shader.cl + samples/sbt.json as shipped, Cornell-like scene (%d triangles), 800x600, 16 spp, depth 8" % b.triangle_count(),
           "rays_per_frame": rays, "value": rays / ms / 1e3, "unit": "Mrays/s", "ms_per_frame": ms, "raygen": module.info(),
           "parity": "tests/test_gpu_c1.py compares this frame with the reference runtime (profiles/parity_r2.json)"}
    # the same unmodified file with the two wavefront hooks a host appends (programs/shader_wavefront_hooks.cl): stage launches
    hooks = os.path.join(ROOT, "radiance-ray-tracing_b200", "programs", "shader_wavefront_hooks.cl")
    try:
        pipe, module = scenes.make_pipeline(plt, source=src + "\n" + open(hooks).read(), sbt=sbt_text)
        rd.BindPipeline(plt, pipe)
        plt.set_option("wavefront", 1)
        wms = time_frames(torch, plt, rd, stream, flush, w, h, reps=5)
        out["wavefront_with_hooks"] = {"ms_per_frame": wms, "value": rays / wms / 1e3,
                                       "program": "samples/shader.cl + programs/shader_wavefront_hooks.cl (two functions appended by the host)"}
    except Exception as e:
        out["wavefront_with_hooks"] = {"error": repr(e)[:300]}
    finally:
        plt.set_option("wavefront", 0)
    for hnd in [tlas] + list(blas):
        hnd.free()
    return out


_UNIT_FRAME = [1000]


def comm_setup(torch, dist, rd, plt, rank, world):
    """One NCCL communicator inside the library per rank (rd_comm_init): rank 0 creates the id, torch.distributed only
    carries its 128 bytes; the dynamic partition's counter handle travels the same way."""
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid.copy_(torch.from_numpy(rd.CommUniqueId()))
    dist.broadcast(uid, src=0)
    rd.CommInit(plt, rank, world, uid.cpu().numpy())


def multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world, name, bundle, w, h, spp, depth, frames=3, units_per_rank=4):
    """Path tracing across all ranks, wavefront mode: frame ms (max over ranks, device events, gather included) of the
    static interleaved band partition (rd_trace_rays_interleaved + rd_frame_gather) and of the dynamic work-unit
    partition (rd_trace_rays_units + rd_frame_gather_units); Mrays/s with the rays of one frame counted on rank 0."""
    rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
    tlas, blas = scenes.build_scene(plt, bundle)
    ds = scenes.DeviceScene(plt, bundle, tlas, w, h, scenes.rtprop(0, spp, depth, 0))
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
    mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    units = min(64, world * units_per_rank)
    frame_no = _UNIT_FRAME   # counter slots are consumed in order over the whole job (rd_trace_rays_units)

    def timed(render):
        ms = []
        for f in range(frames + 1):
            torch.cuda.synchronize()
            dist.barrier()
            ds.set_rtprop(0, spp, depth)
            plt.sync()
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                render()
                e1.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms.append(float(t[0]))
        return float(np.median(ms[1:]))

    def static_frame():
        rd.TraceRaysInterleaved(plt, w, h, rank, world, 1)
        rd.FrameGather(plt, ds.rdImage, w, h, 4, 1)

    owned = [0]

    def dynamic_frame():
        frame_no[0] += 1
        owned[0] = rd.TraceRaysUnits(plt, w, h, units, frame_no[0])
        rd.FrameGatherUnits(plt, ds.rdImage, w, h, 4, units, owned[0])

    plt.set_option("wavefront", 1)
    try:
        static_ms = timed(static_frame)
        img_static = ds.read_image() if rank == 0 else None
        dynamic_ms = timed(dynamic_frame)
        img_dynamic = ds.read_image() if rank == 0 else None
        # per-rank stage timeline of one static frame (what the frame's time is made of on every rank)
        plt.set_option("wf_timeline", 1)
        ds.set_rtprop(0, spp, depth)
        plt.sync()
        rd.TraceRaysInterleaved(plt, w, h, rank, world, 1)
        tl = rd.WavefrontTimeline(plt)
        plt.set_option("wf_timeline", 0)
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_timeline", 0)
    stage = {}
    for nme, t in tl:
        stage[nme] = stage.get(nme, 0.0) + t
    mine = torch.tensor([stage.get(k, 0.0) for k in ("generate", "trace", "sort", "shade", "accumulate", "resolve")] +
                        [float(bin(owned[0]).count("1"))], dtype=torch.float64, device="cuda")
    allt = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allt, mine)
    for hnd in [tlas] + list(blas):
        hnd.free()
    if rank != 0:
        return None
    same = bool(np.abs(img_static[..., :3].astype(np.int32) - img_dynamic[..., :3].astype(np.int32)).max() <= 1)
    best = min(static_ms, dynamic_ms)
    return {"workload": name, "triangles": bundle.triangle_count(), "spp": spp, "depth": depth, "rays_per_frame": rays,
            "frame_ms": best, "value": rays / best / 1e3 if rays else None, "unit": "Mrays/s",
            "wavefront_interleaved": {"frame_ms": static_ms, "partition": "interleaved tile-row bands + rd_frame_gather (one packed NCCL send per peer)"},
            "wavefront_units": {"frame_ms": dynamic_ms, "units": units, "units_per_rank_last_frame": [int(t[6]) for t in allt],
                                "partition": "dynamic work units from the job-wide counter + rd_frame_gather_units",
                                "frame_equals_static": same},
            "stage_ms_per_rank_static": {k: [round(float(t[i]), 3) for t in allt]
                                         for i, k in enumerate(("generate", "trace", "sort", "shade", "accumulate", "resolve"))}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--partition", default="interleave", choices=["interleave", "units", "static"],
                    help="N>1 headline: interleaved tile-row bands + rd_frame_gather (default), dynamic work units + "
                         "rd_frame_gather_units, or contiguous tile rows + rd_frame_gather_ranges")
    ap.add_argument("--band", type=int, default=1, help="tile rows per band of the interleaved partition")
    ap.add_argument("--units-per-rank", type=int, default=4, help="work units per rank of the dynamic partition")
    ap.add_argument("--extra", default="all", help="comma list of extra configs (P10M,C1,C3,C4,C5), 'all' or 'none'")
    ap.add_argument("--readback-bands", type=int, default=24,
                    help="N=1 end-to-end leg: bands of the streamed read-back (rd_buffer_stream_to_host); 0 = plain ReadBuffer")
    ap.add_argument("--no-extra", action="store_true", help="same as --extra none")
    args = ap.parse_args()
    claim_stdout()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        line = reference_arm(args, rank, world)
        if line:
            emit(line)
        return
    want = set() if (args.no_extra or args.extra == "none") else (
        {"P10M", "C1", "C3", "C4", "C5"} if args.extra == "all" else {x.strip().upper() for x in args.extra.split(",")})

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
    stream = torch.cuda.ExternalStream(plt.stream(), device=local)
    if world > 1:
        comm_setup(torch, dist, rd, plt, rank, world)
        mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")

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
    n = WIDTH * HEIGHT
    streamed = world == 1 and args.readback_bands > 0
    if streamed:   # the module also gets the band-reporting raygen; it only runs once a host mirror is registered (e2e leg)
        plt.set_option("stream_image", 1)
    pipe, module = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    firsts = [mg.tile_row_range(r, world, WIDTH, HEIGHT)[0] for r in range(world)]
    counts = [mg.tile_row_range(r, world, WIDTH, HEIGHT)[1] for r in range(world)]
    units = min(64, world * args.units_per_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    frame_no = [0]

    def render_and_gather():
        """One frame of this rank's share + the frame's only collective; returns the event between them."""
        e1 = torch.cuda.Event(enable_timing=True)
        if world == 1:
            rd.TraceRaysAsync(plt, WIDTH, HEIGHT)
            e1.record(stream)
        elif args.partition == "interleave":
            rd.TraceRaysInterleaved(plt, WIDTH, HEIGHT, rank, world, args.band)
            e1.record(stream)
            rd.FrameGather(plt, ds.rdImage, WIDTH, HEIGHT, 4, args.band)
        elif args.partition == "units":
            frame_no[0] += 1
            owned = rd.TraceRaysUnits(plt, WIDTH, HEIGHT, units, frame_no[0])
            e1.record(stream)
            rd.FrameGatherUnits(plt, ds.rdImage, WIDTH, HEIGHT, 4, units, owned)
        else:
            rd.TraceRaysRange(plt, WIDTH, HEIGHT, firsts[rank], counts[rank])
            e1.record(stream)
            rd.FrameGatherRanges(plt, ds.rdImage, 4, firsts, counts)
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
    # WriteBuffer(RTProp) from pinned memory, TraceRays (blocking), ReadBuffer(image) into pinned memory.
    # Cache state: the device-timed loop above flushes L2 before every frame; this loop does not (a host program renders frame
    # after frame), so e2e frames see whatever of the 56 MB scene the previous frame left in the 126 MB L2 -- labelled below.
    host_rt = torch.zeros(4, dtype=torch.int32).pin_memory()
    host_rt[1], host_rt[2] = 1, 1
    host_full = torch.zeros(n * 4, dtype=torch.uint8).pin_memory()

    def e2e_step(cold=False):
        rd.WriteBuffer(plt, ds.rdRTProp, 16, host_rt.data_ptr())
        with torch.cuda.stream(stream):
            if cold:
                flush.fill_(rank + 1)
            render_and_gather()
        if rank == 0:
            rd.ReadBuffer(plt, ds.rdImage, n * 4, host_full.data_ptr(), 0)
        else:
            plt.sync()

    def e2e_loop(cold):
        for _ in range(3):
            e2e_step(cold)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step(cold)
        barrier()
        ms = (time.perf_counter() - t0) * 1e3 / args.steps
        tt = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt[0])

    if streamed:
        # same two API calls per frame; the bands of the frame travel to host_full while the later bands render
        rd.StreamBufferToHost(plt, ds.rdImage, host_full.data_ptr(), 4, args.readback_bands)
    e2e_ms = e2e_loop(False)
    # the same loop with the 256 MiB flush in front of every frame, minus the flush's own time: the cold-cache figure
    with torch.cuda.stream(stream):
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(5):
            flush.fill_(7)
        f1.record(stream)
    torch.cuda.synchronize()
    flush_ms = f0.elapsed_time(f1) / 5
    e2e_cold_ms = e2e_loop(True) - flush_ms
    if streamed:
        rd.StreamBufferToHost(plt, None, None)
        with torch.cuda.stream(stream):   # the streamed frame must be the frame the plain path delivers
            render_and_gather()
        check = torch.zeros(n * 4, dtype=torch.uint8).pin_memory()
        rd.ReadBuffer(plt, ds.rdImage, n * 4, check.data_ptr(), 0)
        if not torch.equal(check, host_full):
            raise SystemExit("bench: streamed read-back delivered a different frame than TraceRays + ReadBuffer")
    if plt.status() != (0, 0):
        raise SystemExit("bench: device status flags raised %r" % (plt.status(),))

    multi = {}
    if world > 1 and want:
        # frame ms of the path-traced configs across the ranks (the "frame ms at 1/2/4/8 B200" half of the metric)
        try:
            if "C3" in want:
                multi["C3"] = multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world,
                    "C3: path.cl, 1080p 16 spp depth 5, blob-field n=2237", scenes.blob_field_scene(2237, WIDTH, HEIGHT), WIDTH, HEIGHT, 16, 5,
                    units_per_rank=args.units_per_rank)
            if "C5" in want:
                multi["C5"] = multi_gpu_path_config(torch, dist, pkg, plt, stream, rank, world,
                    "C5: path.cl, 3840x2160 64 spp depth 5, 5 instances of the n=2237 height-field (50 M triangles)",
                    scenes.replicated_scene(2237, 5, 3840, 2160), 3840, 2160, 64, 5, frames=2, units_per_rank=args.units_per_rank)
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
        rays_per_launch = n // world if world > 1 and args.partition != "static" else counts[0]
        achieved = rays_per_launch * bytes_per_ray / (kern_ms * 1e-3) / 1e9
        ncu = (ncu_reference("C2") or {}) if world == 1 else {}
        partition = ("single GPU" if world == 1 else
                     "interleaved bands of %d 8x4-tile row(s), round robin over ranks + rd_frame_gather (one packed NCCL send per peer to rank 0)" % args.band
                     if args.partition == "interleave" else
                     "dynamic work units (%d, job-wide counter over CUDA IPC / NVLink) + rd_frame_gather_units" % units if args.partition == "units"
                     else "contiguous 8x4-tile rows per GPU + rd_frame_gather_ranges (NCCL send/recv in place)")
        out = {
            "metric": "Mrays/s (primary closest-hit, 1080p)", "value": n / (frame_ms * 1e-3) / 1e6, "unit": "Mrays/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": frame_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "l2": L2_NOTE, "rays_per_step": n, "partition": partition},
            "details": {"build_ms_incl_upload": build_ms, "build_ms_incl_upload_warm": build_warm_ms, "bvh": blas[0].stats(),
                        "raygen": module.info(), "node_bytes": NODE_BYTES},
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": n / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 16,
                    "d2h_bytes_per_step": n * 4, "ms_per_step": e2e_ms,
                    "cache_state": "frames back to back, no L2 flush (the render loop of samples/sample1.cpp:466-490)",
                    "cold": {"value": n / (e2e_cold_ms * 1e-3) / 1e6, "ms_per_step": e2e_cold_ms,
                             "cache_state": "256 MiB L2 flush before every frame, the flush's own %.3f ms subtracted" % flush_ms},
                    "readback": ("streamed in %d bands while the frame renders (rd_buffer_stream_to_host)" % args.readback_bands)
                    if streamed else "ReadBuffer after the frame (rank 0, after the gather)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu.get("dram_bytes_per_launch"), "traffic_source": ncu.get("source"),
                         "traffic_unit": "bytes per launch (ncu dram read + write)",
                         "algorithmic_bytes_per_launch": rays_per_launch * bytes_per_ray, "peak_kind": peak_kind,
                         "kernel": "raygen (NVRTC, programs/primary.cl + inlined traversal)",
                         "kernel_ms": kern_ms, "bytes_per_ray": bytes_per_ray, "nodes_per_ray": nodes_per_ray,
                         "tris_per_ray": tris_per_ray},
        }
        if ncu.get("warp_inst_per_launch"):
            wi = ncu["warp_inst_per_launch"] / n
            out["issue"] = {"warp_inst_per_ray": wi, "source": ncu.get("source"), "issue_peak_warp_inst_per_s": ISSUE_PEAK,
                            "ceiling_mrays_at_full_issue": ISSUE_PEAK / wi / 1e6,
                            "frac_of_issue_ceiling": (n / (kern_ms * 1e-3) / 1e6) / (ISSUE_PEAK / wi / 1e6)}
        if multi:
            out["extra"] = {k: v for k, v in multi.items() if v is not None}
        if world == 1 and want:
            for h in [tlas] + list(blas):   # the extras build their own scenes
                h.free()
            inflight = int(os.environ.get("RD_BENCH_WF_PATHS", 1 << 26))   # paths in flight per wavefront frame over both lanes (the library default: all 16 samples of a 1080p frame, 8 of a 4K frame); developer knob for A/B runs
            modes = (("inline", 0, 1, inflight), ("wavefront_sorted", 1, 1, inflight), ("wavefront_unsorted", 1, 0, inflight))
            out["extra"] = {}
            todo = [
                ("P10M", lambda: primary_config(torch, pkg, plt, stream, flush,
                    "P10M: primary-ray closest-hit, 1920x1080, 1 spp, height-field n=2237 (10008338 triangles), programs/primary.cl", 2237, "P10M")),
                ("C1", lambda: c1_config(torch, pkg, plt, stream, flush)),
                ("C3", lambda: extra_config(torch, pkg, plt, stream, flush,
                    "C3: 4-bounce path tracing (depth 5) with any-hit shadow rays, 1080p 16 spp, blob-field n=2237, programs/path.cl",
                    scenes.blob_field_scene(2237, WIDTH, HEIGHT), "path", 16, 5, modes)),
                ("C4", lambda: extra_config(torch, pkg, plt, stream, flush,
                    "C4: 64 SBT hit groups, 4096 instances, n=2240 height-field, 1080p 16 spp depth 5, programs/groups.cl",
                    scenes.groups_scene(2240, 64, 64, WIDTH, HEIGHT), "groups", 16, 5, modes[1:])),
                # single-GPU leg of BASELINE configs[4] (4K, 64 spp, 50 M triangles): the denominator of its 2/4/8-GPU frame ms
                ("C5", lambda: extra_config(torch, pkg, plt, stream, flush,
                    "C5: path.cl, 3840x2160 64 spp depth 5, 5 instances of the n=2237 height-field (50 M triangles)",
                    scenes.replicated_scene(2237, 5, 3840, 2160), "path", 64, 5, (("wavefront_sorted", 1, 1, inflight),), 3840, 2160)),
            ]
            for key, fn in todo:
                if key not in want:
                    continue
                try:
                    out["extra"][key] = fn()
                except Exception as e:   # the headline line must survive a failure here
                    out["extra"][key] = {"error": repr(e)[:300]}
        if world == 1 and not args.no_cpu_baseline:
            ref = entry.load_oracle()
            if ref.available():
                out["cpu_baseline"] = cpu_baseline_leg(pkg, ref, bundle)
            else:
                out["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "reference",
                                       "sample": "oracle/_ref not built on this box"}
        emit(json.dumps(out))
    if world > 1:
        rd.CommDestroy(plt)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
