"""Developer tool: times a path-tracing user program (programs/path.cl) through RD::TraceRays and counts
its traceRay invocations. usage: python tools/exp_path.py [scene: cornell|hf|c3] [n_grid] [spp] [depth]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
scene = sys.argv[1] if len(sys.argv) > 1 else "c3"
n_grid = int(sys.argv[2]) if len(sys.argv) > 2 else 2237
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 16
depth = int(sys.argv[4]) if len(sys.argv) > 4 else 5
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
stream = torch.cuda.ExternalStream(plt.stream(), device=0)
t0 = time.perf_counter()
bundle = {"cornell": lambda: scenes.cornell_like(), "hf": lambda: scenes.heightfield_scene(n_grid, W, H),
          "c3": lambda: scenes.blob_field_scene(n_grid, W, H),
          "c4": lambda: scenes.groups_scene(n_grid, int(os.environ.get("RD_EXP_SIDE", "64")), 64, W, H)}[scene]()
print("scene %s host generation %.1f s, %d triangles" % (scene, time.perf_counter() - t0, bundle.triangle_count()), flush=True)
t0 = time.perf_counter()
tlas, blas = scenes.build_scene(plt, bundle)
plt.sync()
print("build+upload %.1f ms" % ((time.perf_counter() - t0) * 1e3), [b.stats() for b in blas[:8]], len(blas), flush=True)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, spp, depth, 0))
for program in sys.argv[5:] or ["path"]:
    plt.set_option("count_rays", 1)
    pipe, module = scenes.make_pipeline(plt, program)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    plt.ray_count()
    rd.TraceRays(plt, 0, 0, 0, W, H)
    rays = plt.ray_count()
    plt.set_option("count_rays", 0)
    pipe, module = scenes.make_pipeline(plt, program)
    rd.BindPipeline(plt, pipe)
    base_img = None
    for mode, wf, srt in (("inline", 0, 0), ("wavefront+sort", 1, 1), ("wavefront", 1, 0)):
        plt.set_option("wavefront", wf)
        plt.set_option("wf_sort", srt)
        ms = []
        l0 = plt.launch_count()
        for it in range(4):
            ds.set_rtprop(0, spp, depth)
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                rd.TraceRaysAsync(plt, W, H)
                e1.record(stream)
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        t = float(np.median(ms[1:]))
        img = ds.read_image()
        if base_img is None:
            base_img = img
        d = img[..., :3].astype(np.float64) - base_img[..., :3].astype(np.float64)
        print("%s [%s]: %d spp depth %d: %.2f ms/frame, %d rays/frame (%.2f per sample), %.0f Mrays/s, launches/frame %d, regs %s, mean rgb %s, rmse vs inline %.3f" % (
            program, mode, spp, depth, t, rays, rays / (W * H * spp), rays / t / 1e3, (plt.launch_count() - l0) // 4, module.info(),
            img[..., :3].mean(axis=(0, 1)), np.sqrt((d * d).mean())), flush=True)
    plt.set_option("wavefront", 0)
