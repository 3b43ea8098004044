"""Developer tool: times the wavefront building blocks (primary ray generation, static-batch and
dynamic-fetch trace kernels) on the C2 / C3-size height-fields and checks they agree."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
stream = torch.cuda.ExternalStream(plt.stream(), device=0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=10):
    ms = []
    for it in range(reps + 3):
        with torch.cuda.stream(stream):
            flush.fill_(it)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms[3:]))


for n_grid in [int(a) for a in sys.argv[1:]] or [708]:
    bundle = scenes.heightfield_scene(n_grid, W, H)
    import time
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    print("n_grid", n_grid, "build+upload %.1f ms" % ((time.perf_counter() - t0) * 1e3), blas[0].stats(), flush=True)
    n = W * H
    cam = rd.CreateBuffer(plt, 48)
    rd.WriteBuffer(plt, cam, 48, np.array(bundle.camera).reshape(1))
    ra, rb, ha, hb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 4)
    t_gen = timed(lambda: rd.GeneratePrimaryRays(plt, cam, 0, 0, W, H, ra, rb))
    print("  primary ray generation %.4f ms" % t_gen, flush=True)
    res = {}
    for name, flags in [("static", 0), ("dynamic", rd.TRACE_DYNAMIC_FETCH), ("static_first", 1), ("dynamic_first", rd.TRACE_DYNAMIC_FETCH | 1)]:
        t = timed(lambda: rd.TraceRayList(plt, tlas, ra, rb, ha, hb, n, flags))
        out = np.zeros(n, rd.HIT_DTYPE)
        rd.ReadBuffer(plt, ha, n * 16, out)
        res[name] = out
        print("  trace %-14s %.4f ms  %.0f Mrays/s  hits %d" % (name, t, n / t / 1e3, int((out["t"] < 3e38).sum())), flush=True)
    print("  static vs dynamic: prim mismatches", int((res["static"]["prim"] != res["dynamic"]["prim"]).sum()), "t mismatches", int((res["static"]["t"] != res["dynamic"]["t"]).sum()), flush=True)
    assert np.array_equal(res["static_first"]["t"] < 3e38, res["dynamic_first"]["t"] < 3e38)
    rd.TraceRayList(plt, tlas, ra, rb, ha, hb, n, rd.TRACE_COUNTERS)
    plt.sync()
    c = rd.TraceCounters(plt)
    print("  per ray: nodes %.2f tris %.2f" % (c["nodes"] / c["rays"], c["tris"] / c["rays"]), flush=True)
    # megakernel through TraceRays for comparison
    ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, 1, 1, 0))
    pipe, module = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    t = timed(lambda: rd.TraceRaysAsync(plt, W, H))
    print("  megakernel TraceRays     %.4f ms  %.0f Mrays/s  %s" % (t, n / t / 1e3, module.info()), flush=True)
    for h in (cam, ra, rb, ha, hb, tlas):
        h.free()
