"""Developer tool: latency of each 8x4-tile row of the C2 frame launched on its own (240 warps: less than one
warp per SM scheduler), i.e. the critical path of the slowest warp of that row, and node counts per ray there."""
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
bundle = scenes.heightfield_scene(int(sys.argv[1]) if len(sys.argv) > 1 else 708, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, 1, 1, 0))
pipe, module = scenes.make_pipeline(plt, "primary")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
row_items = 4 * W
times = []
for r in range(H // 4):
    ms = []
    for it in range(4):
        with torch.cuda.stream(stream):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rd.TraceRaysRange(plt, W, H, r * row_items, row_items)
            e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    times.append(min(ms[1:]))
times = np.array(times)
print("tile-row latency (ms): max %.4f at row %d, median %.4f, p90 %.4f" % (times.max(), times.argmax(), np.median(times), np.percentile(times, 90)))
print("rows above 0.1 ms:", [(int(i), round(float(t), 3)) for i, t in enumerate(times) if t > 0.1][:40])
# node / triangle fetches of the slowest row's rays
n = W * H
cam = rd.CreateBuffer(plt, 48)
rd.WriteBuffer(plt, cam, 48, np.array(ds.cam).reshape(1))
ra, rb, ha, hb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 4)
rd.GeneratePrimaryRays(plt, cam, 0, 0, W, H, ra, rb)
plt.sync()
rays_a = np.zeros((n, 4), np.float32); rays_b = np.zeros((n, 4), np.float32)
rd.ReadBuffer(plt, ra, n * 16, rays_a); rd.ReadBuffer(plt, rb, n * 16, rays_b)
worst = int(times.argmax())
sel = slice(worst * row_items, (worst + 1) * row_items)
rays = np.concatenate([rays_a[sel], rays_b[sel]], axis=1)
per = []
for k in range(0, row_items, 32):   # per warp-sized group: counters of 32 rays
    rd.trace_host_rays(plt, tlas, rays[k:k + 32], rd.TRACE_COUNTERS)
    c = rd.TraceCounters(plt)
    per.append((c["nodes"] / 32.0, c["tris"] / 32.0))
per = np.array(per)
print("slowest row %d: nodes/ray per warp: mean %.1f max %.1f; tris/ray mean %.1f max %.1f" % (worst, per[:, 0].mean(), per[:, 0].max(), per[:, 1].mean(), per[:, 1].max()))
