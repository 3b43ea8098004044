"""Developer tool: C2 frame with the RGBA8 framebuffer (a) in HBM + ReadBuffer, (b) written by the kernel straight into
pinned host memory (zero copy over PCIe). Prints ms per frame end to end (host buffer valid at the end of each step)."""
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
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
bundle = scenes.heightfield_scene(708, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, 1, 1, 0))
pipe, module = scenes.make_pipeline(plt, "primary")
rd.BindPipeline(plt, pipe)
n = W * H
host = torch.empty(n * 4, dtype=torch.uint8).pin_memory()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(label, setup, step, reps=20):
    setup()
    ts = []
    for it in range(reps + 3):
        flush.fill_(it)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step()
        ts.append((time.perf_counter() - t0) * 1e3)
    t = float(np.median(ts[3:]))
    print("%-40s %.4f ms  %.0f Mrays/s e2e" % (label, t, n / t / 1e3), flush=True)
    return host.numpy().copy()


rt = np.array(scenes.rtprop(0, 1, 1, 0)).reshape(1)


def setup_dev():
    rd.BindDescriptorSet(plt, ds.descriptors)


def step_dev():
    rd.WriteBuffer(plt, ds.rdRTProp, 16, rt)
    rd.TraceRays(plt, 0, 0, 0, W, H)
    rd.ReadBuffer(plt, ds.rdImage, n * 4, host.numpy())


a = run("HBM image + ReadBuffer", setup_dev, step_dev)
himg = rd.WrapBuffer(plt, host.data_ptr(), n * 4, owner=host)
dsc = list(ds.descriptors)
dsc[2] = himg


def setup_host():
    host.zero_()
    rd.BindDescriptorSet(plt, rd.CreateDescriptorSet(dsc))


def step_host():
    rd.WriteBuffer(plt, ds.rdRTProp, 16, rt)
    rd.TraceRays(plt, 0, 0, 0, W, H)


b = run("image in pinned host memory (zero copy)", setup_host, step_host)
print("identical:", bool(np.array_equal(a, b)))
