"""Smoke-sized run of every launch mode for compute-sanitizer (one tool per gpurun call):
inline raygen, wavefront stages (sorted), chunk-stealing raygen, dynamic work units, banded read-back, the frame-gather
pack / unpack kernels, a deferred small build, a general (multi-launch) build and an explicit ray list."""
import os
import sys

import faulthandler

import numpy as np

if os.environ.get("RD_SMOKE_WATCHDOG"):
    faulthandler.dump_traceback_later(int(os.environ["RD_SMOKE_WATCHDOG"]), exit=True)


def stage(msg):
    print("[smoke] " + msg, flush=True)


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
plt = rd.Platform.GetPlatform(0)
b = scenes.cornell_like(2)
tlas, _ = scenes.build_scene(plt, b)
w, h = 96, 72
ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
plt.set_option("stealing", 1)
plt.set_option("stream_image", 1)
pipe, _ = scenes.make_pipeline(plt, "path")
plt.set_option("stealing", 0)
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
rd.TraceRays(plt, 0, 0, 0, w, h)
plain = ds.read_image()
stage("inline done")
plt.set_option("wavefront", 1)
ds.set_rtprop(0, 2, 3)
rd.TraceRays(plt, 0, 0, 0, w, h)
wave = ds.read_image()
stage("wavefront done")
mask = rd.TraceRaysUnits(plt, w, h, 6, 300)
stage("units wavefront done")
plt.sync()
plt.set_option("wavefront", 0)
if "steal" not in os.environ.get("RD_SMOKE_SKIP", ""):
    rd.TraceRaysSteal(plt, w, h, 301, 64)
stage("steal done")
plt.sync()
mask2 = rd.TraceRaysUnits(plt, w, h, 9, 302)
stage("units inline done")
plt.sync()
dst = rd.CreateBuffer(plt, w * h * 4)
for r in range(3):
    rd._check(rd.lib().rd_debug_band_roundtrip(plt.handle, ds.rdImage.handle, dst.handle, w, h, 4, 2, r, 3))
import torch
host = torch.zeros(w * h * 4, dtype=torch.uint8).pin_memory()
rd.StreamBufferToHost(plt, ds.rdImage, host.data_ptr(), 4, 4)
ds.set_rtprop(0, 2, 3)
rd.TraceRays(plt, 0, 0, 0, w, h)
rd.ReadBuffer(plt, ds.rdImage, w * h * 4, host.data_ptr())
rd.StreamBufferToHost(plt, None, None)
stage("banded done")
hf = scenes.heightfield_scene(72, 64, 48)       # 10 368 triangles: the general multi-launch build
t2, _ = scenes.build_scene(plt, hf)
rays = scenes.pinhole_rays(hf.camera, stride=2)
hitA, hitB = rd.trace_host_rays(plt, t2, rays)
first, _ = rd.trace_host_rays(plt, t2, rays, rd.TRACE_TERMINATE_ON_FIRST_HIT)
stage("ray lists done")
# determinism: the wavefront frame (atomics in the queues, the sort and the compaction) rendered five more times
plt.set_option("wavefront", 1)
frames = []
for k in range(5):
    ds.set_rtprop(0, 2, 3)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    frames.append(ds.read_scratch().tobytes())
plt.set_option("wavefront", 0)
viol, checked = plt.bounds_violations()
print("bounds-checked build: %d, violations: %d, wavefront frames identical over 5 runs: %s, device status %r" % (
    checked, viol, all(f == frames[0] for f in frames), plt.status()))
print("sanitize smoke done: inline mean %.2f, wavefront mean %.2f, units %x %x, hits %d" % (
    plain[..., :3].mean(), wave[..., :3].mean(), mask, mask2, int((hitA["t"] < 3e38).sum())))
