"""Ablation harness (developer tool): times the C2 frame for several NVRTC flag sets / launch options.
usage: python tools/exp_variants.py [n_grid] ; prints one line per variant."""
import os
import sys
import json

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
W, H = 1920, 1080
n_grid = int(sys.argv[1]) if len(sys.argv) > 1 else 708
plt = rd.Platform.GetPlatform(0)
stream = torch.cuda.ExternalStream(plt.stream(), device=0)
plt.set_option("l2_persist", int(os.environ.get("RD_EXP_L2_PERSIST", "0")))
bundle = scenes.heightfield_scene(n_grid, W, H)
import time
tlas, blas = scenes.build_scene(plt, bundle)
plt.sync()
for _k in range(3):
    t0 = time.perf_counter()
    tlas, blas = scenes.build_scene(plt, bundle)
    plt.sync()
    print("bvh", blas[0].stats(), "build_ms(rebuild %d, incl upload) %.1f" % (_k, (time.perf_counter() - t0) * 1e3), flush=True)
def fetch_counts():
    """wide nodes / triangle records fetched per primary ray (counters-on build of the library's own traversal kernel)"""
    n = W * H
    ra, rb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 16)
    ha, hb = rd.CreateBuffer(plt, n * 16), rd.CreateBuffer(plt, n * 4)
    rd.GeneratePrimaryRays(plt, ds.rdCam, 0, 0, W, H, ra, rb)
    rd.TraceRayList(plt, tlas, ra, rb, ha, hb, n, rd.TRACE_COUNTERS)
    plt.sync()
    c = rd.TraceCounters(plt)
    for hnd in (ra, rb, ha, hb):
        hnd.free()
    return c["nodes"] / c["rays"], c["tris"] / c["rays"]


ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, int(os.environ.get('RD_EXP_SPP', '1')), int(os.environ.get('RD_EXP_DEPTH', '1')), 0))
print("fetches per primary ray: %.2f wide nodes, %.2f triangle records" % fetch_counts(), flush=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
VARIANTS = json.loads(os.environ.get("RD_EXP_VARIANTS", "null")) or [
    {"flags": ""}, {"flags": "-DRD_DEQUANT_I2F=1"}, {"flags": "-DRD_POSTPONE=5"}, {"flags": "-DRD_POSTPONE=3"},
    {"flags": "", "block": 64}, {"flags": "", "block": 256}, {"flags": "", "tile": 0}, {"flags": "-DRD_STACK_ENTRIES=24"},
]
base = None
for v in VARIANTS:
    os.environ["RD_NVRTC_FLAGS"] = v.get("flags", "")
    plt.set_option("block", v.get("block", 128))
    plt.set_option("tile", v.get("tile", 1))
    plt.set_option("fmad", v.get("fmad", 1))
    plt.set_option("stream_image", 1 if v.get("banded") else 0)
    pipe, module = scenes.make_pipeline(plt, v.get("program", "primary"))
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    if v.get("banded"):   # time the band-reporting raygen (its copies run on the copy stream, outside the events)
        host = torch.zeros(W * H * 4, dtype=torch.uint8).pin_memory()
        rd.StreamBufferToHost(plt, ds.rdImage, host.data_ptr(), 4, int(v["banded"]))
    ms = []
    for it in range(13):
        with torch.cuda.stream(stream):
            flush.fill_(it)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rd.TraceRaysAsync(plt, W, H)
            e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        if v.get("banded"):
            rd.ReadBuffer(plt, ds.rdImage, W * H * 4, host.data_ptr())
    if v.get("banded"):
        rd.StreamBufferToHost(plt, None, None)
    img = ds.read_image()
    if base is None:
        base = img
    diff = int((img != base).any(axis=-1).sum())
    t = float(np.median(ms[3:]))
    import zlib
    print("%-50s %s  %.4f ms  %.0f Mrays/s  min %.4f  regs %s  px_diff %d  crc %08x" % (
        json.dumps(v), "", t, W * H / t / 1e3, min(ms[3:]), module.info(), diff, zlib.crc32(img.tobytes())), flush=True)
