"""Developer tool: end-to-end frame time (WriteBuffer RTProp + TraceRays + ReadBuffer into pinned memory) of the C2 frame
with the plain read-back and with rd_buffer_stream_to_host for several band counts.
usage: python tools/exp_readback.py [n_grid] ; prints one line per variant."""
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
n_grid = int(sys.argv[1]) if len(sys.argv) > 1 else 708
plt = rd.Platform.GetPlatform(0)
bundle = scenes.heightfield_scene(n_grid, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, 1, 1, 0))
plt.set_option("stream_image", 1)
pipe, module = scenes.make_pipeline(plt, "primary")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
host_rt = torch.zeros(4, dtype=torch.int32).pin_memory()
host_rt[1], host_rt[2] = 1, 1
host = torch.zeros(W * H * 4, dtype=torch.uint8).pin_memory()
base = None
for bands in [int(x) for x in os.environ.get('RD_EXP_BANDS', '0,1,4,8,12,16,24,32,48,64,0').split(',')]:
    rd.StreamBufferToHost(plt, ds.rdImage if bands else None, host.data_ptr() if bands else None, 4, max(1, bands))

    def frame():
        rd.WriteBuffer(plt, ds.rdRTProp, 16, host_rt.data_ptr())
        rd.TraceRaysAsync(plt, W, H)
        rd.ReadBuffer(plt, ds.rdImage, W * H * 4, host.data_ptr())
    for _ in range(5):
        frame()
    ts = []
    for _ in range(40):
        t0 = time.perf_counter()
        frame()
        ts.append((time.perf_counter() - t0) * 1e3)
    img = host.numpy().copy()
    if base is None:
        base = img
    print("bands %2d  median %.4f ms  min %.4f ms  %.0f Mrays/s  identical %s" % (
        bands, float(np.median(ts)), min(ts), W * H / float(np.median(ts)) / 1e3, bool((img == base).all())), flush=True)
