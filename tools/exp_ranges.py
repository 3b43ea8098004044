"""Developer tool: where a C2 frame's time goes by screen region, and the cost of the persistent stealing launch."""
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
bundle = scenes.heightfield_scene(708, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, 1, 1, 0))
plt.set_option("stealing", 1)
pipe, module = scenes.make_pipeline(plt, "primary")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)


def timed(fn, do_flush=True):
    ms = []
    for it in range(9):
        with torch.cuda.stream(stream):
            if do_flush:
                flush.fill_(it)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn(it)
            e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms[3:]))


n = W * H
print("full frame        %.4f ms (no flush %.4f)" % (timed(lambda i: rd.TraceRaysRange(plt, W, H, 0, n)), timed(lambda i: rd.TraceRaysRange(plt, W, H, 0, n), False)))
for parts in (2, 4, 8):
    ts = [timed(lambda i, k=k: rd.TraceRaysRange(plt, W, H, k * (n // parts), n // parts)) for k in range(parts)]
    print("split in %d ranges: %s  sum %.4f max %.4f" % (parts, ["%.4f" % t for t in ts], sum(ts), max(ts)))
frame = [0]


def steal(chunk):
    rd.TraceRaysSteal(plt, W, H, frame[0], chunk)
    frame[0] += 1


for chunk in (32, 128, 512, 2048):
    print("steal chunk %4d   %.4f ms (no flush %.4f)" % (chunk, timed(lambda i: steal(chunk)), timed(lambda i: steal(chunk), False)))
