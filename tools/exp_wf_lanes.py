"""Developer tool: wavefront frame time of programs/path.cl with one and two lanes (option wf_lanes), on the whole frame and on
a 1/8 share of it (what one rank of an 8-GPU job renders: rd_trace_rays_interleaved with world = 8).
usage: python tools/exp_wf_lanes.py [n_grid] [spp] [depth]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
n_grid = int(sys.argv[1]) if len(sys.argv) > 1 else 2237
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 5
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
stream = torch.cuda.ExternalStream(plt.stream(), device=0)
bundle = scenes.blob_field_scene(n_grid, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, spp, depth, 0))
pipe, module = scenes.make_pipeline(plt, "path")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
plt.set_option("wavefront", 1)
L = rd.lib()
base = {}
for world in (1, 8):
    for lanes in (1, 2, 3, 4):
        plt.set_option("wf_lanes", lanes)
        ms = []
        for it in range(5):
            ds.set_rtprop(0, spp, depth)
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                if world == 1:
                    rd.TraceRaysAsync(plt, W, H)
                else:
                    rd._check(L.rd_trace_rays_interleaved(plt.handle, W, H, 3, world, 1))
                e1.record(stream)
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        img = ds.read_image()
        same = base.setdefault(world, img) is img or np.array_equal(base[world], img)
        print("share 1/%d  lanes %d  %.3f ms (min %.3f)  image %s" % (world, lanes, float(np.median(ms[1:])), min(ms[1:]), "same" if same else "DIFFERENT"), flush=True)
