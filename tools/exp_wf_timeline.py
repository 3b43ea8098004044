"""Developer tool: per-stage device times (option wf_timeline) of one wavefront frame of programs/path.cl on the C3 scene.
usage: python tools/exp_wf_timeline.py [n_grid] [spp] [depth] [world]   (world > 1: the share of rank 3 of an interleaved world-GPU frame)"""
import os
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
n_grid = int(sys.argv[1]) if len(sys.argv) > 1 else 2237
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 5
world = int(sys.argv[4]) if len(sys.argv) > 4 else 1
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
bundle = scenes.blob_field_scene(n_grid, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, spp, depth, 0))
pipe, module = scenes.make_pipeline(plt, "path")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
plt.set_option("wavefront", 1)
plt.set_option("wf_timeline", 1)
plt.set_option("wf_dynamic", int(os.environ.get("RD_WF_DYNAMIC", "1")))
for rep in range(3):
    ds.set_rtprop(0, spp, depth)
    if world == 1:
        rd.TraceRays(plt, 0, 0, 0, W, H)
    else:
        rd._check(rd.lib().rd_trace_rays_interleaved(plt.handle, W, H, 3, world, 1))
        plt.sync()
    tl = rd.WavefrontTimeline(plt)
print("stages in launch order (ms):")
print("  " + "  ".join("%s %.2f" % (n, ms) for n, ms in tl))
tot = OrderedDict()
for n, ms in tl:
    tot[n] = tot.get(n, 0.0) + ms
print("per stage kind:", "  ".join("%s %.2f" % kv for kv in tot.items()), " sum %.2f ms" % sum(tot.values()))
