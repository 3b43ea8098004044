"""Developer tool: one wavefront frame of programs/path.cl on the C3 scene (for ncu launch lists).
usage: python tools/exp_wf_profile.py [n_grid] [spp] [depth] [sort]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
n_grid = int(sys.argv[1]) if len(sys.argv) > 1 else 2237
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 5
sort = int(sys.argv[4]) if len(sys.argv) > 4 else 1
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(0)
bundle = scenes.blob_field_scene(n_grid, W, H)
tlas, blas = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, spp, depth, 0))
pipe, module = scenes.make_pipeline(plt, "path")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
plt.set_option("wavefront", 1)
plt.set_option("wf_sort", sort)
plt.set_option("wf_dynamic", int(os.environ.get("RD_WF_DYNAMIC", "1")))
plt.set_option("wf_ray_sort", int(os.environ.get("RD_WF_RAYSORT", "0")))
if "RD_WF_PATHS" in os.environ:
    plt.set_option("wf_paths", int(os.environ["RD_WF_PATHS"]))
import time
for _ in range(3):
    ds.set_rtprop(0, spp, depth)
    t0 = time.perf_counter()
    rd.TraceRays(plt, 0, 0, 0, W, H)
    print("frame %.2f ms" % ((time.perf_counter() - t0) * 1e3))
print("done", ds.read_image()[..., :3].mean())
