"""Developer tool / multi-GPU check: torchrun --nproc-per-node N tools/exp_steal.py [scene] [spp] [depth]
Renders one frame of programs/path.cl with dynamic chunk stealing across N GPUs and with the static row
partition, checks both against a single-GPU render on rank 0, prints per-rank chunk counts and timings."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pkg = entry.load_package()
rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
scene = sys.argv[1] if len(sys.argv) > 1 else "cornell"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 4
W, H = 1920, 1080
plt = rd.Platform.GetPlatform(local)
stream = torch.cuda.ExternalStream(plt.stream(), device=local)
bundle = scenes.cornell_like() if scene == "cornell" else scenes.blob_field_scene(int(scene), W, H)
tlas, _ = scenes.build_scene(plt, bundle)
ds = scenes.DeviceScene(plt, bundle, tlas, W, H, scenes.rtprop(0, spp, depth, 0))
n = W * H
image_t = torch.zeros(n * 4, dtype=torch.uint8, device="cuda")
ds.rdImage = rd.WrapBuffer(plt, image_t.data_ptr(), n * 4, owner=image_t)
ds.descriptors[2] = ds.rdImage
plt.set_option("stealing", 1)
pipe, _ = scenes.make_pipeline(plt, "path")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


# single-GPU truth on every rank (rank 0's is used)
rd.TraceRays(plt, 0, 0, 0, W, H)
truth = image_t.clone()
for mode in ("steal", "static"):
    times = []
    for frame in range(6):
        sync()
        with torch.cuda.stream(stream):
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record(stream)
            if mode == "steal":
                image_t.zero_()
                rd.TraceRaysSteal(plt, W, H, frame, int(os.environ.get('RD_CHUNK', '32')))
                e1.record(stream)
                mg.combine_sum(image_t, world, dist)
            else:
                first, count, _, _ = mg.tile_row_range(rank, world, W, H)
                rd.TraceRaysRange(plt, W, H, first, count)
                e1.record(stream)
                mg.gather_rows(image_t, rank, world, W, H, dist)
            e2.record(stream)
        sync()
        t = torch.tensor([e0.elapsed_time(e2), e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(t.cpu().numpy())
    owned = len(rd.StealOwned(plt)) if mode == "steal" else -1
    counts = torch.tensor([owned], device="cuda")
    allc = [torch.zeros_like(counts) for _ in range(world)]
    if world > 1:
        dist.all_gather(allc, counts)
    else:
        allc = [counts]
    if rank == 0:
        med = np.median(np.array(times[2:]), axis=0)
        same = bool(torch.equal(image_t, truth))
        print("%s: world %d frame %.3f ms (render %.3f ms), image == single-GPU: %s, chunks per rank %s" % (
            mode, world, med[0], med[1], same, [int(c) for c in allc]), flush=True)
if world > 1:
    dist.destroy_process_group()
