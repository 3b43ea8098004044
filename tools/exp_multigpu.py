"""Multi-GPU correctness check (developer tool, run under torchrun on N GPUs of one box):
   torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/exp_multigpu.py
Every partition + C-ABI collective pair renders the frame rank 0 renders alone, bit for bit (inline) or within the
wavefront-vs-inline bar: interleaved bands + rd_frame_gather (RGBA8 and float scratch), contiguous ranges +
rd_frame_gather_ranges, dynamic work units + rd_frame_gather_units (inline and wavefront mode)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pkg = entry.load_package()
rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
plt = rd.Platform.GetPlatform(local)
uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    uid.copy_(torch.from_numpy(rd.CommUniqueId()))
dist.broadcast(uid, src=0)
rd.CommInit(plt, rank, world, uid.cpu().numpy())
mg.exchange_steal_handle(rd, plt, rank, world, dist, "cuda")

b = scenes.cornell_like()
b.camera["x"], b.camera["y"] = 0.013, 1.037
tlas, _ = scenes.build_scene(plt, b)
w, h, spp, depth = 640, 480, 2, 4
ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, spp, depth, 0))
pipe, _ = scenes.make_pipeline(plt, "path")
rd.BindPipeline(plt, pipe)
rd.BindDescriptorSet(plt, ds.descriptors)
rd.TraceRays(plt, 0, 0, 0, w, h)
whole, whole_scr = ds.read_image(), ds.read_scratch()
ok = True


def clear():
    rd.WriteBuffer(plt, ds.rdImage, whole.nbytes, np.zeros_like(whole))
    rd.WriteBuffer(plt, ds.rdImageScratch, whole_scr.nbytes, np.zeros_like(whole_scr))
    ds.set_rtprop(0, spp, depth)
    plt.sync()
    dist.barrier()


def check(name, exact=True, scratch=False):
    global ok
    plt.sync()
    dist.barrier()
    if rank == 0:
        img = ds.read_image()
        if exact:
            good = np.array_equal(img, whole)
        else:
            d = img[..., :3].astype(np.float64) - whole[..., :3].astype(np.float64)
            good = np.sqrt((d * d).mean()) <= 0.5
        if scratch:
            good = good and np.array_equal(ds.read_scratch()[..., :3], whole_scr[..., :3])
        print("%-70s %s" % (name, "ok" if good else "MISMATCH"), flush=True)
        ok = ok and good


frame = 5000
for band in (1, 3):
    clear()
    rd.TraceRaysInterleaved(plt, w, h, rank, world, band)
    rd.FrameGather(plt, ds.rdImage, w, h, 4, band)
    rd.FrameGather(plt, ds.rdImageScratch, w, h, 16, band)
    check("inline, interleaved bands of %d + rd_frame_gather (image + scratch)" % band, scratch=True)
clear()
firsts = [mg.tile_row_range(r, world, w, h)[0] for r in range(world)]
counts = [mg.tile_row_range(r, world, w, h)[1] for r in range(world)]
rd.TraceRaysRange(plt, w, h, firsts[rank], counts[rank])
rd.FrameGatherRanges(plt, ds.rdImage, 4, firsts, counts)
check("inline, contiguous ranges + rd_frame_gather_ranges")
for units in (world, 4 * world, 64):
    clear()
    frame += 1
    owned = rd.TraceRaysUnits(plt, w, h, units, frame)
    rd.FrameGatherUnits(plt, ds.rdImage, w, h, 4, units, owned)
    check("inline, %d dynamic units + rd_frame_gather_units (rank %d took %d)" % (units, rank, bin(owned).count("1")))
plt.set_option("wavefront", 1)
clear()
rd.TraceRaysInterleaved(plt, w, h, rank, world, 1)
rd.FrameGather(plt, ds.rdImage, w, h, 4, 1)
check("wavefront, interleaved bands + rd_frame_gather", exact=False)
for units in (2 * world, 64):
    clear()
    frame += 1
    owned = rd.TraceRaysUnits(plt, w, h, units, frame)
    rd.FrameGatherUnits(plt, ds.rdImage, w, h, 4, units, owned)
    check("wavefront, %d dynamic units + rd_frame_gather_units" % units, exact=False)
plt.set_option("wavefront", 0)
rd.CommDestroy(plt)
dist.destroy_process_group()
if rank == 0:
    print("MULTIGPU", "PASS" if ok else "FAIL", flush=True)
sys.exit(0 if ok else 1)
