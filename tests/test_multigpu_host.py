"""Host-side logic of the multi-GPU path under gloo on CPU, world size 2 (no GPU needed): the static row
partition covers the frame exactly once and its send/recv gather rebuilds the frame on rank 0; the
sum-combine of the stealing partition rebuilds a frame whose chunks were split arbitrarily."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 64, 36


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    mg = entry.load_package().multigpu
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    truth = torch.arange(W * H * 4, dtype=torch.int64).remainder(251).to(torch.uint8)
    # static partition
    img = torch.zeros(W * H * 4, dtype=torch.uint8)
    first, count, y0, ny = mg.tile_row_range(rank, world, W, H)
    assert first == y0 * W and count == ny * W
    img[first * 4:(first + count) * 4] = truth[first * 4:(first + count) * 4]
    mg.gather_rows(img, rank, world, W, H, dist)
    ok_static = bool(torch.equal(img, truth)) if rank == 0 else True
    # stealing partition: chunks assigned by a fixed pseudo-random rule, each to exactly one rank
    chunk_items = 96
    n_chunks = (W * H + chunk_items - 1) // chunk_items
    img = torch.zeros(W * H * 4, dtype=torch.uint8)
    mine = [c for c in range(n_chunks) if (c * 7 + 3) % world == rank]
    for c in mine:
        px = torch.from_numpy(mg.chunk_pixels(c, chunk_items, W, H))
        idx = (px[:, None] * 4 + torch.arange(4)[None, :]).reshape(-1)
        img[idx] = truth[idx]
    mg.combine_sum(img, world, dist)
    ok_steal = bool(torch.equal(img, truth)) if rank == 0 else True
    out[rank] = (ok_static, ok_steal, len(mine))
    dist.barrier()
    dist.destroy_process_group()


def test_partitions_and_gathers_world_size_2():
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, 29500 + os.getpid() % 2000, out), nprocs=world, join=True)
        res = dict(out)
    assert res[0][0] and res[0][1], res
    assert sum(v[2] for v in res.values()) == (W * H + 95) // 96


def test_row_partition_covers_every_row_once():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    mg = entry.load_package().multigpu
    for world in (1, 2, 3, 4, 8):
        for h in (1080, 2160, 36, 6):
            rows = np.zeros(h, np.int32)
            for r in range(world):
                _, _, y0, ny = mg.tile_row_range(r, world, 1920, h)
                rows[y0:y0 + ny] += 1
            assert (rows == 1).all(), (world, h)
            for band in (1, 2, 7):
                rows = np.zeros(h - h % 4, np.int32)
                for r in range(world):
                    for y0, ny in mg.interleaved_rows(r, world, h, band):
                        rows[y0:y0 + ny] += 1
                assert (rows == 1).all(), (world, h, band)
    # the chunk -> pixel map is a bijection of the frame
    px = np.concatenate([mg.chunk_pixels(c, 1024, 64, 36) for c in range((64 * 36 + 1023) // 1024)])
    assert np.array_equal(np.sort(px), np.arange(64 * 36))
