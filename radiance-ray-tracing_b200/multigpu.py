"""Host-side plumbing of the multi-GPU path: one process per GPU, the scene replicated,
the frame sharded by screen tiles, ONE collective per frame (the framebuffer gather to rank 0).

Three partitions:
  interleave  bands of 8x4-tile rows dealt round robin (rd_trace_rays_interleaved): a static partition that
           spreads the expensive screen regions over all ranks; frame = byte-wise SUM over ranks (NCCL reduce)
  static   contiguous runs of 8x4-tile rows per rank (rd_trace_rays_range) + NCCL send/recv of each rank's rows
  steal    dynamic chunk stealing from one counter shared over CUDA IPC / NVLink (rd_trace_rays_steal); every
           rank zeroes its framebuffer, renders the chunks it wins, and the frame is the byte-wise SUM over
           ranks (each pixel is written by exactly one rank), reduced to rank 0 with NCCL
The functions take torch tensors and a process group, so the same code runs under gloo on CPU in the tests.
"""
import numpy as np


def tile_row_range(rank, world, width, height, tile_h=4):
    """Contiguous run of tile rows owned by `rank`: (first work-item, count, first pixel row, rows)."""
    tile_rows = height // tile_h
    r0, r1 = tile_rows * rank // world, tile_rows * (rank + 1) // world
    if rank == world - 1:
        r1 = tile_rows
    y0, y1 = r0 * tile_h, (height if rank == world - 1 else r1 * tile_h)
    return y0 * width, (y1 - y0) * width, y0, y1 - y0


def interleaved_rows(rank, world, height, band_tile_rows=1, tile_h=4):
    """Pixel-row runs (y0, rows) owned by `rank` under the interleaved band partition."""
    tile_rows, out, b = height // tile_h, [], rank
    while b * band_tile_rows < tile_rows:
        r0 = b * band_tile_rows
        r1 = min(r0 + band_tile_rows, tile_rows)
        out.append((r0 * tile_h, (r1 - r0) * tile_h))
        b += world
    return out


def gather_rows(image, rank, world, width, height, dist, channels=4):
    """Static partition: rank r sends its rows of the flat (H*W*channels) uint8 framebuffer to rank 0."""
    if world == 1:
        return
    ops = []
    if rank == 0:
        for r in range(1, world):
            _, _, y0, ny = tile_row_range(r, world, width, height)
            if ny:
                ops.append(dist.P2POp(dist.irecv, image[y0 * width * channels:(y0 + ny) * width * channels], r))
    else:
        _, _, y0, ny = tile_row_range(rank, world, width, height)
        if ny:
            ops.append(dist.P2POp(dist.isend, image[y0 * width * channels:(y0 + ny) * width * channels], 0))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def combine_sum(image, world, dist):
    """Stealing partition: frame = sum over ranks of framebuffers that are zero outside the owned chunks."""
    if world > 1:
        dist.reduce(image, dst=0, op=dist.ReduceOp.SUM)


def exchange_steal_handle(rd, plt, rank, world, dist, device):
    """Rank 0 exports its chunk counter, the others import it (64-byte CUDA IPC handle, one broadcast)."""
    import torch
    if world == 1 or getattr(plt, "_steal_exchanged", False):
        return
    plt._steal_exchanged = True
    h = torch.zeros(64, dtype=torch.uint8, device=device)
    if rank == 0:
        h.copy_(torch.from_numpy(rd.StealCounterExport(plt)))
    dist.broadcast(h, src=0)
    if rank != 0:
        rd.StealCounterImport(plt, h.cpu().numpy())


def chunk_pixels(chunk, chunk_items, width, height, tile=True):
    """Pixel indices (row-major) of the work-items of one stolen chunk (8x4 tile remap of cl_compat.h)."""
    t = np.arange(chunk * chunk_items, min((chunk + 1) * chunk_items, width * height), dtype=np.int64)
    if not (tile and width % 8 == 0 and height % 4 == 0):
        return t
    tiles_x = width // 8
    tl, lane = t >> 5, t & 31
    ty, tx = tl // tiles_x, tl % tiles_x
    return ((ty << 2) + (lane >> 3)) * width + (tx << 3) + (lane & 7)
