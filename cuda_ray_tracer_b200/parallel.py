"""Multi-GPU plumbing: scene replicated, image split into interleaved tiles, NCCL (or gloo in tests) only for the
scene broadcast and the framebuffer gather.  No data-path collective inside a frame."""
import numpy as np

from . import api

TILE_W, TILE_H = 32, 8


def tile_shape(world, width=None):
    """tile shape for `world` ranks: rt_tile_shape (csrc/comm.cu) — 8-row tiles whose count per image row shares the
    fewest factors with world, so that no rank owns whole columns of tiles (256x8 at 3840 wide: a diagonal pattern)."""
    if width is None or world <= 1:
        return TILE_W, TILE_H
    return api.tile_shape(world, width)


def tile_set(world, rank, tile_w=None, tile_h=None, width=None):
    """rt_tile_set for `rank` of `world` (None for a single rank = whole image)"""
    if world <= 1:
        return None
    if tile_w is None or tile_h is None:
        tile_w, tile_h = tile_shape(world, width)
    return api.TileSet(tile_w, tile_h, world, rank)


def tile_owner(w, h, world, tile_w=TILE_W, tile_h=TILE_H):
    """[h][w] int32 rank that renders each pixel: tile t (row-major over ceil(w/tw) x ceil(h/th) tiles) -> t % world.
    Host mirror of slot_to_pixel / pixel_to_slot in csrc/render.cu."""
    tw, th = min(tile_w, w), min(tile_h, h)
    tiles_x = (w + tw - 1) // tw
    ys, xs = np.mgrid[0:h, 0:w]
    return (((ys // th) * tiles_x + xs // tw) % world).astype(np.int32)


def broadcast_scene(tensors, src=0):
    """replicate the scene arrays (torch tensors on each rank's device) from rank `src`"""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        for t in tensors:
            dist.broadcast(t, src)


def upload_sharded(host_pinned, dev_full, dev_part, rank, world):
    """host -> every GPU, for a replicated array that all ranks hold in pinned host memory: each rank copies only
    its 1/world slice over its own PCIe link and NCCL all-gathers the slices over NVLink (36 MB of vertices at 8
    GPUs: 4.5 MB of PCIe per rank instead of 36).  dev_full: [world * rows_per_rank, ...] device tensor (rows past the
    array's length are padding), dev_part: [rows_per_rank, ...] staging.  Returns dev_full."""
    import torch.distributed as dist
    rows = dev_part.shape[0]
    lo = rank * rows; hi = min(lo + rows, host_pinned.shape[0])
    if hi > lo:
        dev_part[: hi - lo].copy_(host_pinned[lo:hi], non_blocking=True)
    dist.all_gather_into_tensor(dev_full, dev_part)
    return dev_full


_comm = None


def comm(device=None):
    """the process's rt_comm over torch.distributed's ranks (one GPU per process), created on first use"""
    global _comm
    if _comm is None:
        import torch
        _comm = api.Comm.from_torch_distributed(torch.cuda.current_device() if device is None else device)
    return _comm


def gather_framebuffer(img, dst=0, width=None, height=None, tile_w=None, tile_h=None):
    """Every rank rendered only its own tiles into a zeroed image; afterwards rank `dst` holds the whole image.
    CUDA tensors: rt_framebuffer_gather (each rank packs its tiles and sends 1/world of the image over NCCL).
    Host tensors (the gloo tests of the partition logic): the sum over uint8 is the gather."""
    import torch.distributed as dist
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return img
    if img.is_cuda:
        h, w = (img.shape[0], img.shape[1]) if height is None else (height, width)
        if tile_w is None or tile_h is None:
            tile_w, tile_h = tile_shape(dist.get_world_size(), w)
        comm(img.device.index).gather_framebuffer([img], w, h, tile_w, tile_h, dst)
    else:
        dist.reduce(img, dst, op=dist.ReduceOp.SUM)
    return img
