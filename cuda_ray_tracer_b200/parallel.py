"""Multi-GPU plumbing: scene replicated, image split into interleaved tiles, NCCL (or gloo in tests) only for the
scene broadcast and the framebuffer gather.  No data-path collective inside a frame."""
import numpy as np

from . import api

TILE_W, TILE_H = 32, 8


def tile_set(world, rank, tile_w=TILE_W, tile_h=TILE_H):
    """rt_tile_set for `rank` of `world` (None for a single rank = whole image)"""
    return api.TileSet(tile_w, tile_h, world, rank) if world > 1 else None


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


def gather_framebuffer(img, dst=0):
    """sum the per-rank images onto rank `dst`: every rank wrote only its own tiles into a zeroed image, so the
    sum over uint8 is the gather (33 MB at 4K: tens of microseconds over NVLink)"""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(img, dst, op=dist.ReduceOp.SUM)
    return img
