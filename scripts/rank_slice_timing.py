"""Where does a 1/8 image slice lose efficiency against 1/8 of the full frame?  One GPU renders rank r's tiles of an
8-rank split; prints render_ms and the traversal share for the slice and for the whole image."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth, parallel
W, H, SPP, N = 3840, 2160, 16, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, SPP); arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED | (api.BUILD_WIDE4 if os.environ.get('RT_WIDE4') else 0))  # RT_WIDE4=1: the 4-wide tree
img = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
def run(tiles, label):
    best = None
    for _ in range(4):
        c = api.Counters(); sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE | api.TIME_TRAVERSE, tiles=tiles, counters=c)
        if best is None or c.render_ms < best.render_ms: best = c
    print("%-28s render %.3f ms, traverse %.3f ms (%.1f%%), %d rays, %d passes, %d iters, %d launches -> %.0f Mrays/s" % (
        label, best.render_ms, best.traverse_ms, 100 * best.traverse_ms / best.render_ms, best.rays, best.passes, best.wavefront_iters, best.kernel_launches, best.rays / best.render_ms / 1e3))
    return best
full = run(None, "whole image")
for g in (2, 4, 8):
    tw, th = parallel.tile_shape(g, W)
    c = run(api.TileSet(tw, th, g, 0), "rank 0 of %d (%dx%d tiles)" % (g, tw, th))
    print("   vs 1/%d of the whole image: %.3f ms -> slice efficiency %.1f%%" % (g, full.render_ms / g, 100 * full.render_ms / g / c.render_ms))

print("tile shapes at 8 ranks, all ranks (max over ranks decides the frame):")
for tw, th in ((256, 1), (256, 2), (256, 4), (256, 8), (256, 16), (256, 32), (768, 8), (1280, 8)):
    ms = []
    for r in range(8):
        best = 1e9
        for _ in range(2):
            c = api.Counters(); sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE, tiles=api.TileSet(tw, th, 8, r), counters=c); best = min(best, c.render_ms)
        ms.append(best)
    print("  %4dx%-2d  max %.3f  mean %.3f  min %.3f" % (tw, th, max(ms), sum(ms) / 8, min(ms)))
