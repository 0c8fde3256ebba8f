"""Render time of every rank of an 8-way (or G-way) split of the 4K x 16 spp frame for a list of tile shapes, on ONE GPU:
the slowest rank decides the multi-GPU frame.  usage: python scripts/tile_shape_scan.py [G]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth
G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W, H, SPP, N = 3840, 2160, 16, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, SPP); arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
img = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
c = api.Counters(); sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE, counters=c)
best1 = min(c.render_ms for _ in range(3) if not sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE, counters=c))
print("whole image %.3f ms -> ideal per rank %.3f" % (best1, best1 / G))
shapes = [(256, 1), (32, 8), (128, 2), (3840, 1), (3840, 4), (1280, 4), (1280, 8), (1280, 1), (768, 4), (768, 8), (768, 1), (256, 8), (256, 4), (640, 4), (1920, 4), (96, 8), (160, 8), (480, 4), (480, 8)]
for tw, th in shapes:
    tx = (W + tw - 1) // tw
    ms = []
    for r in range(G):
        best = 1e9
        for _ in range(3):
            c = api.Counters(); sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE, tiles=api.TileSet(tw, th, G, r), counters=c); best = min(best, c.render_ms)
        ms.append(best)
    print("  %4dx%-2d tiles/row %3d  max %.3f  mean %.3f  min %.3f   (max/ideal %.3f)" % (tw, th, tx, max(ms), sum(ms) / G, min(ms), max(ms) / (best1 / G)))
