"""One rank's whole step of the 8-GPU frame on ONE GPU: update_triangles -> build -> render of rank 0's tiles (1/8 of the 4K x 16 spp
frame), as bench.py runs it at N = 8 (RT_RENDER_OVERLAP_BUILD).  CUDA events around the step, mean and best of 12 after 3 warm-ups.
Variants come from the environment (read by the library): RT_MIN_PASSES, RT_TRAV_CTAS_PER_SM, RT_NO_OVERLAP.
usage: [RT_MIN_PASSES=2 RT_TRAV_CTAS_PER_SM=5] python scripts/slice_step_timing.py [label] [ranks]"""
import os, sys, statistics
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth, parallel
label = sys.argv[1] if len(sys.argv) > 1 else "default"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
W, H, SPP, N = 3840, 2160, 16, 1_000_000
cfg, a = synth.soup(N, 1, W, H, SPP)
p0, p1, p2 = a["tri"][:3]
nor, e1, e2 = api.triangle_setup(p0, p1, p2)
a["tri"] = [torch.from_numpy(x).cuda() for x in (p0, p1, p2, nor, e1, e2)]
desc, keep = api.make_desc(cfg, a)
sc = api.Scene(desc, 0)
img = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tw, th = parallel.tile_shape(G, W)
tiles = api.TileSet(tw, th, G, 0) if G > 1 else None
flags = api.SPP_TRUE | api.RENDER_OVERLAP_BUILD
bflags = api.BOUNDS_COMPUTED | (api.BUILD_WIDE4 if os.environ.get("RT_WIDE4") else 0)
ms = []; rms = []
for k in range(15):
    flush.fill_(k); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1_ = torch.cuda.Event(enable_timing=True)
    c = api.Counters()
    e0.record()
    sc.update_triangles(a["tri"][0], a["tri"][1], a["tri"][2])
    sc.build_lbvh_karas(30, bflags, timed=False)
    sc.render_device(W, H, SPP, img, flags=flags, tiles=tiles, counters=c)
    e1_.record(); torch.cuda.synchronize()
    if k >= 3:
        ms.append(e0.elapsed_time(e1_)); rms.append(c.render_ms)
print("%-28s ranks %d  step mean %.3f best %.3f ms | render_ms mean %.3f | passes %d launches %d rays %d | build %.3f" % (
    label, G, statistics.mean(ms), min(ms), statistics.mean(rms), c.passes, c.kernel_launches, c.rays, sc.build_ms()))
