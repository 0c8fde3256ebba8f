"""Traversal micro-benchmark: 1M-triangle soup, camera rays of the headline frame at reduced size.
Prints per-lane traverse-kernel time (CUDA events inside rt_render) and whole-render time.
RT_LIB=<path> selects an experiment build of the library."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from cuda_ray_tracer_b200 import api, synth
n = int(os.environ.get("TB_TRIS", 1000000)); W, H, spp = 1920, 1080, int(os.environ.get("TB_SPP", 8))
cfg, a = synth.soup(n, 1, W, H, spp); a = synth.finish_triangles(a)
desc, keep = api.make_desc(cfg, a)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
img = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
out = {"lib": os.path.basename(api.LIB_PATH)}
for name, fl in (("exact", 0), ("near", api.TRAVERSE_NEAR_FIRST), ("ref", api.TRAVERSE_REF_ORDER)):
    if name == "ref" and os.environ.get("TB_SKIP_REF"): continue
    best = None
    for rep in range(3):
        c = api.Counters()
        sc.render_device(W, H, spp, img, flags=api.SPP_TRUE | api.TIME_TRAVERSE | fl, counters=c)
        if best is None or c.traverse_ms < best[0]: best = (c.traverse_ms, c.render_ms, c.rays)
    out[name] = {"traverse_ms": round(best[0], 3), "render_ms": round(best[1], 3), "Mrays/s(traverse)": round(best[2] / best[0] / 1e3, 1)}
    out.setdefault("checksum", {})[name] = int(img.to(torch.int64).sum().item())
print(json.dumps(out))
