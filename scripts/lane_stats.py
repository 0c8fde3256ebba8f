import os, sys
sys.path.insert(0, "/root/repo")
import torch
from cuda_ray_tracer_b200 import api, synth
W, H, SPP, N = 3840, 2160, 16, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, SPP); arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
img = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
c = api.Counters(); sc.render_device(W, H, SPP, img, flags=api.SPP_TRUE | api.COUNTERS, counters=c)
print(c.rays, c.nodes_visited, c.leaf_tests)
