"""Three builds of the 1M soup (update_triangles -> rt_build_lbvh), meant to run under
`ncu --metrics gpu__time_duration.sum` for a per-kernel launch list of the build chain (profiles/launches_build_r02.csv)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth
cfg, a = synth.soup(1_000_000, 1, 64, 64, 0)
p0, p1, p2 = a["tri"][:3]
nor, e1, e2 = api.triangle_setup(p0, p1, p2)
a["tri"] = [torch.from_numpy(x).cuda() for x in (p0, p1, p2, nor, e1, e2)]
desc, keep = api.make_desc(cfg, a)
sc = api.Scene(desc, 0)
for k in range(3):
    sc.update_triangles(a["tri"][0], a["tri"][1], a["tri"][2])
    print("build %d: %.3f ms" % (k, sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)))
