"""Diagnostic: where does near-first traversal differ from the reference order, and why?"""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import helpers as H, oracle_lib as O
from cuda_ray_tracer_b200 import api
o = O.OracleScene(H.scene_path("redchair")); o.build(O.BOUNDS_COMPUTED, 30)
sc = H.gpu_scene_from_oracle(o, O.BOUNDS_COMPUTED); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
a = sc.render(96, 96, 0, 0, want_ids=True); b = sc.render(96, 96, 0, api.TRAVERSE_NEAR_FIRST, want_ids=True)
ys, xs = np.nonzero(a["ids"] != b["ids"])
out = [dict(x=int(x), y=int(y), ref=int(a["ids"][y, x]), near=int(b["ids"][y, x])) for y, x in zip(ys, xs)]
print(out)
json.dump(out, open("gpurun_out/diag_nearfirst.json", "w"))
