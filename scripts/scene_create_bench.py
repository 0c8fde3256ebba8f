"""Wall time of rt_scene_create (marshalling) for the 1M-triangle soup from pageable host arrays,
pinned host arrays and device arrays.  usage: [RT_LIB=...] python scripts/scene_create_bench.py [n]"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from cuda_ray_tracer_b200 import api, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
api.load_library()
torch.zeros(1, device="cuda"); torch.cuda.synchronize()
cfg, a = synth.soup(n, 1, 64, 64, 0); a = synth.finish_triangles(a)
out = {"n": n, "lib": os.path.basename(api.LIB_PATH)}
def timed(arrays, label):
    ts = []
    for _ in range(4):
        desc, keep = api.make_desc(cfg, arrays)
        torch.cuda.synchronize(); t = time.perf_counter()
        sc = api.Scene(desc, 0)
        torch.cuda.synchronize(); ts.append((time.perf_counter() - t) * 1e3)
        sc.close()
    out[label + "_ms"] = round(min(ts[1:]), 3)
timed(a, "pageable")
pin = dict(a); pin["tri"] = [torch.from_numpy(x).pin_memory() for x in a["tri"]]; pin["tri_mat"] = torch.from_numpy(a["tri_mat"]).pin_memory(); pin["refs"] = torch.from_numpy(a["refs"]).pin_memory()
timed(pin, "pinned")
dev = dict(a); dev["tri"] = [torch.from_numpy(x).cuda() for x in a["tri"]]; dev["tri_mat"] = torch.from_numpy(a["tri_mat"]).cuda(); dev["refs"] = torch.from_numpy(a["refs"]).cuda()
timed(dev, "device")
print(json.dumps(out))
