"""Are the sporadic +7..+17 ms steps seen in bench.py host-side or GPU-side?  40 traverse launches of 24 M rays each are
queued back to back with no host interaction (events only); then the same launches with a host synchronize after each."""
import os, sys, json, time
import numpy as np, torch
ROOT = os.getcwd(); sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth
W, H, N = 3840, 2160, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, 0); arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
dev = torch.device("cuda", 0)
n = 8_000_000
g = torch.Generator(device=dev); g.manual_seed(1)
d = torch.randn(n, 3, device=dev, generator=g); d[:, 2] = -d[:, 2].abs() - 2.0; d = (d / d.norm(dim=1, keepdim=True)).contiguous()
o = (torch.rand(n, 3, device=dev, generator=g) * 0.6 - 0.3); o[:, 2] = 3.5; o = o.contiguous()
ids = torch.empty(n, dtype=torch.int32, device=dev); ts = torch.empty(n, dtype=torch.float32, device=dev)
for _ in range(3): sc.trace_device(n, o, d, ids, ts, flags=0, counters=None)
torch.cuda.synchronize()
K = 60
ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
ev[0].record()
for k in range(K):
    sc.trace_device(n, o, d, ids, ts, flags=0, counters=None); ev[k + 1].record()
torch.cuda.synchronize()
a = [round(ev[k].elapsed_time(ev[k + 1]), 3) for k in range(K)]
b = []
for k in range(K):
    torch.cuda.synchronize(); t = time.perf_counter()
    sc.trace_device(n, o, d, ids, ts, flags=0, counters=None); torch.cuda.synchronize()
    b.append(round((time.perf_counter() - t) * 1e3, 3))
print(json.dumps({"queued_back_to_back_ms": a, "host_sync_each_ms": b, "median": [float(np.median(a)), float(np.median(b))], "max": [max(a), max(b)]}))
