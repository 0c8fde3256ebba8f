"""Fixed cost of one traverse launch: rt_trace over the first n primary rays of the 4K frame, n = 1M .. 24M, L2 warm
(back-to-back) and cold (256 MiB flush before each launch).  T(n) = a + b n: a is what every launch pays whatever
its size (ramp-up from a cold L2 + the drain of the longest rays at the end)."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth
W, H, N = 3840, 2160, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, 0); arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED | (api.BUILD_WIDE4 if os.environ.get('RT_WIDE4') else 0))  # RT_WIDE4=1: the 4-wide tree
dev = torch.device("cuda", 0)
# primary rays of the frame in row-major pixel order (struct.cu:16-62, no jitter), 3 samples per pixel -> 24.9M rays
ys, xs = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32), indexing="ij")
m = float(max(W, H))
def rays(jx, jy):
    sx = (2 * (xs + jx) - W) / m; sy = (H - 2 * (ys + jy)) / m
    d = torch.stack([sx, sy, -torch.ones_like(sx)], -1).reshape(-1, 3)
    return d / d.norm(dim=1, keepdim=True)
# only the central band of the image (rows H/4 .. 3H/4: every ray crosses the soup), in interleaved row order so that any
# prefix of the array is a uniform sample of the band: per-ray cost does not depend on n
band = torch.arange(H // 4, 3 * H // 4, device=dev)
perm_rows = torch.cat([band[k::8] for k in range(8)])
def pick(t):
    return t.reshape(H, W, 3)[perm_rows].reshape(-1, 3)
d = torch.cat([pick(rays(jx, jy)) for jx, jy in ((0, 0), (0.3, -0.2), (-0.4, 0.1), (0.1, 0.4), (-0.2, -0.3), (0.45, 0.25))], 0).contiguous()
o = torch.tensor([0.0, 0.0, 3.5], device=dev).repeat(len(d), 1).contiguous()
ids = torch.empty(len(d), dtype=torch.int32, device=dev); ts = torch.empty(len(d), dtype=torch.float32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {"rays_total": len(d), "warm": [], "cold": []}
for mode in ("warm", "cold"):
    for n in (1, 2, 4, 8, 16, 24):
        n *= 1_000_000
        best = 1e9
        for rep in range(4):
            if mode == "cold": flush.fill_(rep)
            c = api.Counters(); sc.trace_device(n, o, d, ids, ts, flags=0, counters=c); torch.cuda.synchronize()
            best = min(best, c.traverse_ms)
        out[mode].append((n, round(best, 4)))
    x = np.array([a for a, _ in out[mode]], float); y = np.array([b for _, b in out[mode]])
    b, a = np.polyfit(x, y, 1)
    out[mode + "_fit"] = {"fixed_ms_per_launch": round(float(a), 4), "ms_per_Mray": round(float(b * 1e6), 5)}
print(json.dumps(out))
