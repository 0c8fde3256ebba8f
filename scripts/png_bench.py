"""PNG encode time of a 4K RGBA8 frame: row-block-parallel encoder at 1 thread / all threads, deflate levels 4 (default)
and 6 (what a single compress2() call, the previous writer, used).  CPU only.  usage: python scripts/png_bench.py [out.json]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from cuda_ray_tracer_b200 import api

H, W = 2160, 3840
rng = np.random.RandomState(0)
img = np.zeros((H, W, 4), np.uint8); img[..., 3] = 255
for i in range(20000):  # soup-like content: flat-coloured patches over black, plus sampling noise
    x, y = rng.randint(0, W), rng.randint(0, H); s = rng.randint(4, 60)
    img[y:y + s, x:x + s, :3] = rng.randint(0, 255, 3)
img[..., :3] += rng.randint(0, 3, (H, W, 3)).astype(np.uint8)
res = {"image": "%dx%d RGBA8, synthetic patches + noise" % (W, H), "host_cores": os.cpu_count(), "runs": []}
for threads, level in ((1, 6), (1, 4), (os.cpu_count(), 6), (os.cpu_count(), 4)):
    os.environ["RT_PNG_THREADS"] = str(threads); os.environ["RT_PNG_LEVEL"] = str(level)
    t = []
    for _ in range(3):
        t0 = time.perf_counter(); api.png_write("/tmp/png_bench.png", img); t.append(time.perf_counter() - t0)
    ok = bool(np.array_equal(api.png_read("/tmp/png_bench.png"), img))
    res["runs"].append({"threads": threads, "level": level, "seconds": min(t), "file_bytes": os.path.getsize("/tmp/png_bench.png"), "roundtrip_ok": ok})
    print(res["runs"][-1])
if len(sys.argv) > 1:
    json.dump(res, open(sys.argv[1], "w"), indent=1)
