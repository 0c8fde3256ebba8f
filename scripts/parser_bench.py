"""Scene-text parse throughput: this library's chunk-parallel parser (1 thread / all threads) against the
reference's own parseInput (oracle/_ref/libref_host.so, src/parse.cpp compiled -O3) on the same file.
CPU only.  usage: python scripts/parser_bench.py [N_TRIANGLES] [out.json]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cuda_ray_tracer_b200 import api, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
out = sys.argv[2] if len(sys.argv) > 2 else None
path = "/tmp/parser_bench_%d.txt" % n
if not os.path.exists(path):
    cfg, a = synth.soup(n, 1, 3840, 2160, 16)
    synth.write_scene_text(path, cfg, a, "soup.png")
size = os.path.getsize(path)
lines = sum(1 for _ in open(path, "rb"))
res = {"file_bytes": size, "lines": lines, "triangles": n, "host_cores": os.cpu_count()}


def best(fn, reps=3):
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
    return min(t)


for th in (1, os.cpu_count()):
    os.environ["RT_PARSE_THREADS"] = str(th)
    s = best(lambda: api.parse_scene_file(path))
    res["b200rt_%dthr_s" % th] = s
    print("libb200rt rt_parse_scene_file, %2d thread(s): %.3f s  %.0f MB/s  %.2f Mlines/s" % (th, s, size / s / 1e6, lines / s / 1e6))
try:
    import ref_lib
    if ref_lib.available():
        s = best(lambda: ref_lib.RefScene(path), reps=2)
        res["reference_parseInput_s"] = s
        print("reference parseInput (1 thread):          %.3f s  %.0f MB/s  %.2f Mlines/s" % (s, size / s / 1e6, lines / s / 1e6))
        res["speedup_all_threads"] = s / res["b200rt_%dthr_s" % os.cpu_count()]; res["speedup_1_thread"] = s / res["b200rt_1thr_s"]
except Exception as e:  # the reference build is optional
    print("reference parser not available:", e)
if out:
    json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res))
