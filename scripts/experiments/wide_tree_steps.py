"""Dumps the reference LBVH of soup(N) (CPU oracle), the triangles and a sample of the 4K frame's primary rays, and runs
wide_tree_steps.cpp on them.  usage: python scripts/experiments/wide_tree_steps.py [N] [rays]   (CPU only)"""
import os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import helpers as H, oracle_lib as O
from cuda_ray_tracer_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 400_000
W, Hh = 3840, 2160
cfg, a = synth.soup(n, 1, W, Hh, 0); a = synth.finish_triangles(a)
o = H.oracle_from_arrays(cfg, a); b = o.build(O.BOUNDS_COMPUTED, 30)
rng = np.random.RandomState(1)
x = rng.uniform(0, W, m).astype(np.float32); y = rng.uniform(0, Hh, m).astype(np.float32)
mx = np.float32(max(W, Hh)); sx = (2 * x - W) / mx; sy = (Hh - 2 * y) / mx
d = np.stack([sx, sy, -np.ones_like(sx)], 1).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
org = np.tile(np.array([0, 0, 3.5], np.float32), (m, 1))
rays = np.concatenate([org, d.astype(np.float32)], 1).astype(np.float32)
tri = np.concatenate([a["tri"][0], a["tri"][1], a["tri"][2]], 1).astype(np.float32)
here = os.path.dirname(os.path.abspath(__file__)); exe = os.path.join(tempfile.gettempdir(), "wide_tree_steps")
subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(here, "wide_tree_steps.cpp")], check=True)
path = os.path.join(tempfile.gettempdir(), "wide_tree_%d.bin" % n)
with open(path, "wb") as f:
    np.array([n, m], np.int32).tofile(f)
    b["left"].astype(np.uint32).tofile(f); b["right"].astype(np.uint32).tofile(f); b["perm"].astype(np.uint32).tofile(f)
    b["bbox"].astype(np.float32).tofile(f); tri.tofile(f); rays.tofile(f)
print(subprocess.run([exe, path], capture_output=True, text=True, check=True).stdout.strip())
os.remove(path)
