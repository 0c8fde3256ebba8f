"""Experiment: how much of the memory-bound shade/generate time hides behind the issue-bound traversal when two
halves of the image are rendered concurrently on ONE GPU (two scenes, two host threads, two streams)?
Also: does a 2-D tile order of the pixel slots help a single render?"""
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth  # noqa: E402

W, H, SPP, N = 3840, 2160, 16, 1_000_000
cfg, arrays = synth.soup(N, 1, W, H, SPP)
arrays = synth.finish_triangles(arrays)
desc, keep = api.make_desc(cfg, arrays)
dev = torch.device("cuda", 0)
flags = api.SPP_TRUE


def make(pool):
    os.environ["RT_POOL_BYTES"] = str(pool)
    sc = api.Scene(desc, 0); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    return sc


def timed(fn, reps=4):
    fn(); torch.cuda.synchronize()
    t = time.time()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.time() - t) / reps * 1e3


full = make(6 << 30)
img = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
cnt = api.Counters()
print("single render, row-major slots      : %.2f ms" % timed(lambda: full.render_device(W, H, SPP, img, flags=flags, counters=cnt)), cnt.rays)
for tw, th in ():
    ts = api.TileSet(tw, th, 1, 0)
    print("single render, %2dx%-2d tiles           : %.2f ms" % (tw, th, timed(lambda: full.render_device(W, H, SPP, img, flags=flags, tiles=ts, counters=cnt))))

for K in (2, 3):
    scs = [make((6 << 30) // K) for _ in range(K)]
    imgs = [torch.zeros((H, W, 4), dtype=torch.uint8, device=dev) for _ in range(K)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(K)]

    def run_all():
        def work(k):
            scs[k].render_device(W, H, SPP, imgs[k], flags=flags, tiles=api.TileSet(32, 8, K, k), stream=streams[k].cuda_stream)
        th = [threading.Thread(target=work, args=(k,)) for k in range(K)]
        for t in th: t.start()
        for t in th: t.join()
    print("%d concurrent renders of 1/%d image each: %.2f ms" % (K, K, timed(run_all)))
    # same split, one after the other (what the tile split alone costs)
    def run_seq():
        for k in range(K):
            scs[k].render_device(W, H, SPP, imgs[k], flags=flags, tiles=api.TileSet(32, 8, K, k), stream=streams[k].cuda_stream)
    print("%d sequential renders of 1/%d image each: %.2f ms" % (K, K, timed(run_seq)))
    del scs, imgs
