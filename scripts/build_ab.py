"""LBVH build time A/B: update_triangles -> rt_build_lbvh on the soup at 1M and 10M triangles, best and median of 12
(CUDA events inside rt_build_lbvh).  usage: RT_LIB=.../libb200rt_x.so python scripts/build_ab.py [label]"""
import os, statistics, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_ray_tracer_b200 import api, synth
label = sys.argv[1] if len(sys.argv) > 1 else os.path.basename(api.LIB_PATH)
for n in (100_000, 1_000_000, 10_000_000):
    cfg, a = synth.soup(n, 1, 64, 64, 0)
    p0, p1, p2 = a["tri"][:3]
    nor, e1, e2 = api.triangle_setup(p0, p1, p2)
    a["tri"] = [torch.from_numpy(x).cuda() for x in (p0, p1, p2, nor, e1, e2)]
    desc, keep = api.make_desc(cfg, a)
    sc = api.Scene(desc, 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    # climb variants of the fused builder: warp-first (default), every level global (round 1), and - timing only, no tree -
    # the leaf phase alone (RT_BU_LEAFONLY), which splits the kernel into gather/records and climb
    only = os.environ.get("BUILD_AB_MODES")  # e.g. "default" when comparing libraries (RT_LIB) rather than switches
    for mode, fl, env in (("default", 0, None), ("no pdl", 0, "RT_NO_PDL"), ("no peek", 0, "RT_BU_NO_PEEK"), ("no compact", 0, "RT_BU_NO_COMPACT"), ("global", api.BUILD_GLOBAL_CLIMB, None), ("leaf-only", 0, "RT_BU_LEAFONLY")):
        if only and mode not in only.split(","): continue
        for e in ("RT_BU_LEAFONLY", "RT_BU_NO_COMPACT", "RT_BU_NO_PEEK", "RT_NO_PDL"): os.environ.pop(e, None)
        if env: os.environ[env] = "1"
        for bits in (30, 63):
            ms = []
            for k in range(12):
                sc.update_triangles(a["tri"][0], a["tri"][1], a["tri"][2]); flush.fill_(k); torch.cuda.synchronize()
                ms.append(sc.build_lbvh_karas(bits, api.BOUNDS_COMPUTED | fl))
            warm = []
            for k in range(6):  # the frame loop's case: the vertex update has just written the records
                sc.update_triangles(a["tri"][0], a["tri"][1], a["tri"][2])
                warm.append(sc.build_lbvh_karas(bits, api.BOUNDS_COMPUTED | fl))
            print("%-6s %-12s n=%8d bits=%d  build after L2 flush: best %.3f median %.3f ms | right after update_triangles: best %.3f ms" % (
                label, mode, n, bits, min(ms), statistics.median(ms), min(warm)))
    for e in ("RT_BU_LEAFONLY", "RT_BU_NO_COMPACT", "RT_BU_NO_PEEK", "RT_NO_PDL"): os.environ.pop(e, None)
    sc.close(); del sc, a
