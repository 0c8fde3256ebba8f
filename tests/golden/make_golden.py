"""Generates tests/golden/reference_host.json from the REFERENCE's own sources.

Run in the build container only (needs /root/reference, through oracle/_ref/libref_host.so which
oracle/Makefile compiles from the reference tree in place):  python tests/golden/make_golden.py

What is recorded (small: hashes and short vectors):
  * cuRAND XORWOW: state after curand_init(seed, subsequence, 0) and the first raw draws / uniforms —
    cuRAND's own host path, the generator the reference seeds per pixel (draw.cu:105,162)
  * per scene file: sha256 of the arrays parse.cpp produces (reference layouts), of the RGBA image rendered by
    the reference's own shootPrimaryRay path on the host, and of the closest-hit ids of hitNearest, for the
    tree built by the CPU restatement of K1..K6 (the build kernels are __global__ and cannot run on a host)
The oracle (oracle/oracle.cpp) must reproduce every entry bit for bit (tests/test_oracle_cpu.py)."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as O  # noqa: E402
import ref_lib as R  # noqa: E402

# (scene, bounds mode, W, H, aa)
RENDERS = [("tri", O.BOUNDS_SHIPPED, 100, 100, 0), ("tri", O.BOUNDS_COMPUTED, 100, 100, 0), ("tri", O.BOUNDS_COMPUTED, 40, 40, 1),
           ("tri", O.BOUNDS_COMPUTED, 24, 24, 32), ("redchair", O.BOUNDS_COMPUTED, 64, 64, 0), ("redchair", O.BOUNDS_COMPUTED, 20, 20, 32),
           ("redchair", O.BOUNDS_COMPUTED, 24, 24, 20), ("tenthousand", O.BOUNDS_COMPUTED, 40, 40, 0), ("tenthousand", O.BOUNDS_COMPUTED, 12, 12, 40),
           ("spiral", O.BOUNDS_COMPUTED, 32, 32, 0), ("spiral", O.BOUNDS_SHIPPED, 16, 16, 0)]


FEATURE_RENDERS = [(48, 36, 0), (32, 24, 1), (16, 12, 32)]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def rays(n, seed):
    rng = np.random.RandomState(seed)
    org = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    d = rng.uniform(-1, 1, (n, 3)).astype(np.float32) - org
    return org, (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)


def main():
    out = {"curand": [], "parse": {}, "render": [], "trace": []}
    for seed, sub in [(1234, 0), (1234, 1), (1234, 7), (1234, 255), (1234, 256), (1234, 9999), (1234, 8294399), (1234 + 17, 0), (1234 + 17, 31),
                      (1234 + 5000000, 13), (0, 0), (2**40 + 5, 3)]:
        st, raw, uni, nrm = R.curand(seed, sub, 8)
        out["curand"].append({"seed": seed, "subsequence": sub, "state": st.tolist(), "raw": raw.tolist(),
                              "uniform_bits": uni.view(np.uint32).tolist()})
    for name in ("tri", "redchair", "tenthousand", "spiral"):
        p = os.path.join(O.SCENE_DIR, name + ".txt")
        r = R.RefScene(p); a = r.arrays(); ints, fl = r.config()
        out["parse"][name] = {"counts": r.n, "ints": ints, "floats_bits": fl.view(np.uint32).tolist(),
                              "arrays": {k: (sha(v) if not isinstance(v, list) else [sha(x) for x in v]) for k, v in a.items()}}
    for name, mode, W, H, aa in RENDERS:
        p = os.path.join(O.SCENE_DIR, name + ".txt")
        o = O.OracleScene(p); b = o.build(mode, 30)
        r = R.RefScene(p); r.set_bvh(b)
        rgba, wr = r.render(W, H, aa)
        out["render"].append({"scene": name, "bounds": mode, "w": W, "h": H, "aa": aa, "rgba": sha(rgba), "written": int(wr.sum()),
                              "nonzero_alpha": int((rgba[..., 3] > 0).sum())})
    for name, mode in (("tri", O.BOUNDS_SHIPPED), ("redchair", O.BOUNDS_COMPUTED), ("tenthousand", O.BOUNDS_COMPUTED), ("spiral", O.BOUNDS_COMPUTED)):
        p = os.path.join(O.SCENE_DIR, name + ".txt")
        o = O.OracleScene(p); b = o.build(mode, 30)
        r = R.RefScene(p); r.set_bvh(b)
        org, d = rays(4000, 11)
        ids, ts = r.trace_ids(org, d)
        out["trace"].append({"scene": name, "bounds": mode, "n": 4000, "seed": 11, "ids": sha(ids), "t": sha(ts), "hits": int((ids != -1).sum())})
    # generated scene files for the branches no reference scene has: bulbs, fisheye, panorama, DOF on a soup, glass + bulb + GI
    import tempfile
    import feature_scenes as FS
    out["feature"] = []
    with tempfile.TemporaryDirectory() as tmp:
        for kind in FS.KINDS:
            p = FS.write(kind, os.path.join(tmp, kind + ".txt"))
            o = O.OracleScene(p); b = o.build(O.BOUNDS_COMPUTED, 30)
            r = R.RefScene(p); r.set_bvh(b)
            a = r.arrays(); ints, fl = r.config()
            e = {"kind": kind, "counts": r.n, "ints": ints, "floats_bits": fl.view(np.uint32).tolist(),
                 "arrays": {k: (sha(v) if not isinstance(v, list) else [sha(x) for x in v]) for k, v in a.items()}, "render": []}
            for W, H, aa in FEATURE_RENDERS:
                rgba, wr = r.render(W, H, aa)
                e["render"].append({"w": W, "h": H, "aa": aa, "rgba": sha(rgba), "written": int(wr.sum()), "nonzero_alpha": int((rgba[..., 3] > 0).sum())})
            org, d = r.primary_rays(48, 36)
            ids, ts = r.trace_ids(org, d)
            e["primary"] = {"w": 48, "h": 36, "org": sha(org), "dir": sha(d), "ids": sha(ids), "t": sha(ts), "hits": int((ids != -1).sum())}
            out["feature"].append(e)
    json.dump(out, open(os.path.join(HERE, "reference_host.json"), "w"), indent=1)
    print("wrote reference_host.json:", {k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
