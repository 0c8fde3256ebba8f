"""Shared test helpers: the same host inputs go to the oracle and to the CUDA path."""
import os

import numpy as np

import oracle_lib as O
import cuda_ray_tracer_b200 as rt
from cuda_ray_tracer_b200 import api, synth

SCENES = ["tri", "redchair", "tenthousand", "spiral"]


def scene_path(name):
    p = os.path.join(O.SCENE_DIR, name + ".txt")
    return p if os.path.exists(p) else None


def oracle_from_arrays(cfg, arrays):
    o = O.OracleScene()
    o.set_config(**{k: (np.asarray(v, np.float32) if isinstance(v, (tuple, list)) else v) for k, v in cfg.items()})
    refs = arrays["refs"]
    t = arrays.get("tri"); sc = arrays.get("sphere_c")
    # primitives must be added in file order so that refs agree
    if t is not None and sc is None:
        o.add_triangles(t[0], t[1], t[2], arrays["tri_mat"])
    elif sc is not None and t is None:
        o.add_spheres(sc, arrays["sphere_r"], arrays["sphere_mat"])
    else:
        ti = si = 0
        for ty, _ in refs:
            if ty == 1:
                o.add_triangles(t[0][ti:ti + 1], t[1][ti:ti + 1], t[2][ti:ti + 1], arrays["tri_mat"][ti:ti + 1]); ti += 1
            else:
                o.add_spheres(sc[si:si + 1], arrays["sphere_r"][si:si + 1], arrays["sphere_mat"][si:si + 1]); si += 1
    for s in arrays.get("suns", []):
        o.add_sun(s[:3], s[3:])
    for b in arrays.get("bulbs", []):
        o.add_bulb(b[:3], b[3:])
    for p in arrays.get("planes", []):
        o.add_plane(p[:4], p[10:21])
    return o


def gpu_scene_from_oracle(o, bounds_mode=O.BOUNDS_SHIPPED, device=0):
    """rt_scene_create from the oracle's host arrays (identical inputs on both sides)."""
    cfg = o.config(); arrays = o.arrays()
    mn, mx = o.bounds(bounds_mode)
    desc, keep = api.make_desc(cfg, arrays, mn, mx)
    return api.Scene(desc, device)


def mixed_scene(n, seed=3):
    """triangles and spheres interleaved, with duplicates and a degenerate cluster (equal Morton codes)."""
    cfg, a = synth.soup(n, seed, 64, 64, 0)
    a = synth.finish_triangles(a)
    rng = np.random.RandomState(seed)
    ns = n // 3
    sc = rng.uniform(-1, 1, (ns, 3)).astype(np.float32); sr = rng.uniform(0.01, 0.05, ns).astype(np.float32)
    sc[: ns // 4] = sc[0]  # coincident centres -> equal codes
    sm = np.zeros((ns, 11), np.float32); sm[:, :3] = rng.uniform(0.2, 1, (ns, 3)); sm[:, 9] = 1.458
    kinds = np.array([1] * n + [0] * ns, np.int32); rng.shuffle(kinds)
    ids = np.zeros_like(kinds); ids[kinds == 1] = np.arange(n); ids[kinds == 0] = np.arange(ns)
    a.update(sphere_c=sc, sphere_r=sr, sphere_mat=sm, refs=np.stack([kinds, ids], 1).astype(np.int32))
    return cfg, a
