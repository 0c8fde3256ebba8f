"""Synthetic scenes of the named benchmark sizes: counter-based random triangle soups and
sphere fields, emitted either as reference-layout arrays (for rt_scene_create) or as a TinyRaytracer
text scene (for the reference binary and the parser).  Every value is a pure function of
(seed, primitive index, component): SplitMix64 → top 32 bits → u * 2^-32 as float32."""
import numpy as np

from . import api

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def uniform01(seed, idx, comp):
    """float32 in [0,1): hash of (seed, idx, comp)."""
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + idx.astype(np.uint64) * np.uint64(64) + np.uint64(comp)) & _M64
        h = _splitmix(_splitmix(key))
    return ((h >> np.uint64(32)).astype(np.float64) * 2.0 ** -32).astype(np.float32)


def default_config(width, height, aa, **kw):
    cfg = dict(width=width, height=height, color=(1, 1, 1), bounces=4, aa=aa, dof_focus=0.0, dof_lens=0.0, forward=(0, 0, -1),
               right=(1, 0, 0), up=(0, 1, 0), eye=(0, 0, 0), target_up=(0, 1, 0), expose=float("inf"), fisheye=0, panorama=0, ior=1.458,
               rough=0.0, gi=0, trans=(0, 0, 0), shine=(0, 0, 0))
    cfg.update(kw)
    return cfg


def _materials(seed, idx):
    m = np.zeros((len(idx), 11), np.float32)
    for k in range(3):
        m[:, k] = np.float32(0.25) + np.float32(0.75) * uniform01(seed, idx, 16 + k)
    m[:, 9] = 1.458
    return m


def soup(n, seed=1, width=1920, height=1080, aa=16):
    """n random triangles: centre uniform in [-1,1]^3, two edge vectors uniform in [-s,s]^3, s = 2 n^(-1/3);
    eye (0,0,3.5) looking down -z, one white sun (1,1,1), no plane; generation (random) order."""
    idx = np.arange(n)
    s = np.float32(2.0 * n ** (-1.0 / 3.0))
    c = np.stack([np.float32(2) * uniform01(seed, idx, k) - np.float32(1) for k in range(3)], 1)
    e1 = np.stack([(np.float32(2) * uniform01(seed, idx, 3 + k) - np.float32(1)) * s for k in range(3)], 1)
    e2 = np.stack([(np.float32(2) * uniform01(seed, idx, 6 + k) - np.float32(1)) * s for k in range(3)], 1)
    p0 = c.astype(np.float32); p1 = (c + e1).astype(np.float32); p2 = (c + e2).astype(np.float32)
    cfg = default_config(width, height, aa, eye=(0, 0, 3.5))
    arrays = dict(tri=[p0, p1, p2, None, None, None], tri_mat=_materials(seed, idx),
                  refs=np.stack([np.ones(n, np.int32), idx.astype(np.int32)], 1),
                  suns=np.array([[1, 1, 1, 1, 1, 1]], np.float32))
    return cfg, arrays


def spheres(n, seed=1, width=1920, height=1080, aa=16):
    """n spheres, centres uniform in [-1,1]^3, radius 0.25 n^(-1/3)."""
    idx = np.arange(n)
    c = np.stack([np.float32(2) * uniform01(seed, idx, k) - np.float32(1) for k in range(3)], 1).astype(np.float32)
    r = np.full(n, 0.25 * n ** (-1.0 / 3.0), np.float32)
    cfg = default_config(width, height, aa, eye=(0, 0, 3.5))
    arrays = dict(sphere_c=c, sphere_r=r, sphere_mat=_materials(seed, idx),
                  refs=np.stack([np.zeros(n, np.int32), idx.astype(np.int32)], 1), suns=np.array([[1, 1, 1, 1, 1, 1]], np.float32))
    return cfg, arrays


def finish_triangles(arrays):
    """Fill nor/e1/e2 with the Triangle constructor arithmetic (host C++, rt_triangle_setup)."""
    t = arrays.get("tri")
    if t is not None and t[3] is None:
        nor, e1, e2 = api.triangle_setup(t[0], t[1], t[2])
        arrays["tri"] = [t[0], t[1], t[2], nor, e1, e2]
    return arrays


def write_scene_text(path, cfg, arrays, name="out.png"):
    """TinyRaytracer text of a synthetic scene (xyz x3 + `tri -3 -2 -1` per triangle, parse.cpp format).
    %.9g round-trips float32 exactly through stof."""
    with open(path, "w") as f:
        f.write("png %d %d %s\n" % (cfg["width"], cfg["height"], name))
        f.write("eye %.9g %.9g %.9g\n" % tuple(cfg["eye"]))
        f.write("forward %.9g %.9g %.9g\n" % tuple(cfg["forward"]))
        if cfg["aa"]:
            f.write("aa %d\n" % cfg["aa"])
        if cfg["bounces"] != 4:
            f.write("bounces %d\n" % cfg["bounces"])
        for s in arrays.get("suns", []):
            f.write("color %.9g %.9g %.9g\nsun %.9g %.9g %.9g\n" % (s[3], s[4], s[5], s[0], s[1], s[2]))
        t = arrays.get("tri")
        if t is not None:
            m = arrays["tri_mat"]
            fmt = "color %.9g %.9g %.9g\nxyz %.9g %.9g %.9g\nxyz %.9g %.9g %.9g\nxyz %.9g %.9g %.9g\ntri -3 -2 -1\n"
            rows = np.concatenate([m[:, :3], t[0], t[1], t[2]], axis=1).astype(np.float64)  # float32 -> double is exact, %.9g round-trips it
            for lo in range(0, len(rows), 200_000):  # in pieces: a 10M-triangle scene is 1.8 GB of text
                f.write("".join([fmt % tuple(r) for r in rows[lo:lo + 200_000].tolist()]))
        if arrays.get("sphere_c") is not None:
            m = arrays["sphere_mat"]; c = arrays["sphere_c"]; r = arrays["sphere_r"]
            f.write("".join("color %.9g %.9g %.9g\nsphere %.9g %.9g %.9g %.9g\n" % (m[i, 0], m[i, 1], m[i, 2], *c[i], r[i]) for i in range(len(c))))
