"""Generated scene FILES for the branches no scene of the reference exercises (none has a `bulb`, `fisheye`
or `panorama` line): getColorBulb and the bulb shadow rays (src/draw.cu:358-374,649-659), the two projections of
Ray::Ray (src/struct.cu:24-41), depth of field on a soup, and a shiny + transparent + rough + GI scene lit by a bulb.

They are TEXT in the reference's own format (src/parse.cpp:16-222), so the same file goes through the reference's
parser + CUDA program (oracle/_ref/ref_gpu_parity), the reference's sources compiled for the host
(oracle/_ref/libref_host.so), the CPU oracle and this library's parser.  Deterministic: numpy RandomState(seed)."""
import numpy as np


def _f(*v):
    return " ".join("%.9g" % float(np.float32(x)) for x in v)


def _soup_lines(rng, n, scale, with_spheres=True, materials=False, rough=False):
    out = []
    for i in range(n):
        out.append("color " + _f(*rng.uniform(0.2, 1.0, 3)))
        if materials:
            k = rng.randint(0, 6)
            out.append("shininess " + (_f(0.6) if k == 1 else _f(0.3, 0.2, 0.4) if k == 3 else _f(0)))
            out.append("transparency " + (_f(0.7) if k == 2 else _f(0.5, 0.6, 0.4) if k == 3 else _f(0)))
            out.append("ior " + _f(1.3 + 0.1 * (k % 3)))
            if rough:
                out.append("roughness " + _f(0.05 if k == 4 else 0))
        c = rng.uniform(-1, 1, 3)
        if with_spheres and i % 3 == 0:
            out.append("sphere " + _f(*c, rng.uniform(0.03, 0.12)))
        else:
            for _ in range(3):
                out.append("xyz " + _f(*(c + rng.uniform(-scale, scale, 3))))
            out.append("tri -3 -2 -1")
    return out


def scene_text(kind, w=96, h=72, name="out.png"):
    rng = np.random.RandomState({"bulbs": 1, "fisheye": 2, "panorama": 3, "dof": 4, "glassbulb": 5, "bulbonly": 6}[kind])
    L = ["png %d %d %s" % (w, h, name)]
    if kind == "bulbs":  # two bulbs (one inside the cloud) + a sun, plane, exposure
        L += ["eye 0 0.2 3.2", "forward 0 -0.05 -1", "expose 1.7",
              "color 1 0.9 0.8", "sun 1 2 1.5", "color 3 2.5 2", "bulb 0.3 1.1 2.5", "color 1.5 2 3", "bulb -0.2 0.1 0.3",
              "color 0.6 0.6 0.7", "plane 0 1 0 1.3"]
        L += _soup_lines(rng, 400, 0.25)
    elif kind == "bulbonly":  # no sun at all: every shadow ray has a distance limit
        L += ["eye 0 0 3.0", "color 4 4 4", "bulb 0.5 0.8 2.0", "color 2 1 1", "bulb -1.5 -0.5 1.0"]
        L += _soup_lines(rng, 300, 0.3)
    elif kind == "fisheye":
        L += ["eye 0 0 1.6", "fisheye", "color 1 1 1", "sun 1 1 1", "color 2 2 2", "bulb 0 0 1.2"]
        L += _soup_lines(rng, 500, 0.2)
    elif kind == "panorama":
        L += ["eye 0.1 0.05 0.2", "panorama", "up 0.1 1 0", "color 1 1 1", "sun 1 1 1", "color 0.5 0.5 0.6", "plane 0 1 0 1.1"]
        L += _soup_lines(rng, 500, 0.2)
    elif kind == "dof":
        L += ["eye 0 0 3.0", "dof 2.5 0.08", "aa 32", "color 1 1 1", "sun 1 1 1"]
        L += _soup_lines(rng, 400, 0.25)
    elif kind == "glassbulb":  # mirrors, glass, both, rough surfaces, GI, lit by a bulb and a sun
        L += ["eye 0 0 3.2", "bounces 5", "gi 2", "expose 1.5", "color 1 1 0.9", "sun 1 2 1.5", "color 3 2.5 2", "bulb 0.3 1.1 2.5",
              "color 0.6 0.6 0.7", "shininess 0.2", "plane 0 1 0 1.3", "shininess 0"]
        L += _soup_lines(rng, 350, 0.3, materials=True, rough=True)
    else:
        raise KeyError(kind)
    return "\n".join(L) + "\n"


KINDS = ["bulbs", "bulbonly", "fisheye", "panorama", "dof", "glassbulb"]


def write(kind, path, **kw):
    with open(path, "w") as f:
        f.write(scene_text(kind, **kw))
    return path
