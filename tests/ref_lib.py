"""ctypes bindings of oracle/_ref/libref_host.so: the reference's own parse / traversal / shading
sources compiled for the host (oracle/Makefile).  TEST INFRASTRUCTURE; exists only where
oracle/_ref has been built (this container, or shipped prebuilt to the GPU box)."""
import ctypes as C
import os

import numpy as np

from oracle_lib import REF_DIR, _np, _p

_lib = None


def available():
    return os.path.exists(os.path.join(REF_DIR, "libref_host.so"))


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(REF_DIR, "libref_host.so"))
        vp, fp = C.c_void_p, C.POINTER(C.c_float)
        L.ref_load.restype = vp; L.ref_load.argtypes = [C.c_char_p]
        L.ref_free.argtypes = [vp]
        L.ref_new.restype = vp; L.ref_new.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
        L.ref_add_triangles.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.ref_add_sun.argtypes = [vp, vp, vp]
        L.ref_finish.argtypes = [vp]
        L.ref_counts.argtypes = [vp, C.POINTER(C.c_int)]
        L.ref_config.argtypes = [vp, C.POINTER(C.c_int), fp]
        L.ref_tri.restype = fp; L.ref_tri.argtypes = [vp, C.c_int]
        for n in ("ref_tri_mat", "ref_sphere_c", "ref_sphere_r", "ref_sphere_mat", "ref_planes", "ref_suns", "ref_bulbs"):
            getattr(L, n).restype = fp; getattr(L, n).argtypes = [vp]
        L.ref_refs.restype = C.POINTER(C.c_int32); L.ref_refs.argtypes = [vp]
        L.ref_set_bvh.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.ref_render.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
        L.ref_trace_ids.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.ref_primary_rays.argtypes = [vp, C.c_int, C.c_int, vp, vp]
        L.ref_curand.argtypes = [C.c_ulonglong, C.c_ulonglong, vp, C.c_int, vp, vp, vp]
        _lib = L
    return _lib


class RefScene:
    def __init__(self, path=None, handle=None):
        self.h = C.c_void_p(lib().ref_load(path.encode())) if handle is None else handle
        c = (C.c_int * 6)(); lib().ref_counts(self.h, c)
        self.n = dict(zip(("spheres", "triangles", "prims", "suns", "bulbs", "planes"), list(c)))

    def __del__(self):
        try:
            lib().ref_free(self.h)
        except Exception:
            pass

    def config(self):
        ints = (C.c_int * 8)(); fl = (C.c_float * 32)()
        lib().ref_config(self.h, ints, fl)
        return list(ints), np.array(fl[:], np.float32)

    def arrays(self):
        L = lib(); n = self.n
        return dict(
            sphere_c=_np(L.ref_sphere_c(self.h), (n["spheres"], 3), np.float32),
            sphere_r=_np(L.ref_sphere_r(self.h), (n["spheres"],), np.float32),
            sphere_mat=_np(L.ref_sphere_mat(self.h), (n["spheres"], 11), np.float32),
            tri=[_np(L.ref_tri(self.h, k), (n["triangles"], 3), np.float32) for k in range(6)],
            tri_mat=_np(L.ref_tri_mat(self.h), (n["triangles"], 11), np.float32),
            refs=_np(L.ref_refs(self.h), (n["prims"], 2), np.int32),
            suns=_np(L.ref_suns(self.h), (n["suns"], 6), np.float32),
            bulbs=_np(L.ref_bulbs(self.h), (n["bulbs"], 6), np.float32),
            planes=_np(L.ref_planes(self.h), (n["planes"], 21), np.float32))

    def set_bvh(self, built):
        refs = np.ascontiguousarray(built["refs"], np.int32); bbox = np.ascontiguousarray(built["bbox"], np.float32)
        left = np.ascontiguousarray(built["left"], np.uint32); right = np.ascontiguousarray(built["right"], np.uint32)
        lib().ref_set_bvh(self.h, built["n"], _p(refs), _p(bbox), _p(left), _p(right))

    def render(self, W, H, aa, y0=0, y1=None, threads=None):
        rgba = np.zeros((H, W, 4), np.uint8); wr = np.zeros((H, W), np.uint8)
        lib().ref_render(self.h, W, H, aa, y0, H if y1 is None else y1, threads or (os.cpu_count() or 1), _p(rgba), _p(wr))
        return rgba, wr

    def primary_rays(self, W, H):
        o = np.zeros((W * H, 3), np.float32); d = np.zeros((W * H, 3), np.float32)
        lib().ref_primary_rays(self.h, W, H, _p(o), _p(d))
        return o, d

    def trace_ids(self, org, dirs):
        org = np.ascontiguousarray(org, np.float32); dirs = np.ascontiguousarray(dirs, np.float32)
        ids = np.zeros(len(org), np.int32); ts = np.zeros(len(org), np.float32)
        lib().ref_trace_ids(self.h, len(org), _p(org), _p(dirs), _p(ids), _p(ts))
        return ids, ts


def scene_from_arrays(cfg, arrays):
    """triangle-only synthetic scene built with the reference's own constructors (bench.py --impl reference)"""
    L = lib()
    eye = np.asarray(cfg["eye"], np.float32); fwd = np.asarray(cfg["forward"], np.float32)
    h = C.c_void_p(L.ref_new(int(cfg["width"]), int(cfg["height"]), int(cfg["bounces"]), int(cfg["aa"]), _p(eye), _p(fwd)))
    t = arrays["tri"]; m = np.ascontiguousarray(arrays["tri_mat"], np.float32)
    p0, p1, p2 = [np.ascontiguousarray(a, np.float32) for a in t[:3]]
    L.ref_add_triangles(h, len(p0), _p(p0), _p(p1), _p(p2), _p(m))
    for s in arrays.get("suns", []):
        s = np.asarray(s, np.float32)
        L.ref_add_sun(h, _p(s[:3].copy()), _p(s[3:].copy()))
    L.ref_finish(h)
    return RefScene(handle=h)


def curand(seed, subseq, n):
    st = np.zeros(6, np.uint32); raw = np.zeros(n, np.uint32); uni = np.zeros(n, np.float32); nrm = np.zeros(n, np.float32)
    lib().ref_curand(seed, subseq, _p(st), n, _p(raw), _p(uni), _p(nrm))
    return st, raw, uni, nrm
