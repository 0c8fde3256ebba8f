"""ctypes bindings of the CPU oracle (oracle/liboracle.so) and of the reference-derived host
checker (oracle/_ref/libref_host.so).  TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's
cpu_baseline / reference legs import this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
SCENE_DIR = os.path.join(REF_DIR, "scenes")

BOUNDS_SHIPPED, BOUNDS_COMPUTED = 0, 1
SPP_REF, SPP_TRUE = 0, 1


class Counters(C.Structure):
    _fields_ = [("rays", C.c_ulonglong), ("nodes_visited", C.c_ulonglong), ("leaf_tests", C.c_ulonglong),
                ("shaded_hits", C.c_ulonglong), ("max_stack", C.c_int)]


def build_oracle():
    """Compile oracle/liboracle.so (and oracle/_ref when /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"], check=True)
    if os.path.exists("/root/reference/src/draw.cu"):
        subprocess.run(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        p = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(p):
            build_oracle()
        L = C.CDLL(p)
        vp, fp, ip, up = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint32)
        L.o_scene_parse.restype = vp; L.o_scene_parse.argtypes = [C.c_char_p]
        L.o_scene_new.restype = vp
        L.o_scene_free.argtypes = [vp]
        L.o_scene_filename.restype = C.c_char_p; L.o_scene_filename.argtypes = [vp]
        for name in ("o_scene_sphere_c", "o_scene_sphere_r", "o_scene_sphere_mat", "o_scene_tri_mat", "o_scene_suns",
                     "o_scene_bulbs", "o_scene_planes", "o_node_bbox"):
            getattr(L, name).restype = fp; getattr(L, name).argtypes = [vp]
        L.o_scene_tri.restype = fp; L.o_scene_tri.argtypes = [vp, C.c_int]
        for name in ("o_scene_refs", "o_sorted_refs", "o_node_parent"):
            getattr(L, name).restype = ip; getattr(L, name).argtypes = [vp]
        for name in ("o_codes30", "o_perm", "o_node_left", "o_node_right"):
            getattr(L, name).restype = up; getattr(L, name).argtypes = [vp]
        L.o_codes63.restype = C.POINTER(C.c_uint64); L.o_codes63.argtypes = [vp]
        L.o_scene_counts.argtypes = [vp] + [C.POINTER(C.c_int)] * 6
        L.o_scene_get_config.argtypes = [vp, C.POINTER(C.c_int), fp]
        L.o_scene_set_config.argtypes = [vp, C.POINTER(C.c_int), fp]
        L.o_scene_bounds.argtypes = [vp, C.c_int, fp, fp]
        L.o_scene_add_triangles.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.o_scene_add_spheres.argtypes = [vp, C.c_int, vp, vp, vp]
        L.o_scene_add_sun.argtypes = [vp, vp, vp]; L.o_scene_add_bulb.argtypes = [vp, vp, vp]
        L.o_scene_add_plane.argtypes = [vp, vp, vp]
        L.o_triangle_setup.argtypes = [C.c_int] + [vp] * 6
        L.o_build.argtypes = [vp, C.c_int, C.c_int]; L.o_num_nodes.argtypes = [vp]
        L.o_morton_unsorted.argtypes = [vp, C.c_int, C.c_int, vp]
        L.o_xorwow_init.argtypes = [C.c_ulonglong, C.c_ulonglong, vp]
        L.o_xorwow_draws.argtypes = [vp, C.c_int, vp]
        L.o_trace.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.POINTER(Counters)]
        L.o_render.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp,
                               C.POINTER(Counters)]
        _lib = L
    return _lib


def _np(ptr, shape, dtype):
    n = int(np.prod(shape))
    if n == 0:
        return np.zeros(shape, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).view(dtype).reshape(shape).copy()


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleScene:
    """Host scene in the reference's SoA layout + oracle build/trace/render."""

    def __init__(self, path=None):
        L = lib()
        if path is None:
            self.h = C.c_void_p(L.o_scene_new())
        else:
            h = L.o_scene_parse(path.encode())
            if not h:
                raise ValueError("oracle could not parse " + path)
            self.h = C.c_void_p(h)
        self.built = None

    def __del__(self):
        try:
            lib().o_scene_free(self.h)
        except Exception:
            pass

    # ---- config ----
    def config(self):
        ints = (C.c_int * 8)(); fl = (C.c_float * 32)()
        lib().o_scene_get_config(self.h, ints, fl)
        f = np.array(fl[:], np.float32)
        return dict(width=ints[0], height=ints[1], bounces=ints[2], aa=ints[3], gi=ints[4], fisheye=ints[5],
                    panorama=ints[6], color=f[0:3], dof_focus=f[3], dof_lens=f[4], forward=f[5:8], right=f[8:11],
                    up=f[11:14], eye=f[14:17], target_up=f[17:20], expose=f[20], ior=f[21], rough=f[22],
                    trans=f[23:26], shine=f[26:29], filename=lib().o_scene_filename(self.h).decode())

    def set_config(self, **kw):
        c = self.config(); c.update(kw)
        ints = (C.c_int * 8)(c["width"], c["height"], c["bounces"], c["aa"], c["gi"], int(c["fisheye"]), int(c["panorama"]), 0)
        f = np.zeros(32, np.float32)
        f[0:3] = c["color"]; f[3] = c["dof_focus"]; f[4] = c["dof_lens"]; f[5:8] = c["forward"]; f[8:11] = c["right"]
        f[11:14] = c["up"]; f[14:17] = c["eye"]; f[17:20] = c["target_up"]; f[20] = c["expose"]; f[21] = c["ior"]
        f[22] = c["rough"]; f[23:26] = c["trans"]; f[26:29] = c["shine"]
        lib().o_scene_set_config(self.h, ints, f.ctypes.data_as(C.POINTER(C.c_float)))

    def counts(self):
        v = [C.c_int() for _ in range(6)]
        lib().o_scene_counts(self.h, *[C.byref(x) for x in v])
        return dict(zip(("spheres", "triangles", "prims", "suns", "bulbs", "planes"), [x.value for x in v]))

    def arrays(self):
        """Reference-layout host arrays (what copyConfigDataToDevice uploads, config_utils.cu:72-199)."""
        L = lib(); n = self.counts()
        return dict(
            sphere_c=_np(L.o_scene_sphere_c(self.h), (n["spheres"], 3), np.float32),
            sphere_r=_np(L.o_scene_sphere_r(self.h), (n["spheres"],), np.float32),
            sphere_mat=_np(L.o_scene_sphere_mat(self.h), (n["spheres"], 11), np.float32),
            tri=[_np(L.o_scene_tri(self.h, k), (n["triangles"], 3), np.float32) for k in range(6)],
            tri_mat=_np(L.o_scene_tri_mat(self.h), (n["triangles"], 11), np.float32),
            refs=_np(L.o_scene_refs(self.h), (n["prims"], 2), np.int32),
            suns=_np(L.o_scene_suns(self.h), (n["suns"], 6), np.float32),
            bulbs=_np(L.o_scene_bulbs(self.h), (n["bulbs"], 6), np.float32),
            planes=_np(L.o_scene_planes(self.h), (n["planes"], 21), np.float32))

    def bounds(self, mode):
        mn = np.zeros(3, np.float32); mx = np.zeros(3, np.float32)
        fp = C.POINTER(C.c_float)
        lib().o_scene_bounds(self.h, mode, mn.ctypes.data_as(fp), mx.ctypes.data_as(fp))
        return mn, mx

    # ---- construction ----
    def add_triangles(self, p0, p1, p2, mat):
        p0, p1, p2, mat = [np.ascontiguousarray(a, np.float32) for a in (p0, p1, p2, mat)]
        lib().o_scene_add_triangles(self.h, len(p0), _p(p0), _p(p1), _p(p2), _p(mat))

    def add_spheres(self, c, r, mat):
        c, r, mat = [np.ascontiguousarray(a, np.float32) for a in (c, r, mat)]
        lib().o_scene_add_spheres(self.h, len(c), _p(c), _p(r), _p(mat))

    def add_sun(self, d, col=(1, 1, 1)):
        d = np.asarray(d, np.float32); col = np.asarray(col, np.float32)
        lib().o_scene_add_sun(self.h, _p(d), _p(col))

    def add_bulb(self, p, col=(1, 1, 1)):
        p = np.asarray(p, np.float32); col = np.asarray(col, np.float32)
        lib().o_scene_add_bulb(self.h, _p(p), _p(col))

    def add_plane(self, abcd, mat):
        a = np.asarray(abcd, np.float32); m = np.ascontiguousarray(mat, np.float32)
        lib().o_scene_add_plane(self.h, _p(a), _p(m))

    # ---- build ----
    def morton_unsorted(self, mode, bits=30):
        out = np.zeros(self.counts()["prims"], np.uint64)
        lib().o_morton_unsorted(self.h, mode, bits, _p(out))
        return out

    def build(self, mode, bits=30):
        L = lib()
        rc = L.o_build(self.h, mode, bits)
        if rc != 0:
            raise RuntimeError("o_build rc=%d" % rc)
        n = self.counts()["prims"]; nn = 2 * n - 1
        self.built = dict(
            n=n, bits=bits, mode=mode,
            codes=_np(L.o_codes30(self.h), (n,), np.uint32) if bits == 30 else _np(L.o_codes63(self.h), (n,), np.uint64),
            perm=_np(L.o_perm(self.h), (n,), np.uint32),
            refs=_np(L.o_sorted_refs(self.h), (n, 2), np.int32),
            bbox=_np(L.o_node_bbox(self.h), (nn, 6), np.float32),
            left=_np(L.o_node_left(self.h), (max(n - 1, 0),), np.uint32),
            right=_np(L.o_node_right(self.h), (max(n - 1, 0),), np.uint32),
            parent=_np(L.o_node_parent(self.h), (nn,), np.int32))
        return self.built

    # ---- trace / render ----
    def trace(self, org, dirs):
        org = np.ascontiguousarray(org, np.float32); dirs = np.ascontiguousarray(dirs, np.float32)
        n = len(org); ids = np.zeros(n, np.int32); ts = np.zeros(n, np.float32); cnt = Counters()
        lib().o_trace(self.h, n, _p(org), _p(dirs), _p(ids), _p(ts), C.byref(cnt))
        return ids, ts, cnt

    def render(self, w=-1, h=-1, aa=-1, spp_mode=SPP_REF, y0=-1, y1=-1, threads=None, linear=False):
        c = self.config()
        W = c["width"] if w < 0 else w; H = c["height"] if h < 0 else h
        rgba = np.zeros((H, W, 4), np.uint8); ids = np.full((H, W), -9, np.int32); wr = np.zeros((H, W), np.uint8)
        lin = np.zeros((H, W, 4), np.float32) if linear else None
        cnt = Counters()
        lib().o_render(self.h, w, h, aa, spp_mode, y0, y1, threads or (os.cpu_count() or 1), _p(rgba), _p(ids), _p(wr),
                       _p(lin), C.byref(cnt))
        out = dict(rgba=rgba, ids=ids, written=wr, counters=cnt)
        if linear:
            out["linear"] = lin
        return out


def xorwow_init(seed, subseq):
    st = np.zeros(6, np.uint32)
    lib().o_xorwow_init(seed, subseq, _p(st))
    return st


def xorwow_draws(state6, n):
    st = np.array(state6, np.uint32); out = np.zeros(n, np.uint32)
    lib().o_xorwow_draws(_p(st), n, _p(out))
    return out, st


def triangle_setup(p0, p1, p2):
    p0, p1, p2 = [np.ascontiguousarray(a, np.float32) for a in (p0, p1, p2)]
    nor, e1, e2 = [np.zeros_like(p0) for _ in range(3)]
    lib().o_triangle_setup(len(p0), _p(p0), _p(p1), _p(p2), _p(nor), _p(e1), _p(e2))
    return nor, e1, e2
