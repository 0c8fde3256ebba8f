"""Host-side mirror of the reference's interface over the C ABI (include/rt_api.h).

Reference call sequence (main.cu:25-94) and what it maps to here:
    parseInput(argv, stl)                  -> parse_scene_file(path) -> SceneDesc
    initRawConfigFromStl + copyConfigDataToDevice -> Scene(desc, device)
    build_lbvh_karas(raw, 30)              -> Scene.build_lbvh_karas(morton_bits=30, ...)
    render(d_image, w, h, aa, d_cfg)       -> Scene.render(w, h, aa, ...)
    Image.save(name)                       -> png_write(name, rgba)
Nothing here computes: every call lands in libb200rt.so, and load_library() raises if it is missing.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_LIB") or os.path.join(_HERE, "libb200rt.so")  # RT_LIB: experiment builds (Makefile LIB=...)

BOUNDS_FROM_DESC, BOUNDS_COMPUTED, BUILD_TOPDOWN, BUILD_LOCAL_CLIMB, BUILD_WIDE4, BUILD_GLOBAL_CLIMB = 0, 1, 2, 4, 8, 16
SPP_REF, SPP_TRUE, TRAVERSE_NEAR_FIRST, ROUGH_ZYX, COUNTERS, TIME_TRAVERSE, TRAVERSE_REF_ORDER, RENDER_OVERLAP_BUILD, RENDER_SERIAL_PASSES = 0, 1, 2, 4, 8, 16, 32, 64, 128
TRAVERSE_PACKET = 256


class RtError(RuntimeError):
    pass


class SceneDesc(C.Structure):
    _f3 = C.c_float * 3
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("color", C.c_float * 3), ("bounces", C.c_int32), ("aa", C.c_int32),
        ("dof_focus", C.c_float), ("dof_lens", C.c_float),
        ("forward", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("eye", C.c_float * 3),
        ("target_up", C.c_float * 3), ("expose", C.c_float), ("fisheye", C.c_int32), ("panorama", C.c_int32),
        ("ior", C.c_float), ("rough", C.c_float), ("gi", C.c_int32), ("trans", C.c_float * 3), ("shine", C.c_float * 3),
        ("scene_min", C.c_float * 3), ("scene_max", C.c_float * 3),
        ("num_spheres", C.c_int32), ("sphere_c", C.c_void_p), ("sphere_r", C.c_void_p), ("sphere_mat", C.c_void_p),
        ("num_triangles", C.c_int32), ("tri_p0", C.c_void_p), ("tri_p1", C.c_void_p), ("tri_p2", C.c_void_p),
        ("tri_nor", C.c_void_p), ("tri_e1", C.c_void_p), ("tri_e2", C.c_void_p), ("tri_mat", C.c_void_p),
        ("num_prims", C.c_int32), ("refs", C.c_void_p),
        ("num_suns", C.c_int32), ("suns", C.c_void_p), ("num_bulbs", C.c_int32), ("bulbs", C.c_void_p),
        ("num_planes", C.c_int32), ("planes", C.c_void_p),
    ]


class TileSet(C.Structure):
    _fields_ = [("tile_w", C.c_int32), ("tile_h", C.c_int32), ("num_ranks", C.c_int32), ("rank", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("samples", C.c_uint64), ("nodes_visited", C.c_uint64), ("leaf_tests", C.c_uint64),
                ("shaded_hits", C.c_uint64), ("wavefront_iters", C.c_uint32), ("passes", C.c_uint32), ("render_ms", C.c_float),
                ("kernel_launches", C.c_uint32), ("traverse_ms", C.c_float), ("traverse_launches", C.c_uint32),
                ("fallback_rays", C.c_uint64), ("d2h_ms", C.c_float), ("reserved0", C.c_uint32), ("packet_steps", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None

EXPORTS = ["rt_scene_create", "rt_scene_destroy", "rt_scene_update_triangles", "rt_build_lbvh", "rt_scene_build_ms", "rt_scene_begin_frame", "rt_render", "rt_render_host",
           "rt_trace", "rt_export_reference_layout", "rt_scene_num_prims", "rt_parse_scene_file", "rt_host_scene_desc",
           "rt_host_scene_filename", "rt_host_scene_bounds", "rt_host_scene_free", "rt_triangle_setup", "rt_png_write",
           "rt_png_read", "rt_free", "rt_xorwow_init", "rt_last_error", "rt_version",
           "rt_comm_init", "rt_comm_unique_id", "rt_comm_init_rank", "rt_comm_destroy", "rt_comm_size", "rt_comm_num_local", "rt_comm_local_rank",
           "rt_comm_local_device", "rt_tile_shape", "rt_scene_broadcast", "rt_comm_update_triangles", "rt_comm_build_lbvh", "rt_framebuffer_gather",
           "rt_render_multi", "rt_render_reserve", "rt_device_alloc", "rt_device_free", "rt_trace_static", "rt_trace_static_max_rays"]


def load_library():
    """dlopen libb200rt.so; raises RtError when it has not been built (there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtError("libb200rt.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(make -C cuda_ray_tracer_b200); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, ip = C.c_void_p, C.c_int
    L.rt_last_error.restype = C.c_char_p
    L.rt_version.restype = C.c_char_p
    L.rt_scene_create.argtypes = [C.POINTER(SceneDesc), ip, C.POINTER(vp)]
    L.rt_scene_destroy.argtypes = [vp]
    L.rt_scene_update_triangles.argtypes = [vp, vp, vp, vp, vp]
    L.rt_build_lbvh.argtypes = [vp, ip, C.c_uint, vp, C.POINTER(C.c_float)]
    L.rt_scene_begin_frame.argtypes = [vp, vp]
    L.rt_scene_build_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.rt_render.argtypes = [vp, ip, ip, ip, C.c_uint, C.POINTER(TileSet), vp, vp, vp, C.POINTER(Counters), vp]
    L.rt_render_host.argtypes = [vp, ip, ip, ip, C.c_uint, C.POINTER(TileSet), vp, vp, vp, C.POINTER(Counters)]
    L.rt_trace.argtypes = [vp, ip, vp, vp, C.c_uint, vp, vp, C.POINTER(Counters), vp]
    L.rt_export_reference_layout.argtypes = [vp, vp, vp, vp, vp, vp]
    L.rt_scene_num_prims.argtypes = [vp]
    L.rt_parse_scene_file.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.rt_host_scene_desc.argtypes = [vp, C.POINTER(SceneDesc)]
    L.rt_host_scene_filename.argtypes = [vp]; L.rt_host_scene_filename.restype = C.c_char_p
    L.rt_host_scene_bounds.argtypes = [vp, vp, vp]
    L.rt_host_scene_free.argtypes = [vp]
    L.rt_triangle_setup.argtypes = [ip, vp, vp, vp, vp, vp, vp]
    L.rt_png_write.argtypes = [C.c_char_p, ip, ip, vp]
    L.rt_png_read.argtypes = [C.c_char_p, C.POINTER(ip), C.POINTER(ip), C.POINTER(vp)]
    L.rt_free.argtypes = [vp]
    L.rt_xorwow_init.argtypes = [C.c_uint64, C.c_uint64, vp]
    pvp = C.POINTER(vp)
    L.rt_comm_init.argtypes = [ip, C.POINTER(ip), pvp]
    L.rt_comm_unique_id.argtypes = [vp]
    L.rt_comm_init_rank.argtypes = [ip, ip, vp, ip, pvp]
    L.rt_comm_destroy.argtypes = [vp]
    for n in ("rt_comm_size", "rt_comm_num_local"):
        getattr(L, n).argtypes = [vp]
    L.rt_comm_local_rank.argtypes = [vp, ip]; L.rt_comm_local_device.argtypes = [vp, ip]
    L.rt_tile_shape.argtypes = [ip, ip, C.POINTER(ip), C.POINTER(ip)]; L.rt_tile_shape.restype = None
    L.rt_scene_broadcast.argtypes = [vp, C.POINTER(SceneDesc), ip, pvp]
    L.rt_comm_update_triangles.argtypes = [vp, pvp, vp, vp, vp, ip, pvp]
    L.rt_comm_build_lbvh.argtypes = [vp, pvp, ip, C.c_uint, C.POINTER(C.c_float)]
    L.rt_framebuffer_gather.argtypes = [vp, pvp, ip, ip, ip, ip, ip, pvp]
    L.rt_render_reserve.argtypes = [vp, ip, ip]
    L.rt_trace_static.argtypes = [vp, ip, vp, vp, C.c_uint, vp, vp, vp]
    L.rt_trace_static_max_rays.argtypes = []
    L.rt_device_alloc.argtypes = [ip, C.c_size_t]; L.rt_device_alloc.restype = vp
    L.rt_device_free.argtypes = [ip, vp]; L.rt_device_free.restype = None
    L.rt_render_multi.argtypes = [vp, pvp, ip, ip, ip, C.c_uint, ip, ip, ip, pvp, vp, C.POINTER(Counters)]
    _lib = L
    return L


def _check(rc, what):
    if rc != 0:
        raise RtError("%s failed (%d): %s" % (what, rc, load_library().rt_last_error().decode()))


def _ptr(a):
    """host numpy array, torch tensor (host or cuda) or None -> void*"""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return a.ctypes.data_as(C.c_void_p)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, np.float32)
    return a if shape is None else a.reshape(shape)


def make_desc(cfg, arrays, scene_min=None, scene_max=None):
    """Build a SceneDesc from config scalars (dict with the StlConfig field names, include/config.hpp:26-58)
    and reference-layout arrays (dict: sphere_c/r/mat, tri=[p0,p1,p2,nor,e1,e2], tri_mat, refs, suns, bulbs,
    planes).  Arrays may be numpy (host) or torch tensors (host/device).  Returns (desc, keepalive)."""
    d = SceneDesc()
    d.width, d.height, d.bounces, d.aa, d.gi = int(cfg["width"]), int(cfg["height"]), int(cfg["bounces"]), int(cfg["aa"]), int(cfg["gi"])
    d.fisheye, d.panorama = int(cfg["fisheye"]), int(cfg["panorama"])
    d.dof_focus, d.dof_lens, d.expose, d.ior, d.rough = [float(cfg[k]) for k in ("dof_focus", "dof_lens", "expose", "ior", "rough")]
    for k in ("color", "forward", "right", "up", "eye", "target_up", "trans", "shine"):
        setattr(d, k, (C.c_float * 3)(*[float(x) for x in cfg[k]]))
    inf = float("inf")
    d.scene_min = (C.c_float * 3)(*([inf] * 3 if scene_min is None else [float(x) for x in scene_min]))
    d.scene_max = (C.c_float * 3)(*([-inf] * 3 if scene_max is None else [float(x) for x in scene_max]))
    keep = []

    def put(name, a, dtype=np.float32):
        if a is None:
            return 0
        if not hasattr(a, "data_ptr"):
            a = np.ascontiguousarray(a, dtype)
        keep.append(a)
        setattr(d, name, _ptr(a))
        return len(a)

    d.num_spheres = put("sphere_c", arrays.get("sphere_c")); put("sphere_r", arrays.get("sphere_r")); put("sphere_mat", arrays.get("sphere_mat"))
    tri = arrays.get("tri")
    if tri is not None and len(tri[0]):
        for name, a in zip(("tri_p0", "tri_p1", "tri_p2", "tri_nor", "tri_e1", "tri_e2"), tri):
            d.num_triangles = put(name, a)
        put("tri_mat", arrays.get("tri_mat"))
    d.num_prims = put("refs", arrays.get("refs"), np.int32)
    d.num_suns = put("suns", arrays.get("suns")); d.num_bulbs = put("bulbs", arrays.get("bulbs")); d.num_planes = put("planes", arrays.get("planes"))
    return d, keep


class HostScene:
    """parseInput(argv, StlConfig&) — src/parse.cpp:16-222 — behind rt_parse_scene_file."""

    def __init__(self, path):
        L = load_library()
        self.h = C.c_void_p()
        _check(L.rt_parse_scene_file(path.encode(), C.byref(self.h)), "rt_parse_scene_file")
        self.desc = SceneDesc()
        _check(L.rt_host_scene_desc(self.h, C.byref(self.desc)), "rt_host_scene_desc")
        self.filename = L.rt_host_scene_filename(self.h).decode()

    def bounds(self):
        mn = np.zeros(3, np.float32); mx = np.zeros(3, np.float32)
        load_library().rt_host_scene_bounds(self.h, _ptr(mn), _ptr(mx))
        return mn, mx

    def array(self, name, shape, dtype=np.float32):
        p = getattr(self.desc, name)
        n = int(np.prod(shape))
        if n == 0 or not p:
            return np.zeros(shape, dtype)
        ct = C.c_float if dtype == np.float32 else C.c_int32
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(ct)), shape=(n,)).view(dtype).reshape(shape).copy()

    def __del__(self):
        try:
            load_library().rt_host_scene_free(self.h)
        except Exception:
            pass


def parse_scene_file(path):
    return HostScene(path)


class Scene:
    """Device-resident scene (RawConfig + its device arrays, include/config.hpp:75-126)."""

    def __init__(self, desc, device=0, keepalive=None, handle=None):
        L = load_library()
        self.h = C.c_void_p()
        self.desc = desc
        if handle is not None:   # an rt_scene created by rt_scene_broadcast
            self.h = C.c_void_p(handle)
        else:
            _check(L.rt_scene_create(C.byref(desc), device, C.byref(self.h)), "rt_scene_create")
        self.n = L.rt_scene_num_prims(self.h)
        self.device = device

    def close(self):
        if self.h:
            load_library().rt_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # build_lbvh_karas(RawConfig&, int morton_bits = 30), include/lbvh_builder.cuh:14
    def build_lbvh_karas(self, morton_bits=30, flags=BOUNDS_FROM_DESC, stream=None, timed=True):
        ms = C.c_float(0)
        _check(load_library().rt_build_lbvh(self.h, morton_bits, flags, stream, C.byref(ms) if timed else None), "rt_build_lbvh")
        return ms.value

    def build_ms(self):
        """device time of the last build, read after the fact (keeps a frame loop asynchronous)"""
        ms = C.c_float(0)
        _check(load_library().rt_scene_build_ms(self.h, C.byref(ms)), "rt_scene_build_ms")
        return ms.value

    def begin_frame(self, stream=None):
        """fork point of RENDER_OVERLAP_BUILD, when the frame starts before update_triangles (e.g. with an upload)"""
        _check(load_library().rt_scene_begin_frame(self.h, stream), "rt_scene_begin_frame")

    def update_triangles(self, p0, p1, p2, stream=None):
        keep = [a if hasattr(a, "data_ptr") else _f32(a) for a in (p0, p1, p2)]
        _check(load_library().rt_scene_update_triangles(self.h, _ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), stream), "rt_scene_update_triangles")

    def export_reference_layout(self, morton_bits=30):
        n = self.n
        codes = np.zeros(n, np.uint32 if morton_bits == 30 else np.uint64)
        refs = np.zeros((n, 2), np.int32); nodes = np.zeros((2 * n - 1, 11), np.uint32); perm = np.zeros(n, np.uint32)
        _check(load_library().rt_export_reference_layout(self.h, _ptr(codes) if morton_bits == 30 else None,
                                                         _ptr(codes) if morton_bits == 63 else None, _ptr(refs), _ptr(nodes), _ptr(perm)),
               "rt_export_reference_layout")
        return dict(codes=codes, refs=refs, perm=perm, bbox=nodes[:, 0:6].copy().view(np.float32), left=nodes[:, 6].copy(),
                    right=nodes[:, 7].copy(), prim=nodes[:, 8].copy(), nleaf=(nodes[:, 9] & 0xFFFF).astype(np.uint16),
                    counter=nodes[:, 10].copy())

    # render(pixel_t* d_image, int w, int h, int aa, RawConfig*), include/draw.cuh:10 — host buffers
    def render(self, w=None, h=None, aa=None, flags=SPP_REF, tiles=None, want_ids=False, want_written=False):
        w = self.desc.width if w is None else w; h = self.desc.height if h is None else h; aa = self.desc.aa if aa is None else aa
        rgba = np.zeros((h, w, 4), np.uint8)
        ids = np.zeros((h, w), np.int32) if want_ids else None
        wr = np.zeros((h, w), np.uint8) if want_written else None
        cnt = Counters()
        _check(load_library().rt_render_host(self.h, w, h, aa, flags, C.byref(tiles) if tiles else None, _ptr(rgba), _ptr(ids), _ptr(wr),
                                             C.byref(cnt)), "rt_render_host")
        return dict(rgba=rgba, ids=ids, written=wr, counters=cnt)

    def render_host_into(self, rgba, w, h, aa, flags=SPP_REF, tiles=None, counters=None):
        """rt_render_host into a caller-owned host buffer (numpy array or pinned torch tensor, [h][w][4] u8)"""
        _check(load_library().rt_render_host(self.h, w, h, aa, flags, C.byref(tiles) if tiles else None, _ptr(rgba), None, None,
                                             C.byref(counters) if counters is not None else None), "rt_render_host")

    # same with caller-owned device buffers (torch tensors / raw pointers), asynchronous on `stream`
    def render_device(self, w, h, aa, d_rgba, flags=SPP_REF, tiles=None, d_ids=None, d_written=None, counters=None, stream=None):
        _check(load_library().rt_render(self.h, w, h, aa, flags, C.byref(tiles) if tiles else None, _ptr(d_rgba), _ptr(d_ids), _ptr(d_written),
                                        C.byref(counters) if counters is not None else None, stream), "rt_render")

    def trace_device(self, n, d_org, d_dir, d_ids, d_t=None, flags=0, counters=None, stream=None):
        _check(load_library().rt_trace(self.h, n, _ptr(d_org), _ptr(d_dir), flags, _ptr(d_ids), _ptr(d_t),
                                       C.byref(counters) if counters is not None else None, stream), "rt_trace")

    def trace_static(self, n, d_org, d_dir, d_ids, d_t=None, flags=0, stream=None):
        """rt_trace_static: rt_trace with its scratch in a fixed device array (no allocation; n <= rt_trace_static_max_rays())"""
        _check(load_library().rt_trace_static(self.h, n, _ptr(d_org), _ptr(d_dir), flags, _ptr(d_ids), _ptr(d_t), stream), "rt_trace_static")


def triangle_setup(p0, p1, p2):
    p0, p1, p2 = _f32(p0), _f32(p1), _f32(p2)
    nor, e1, e2 = np.zeros_like(p0), np.zeros_like(p0), np.zeros_like(p0)
    load_library().rt_triangle_setup(len(p0), _ptr(p0), _ptr(p1), _ptr(p2), _ptr(nor), _ptr(e1), _ptr(e2))
    return nor, e1, e2


def png_write(path, rgba):
    rgba = np.ascontiguousarray(rgba, np.uint8)
    _check(load_library().rt_png_write(path.encode(), rgba.shape[1], rgba.shape[0], _ptr(rgba)), "rt_png_write")


def png_read(path):
    L = load_library()
    w, h, p = C.c_int(), C.c_int(), C.c_void_p()
    _check(L.rt_png_read(path.encode(), C.byref(w), C.byref(h), C.byref(p)), "rt_png_read")
    a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h.value * w.value * 4,)).reshape(h.value, w.value, 4).copy()
    L.rt_free(p)
    return a


def xorwow_init(seed, subsequence):
    st = np.zeros(6, np.uint32)
    load_library().rt_xorwow_init(seed, subsequence, _ptr(st))
    return st


class Comm:
    """rt_comm: one NCCL communicator over the GPUs that render one image (include/rt_api.h, csrc/comm.cu).
    Comm(ndev=N): this process drives N GPUs.  Comm.from_torch_distributed(device): one process per GPU (torchrun),
    the 128-byte NCCL id travels through torch.distributed."""

    def __init__(self, ndev=None, devices=None, handle=None):
        L = load_library()
        self.h = C.c_void_p()
        if handle is not None:
            self.h = handle
        else:
            dv = (C.c_int * ndev)(*devices) if devices is not None else None
            _check(L.rt_comm_init(ndev, dv, C.byref(self.h)), "rt_comm_init")
        self.size = L.rt_comm_size(self.h); self.num_local = L.rt_comm_num_local(self.h)
        self.ranks = [L.rt_comm_local_rank(self.h, k) for k in range(self.num_local)]
        self.devices = [L.rt_comm_local_device(self.h, k) for k in range(self.num_local)]

    @classmethod
    def from_torch_distributed(cls, device):
        import torch
        import torch.distributed as dist
        L = load_library()
        rank, world = dist.get_rank(), dist.get_world_size()
        idb = np.zeros(128, np.uint8)
        if rank == 0:
            _check(L.rt_comm_unique_id(_ptr(idb)), "rt_comm_unique_id")
        t = torch.from_numpy(idb)
        if dist.get_backend() == "nccl":
            t = t.cuda(device); dist.broadcast(t, 0); idb = t.cpu().numpy()
        else:
            dist.broadcast(t, 0)
        h = C.c_void_p()
        _check(L.rt_comm_init_rank(world, rank, _ptr(np.ascontiguousarray(idb)), device, C.byref(h)), "rt_comm_init_rank")
        return cls(handle=h)

    def close(self):
        if self.h:
            load_library().rt_comm_destroy(self.h); self.h = C.c_void_p()

    def _arr(self, items):
        return (C.c_void_p * self.num_local)(*[(x.h if isinstance(x, Scene) else _ptr(x)) for x in items])

    def broadcast_scene(self, desc, root=0):
        """rt_scene_broadcast: desc on the root's process (None elsewhere) -> one Scene per local GPU"""
        out = (C.c_void_p * self.num_local)()
        _check(load_library().rt_scene_broadcast(self.h, C.byref(desc) if desc is not None else None, root, out), "rt_scene_broadcast")
        return [Scene(desc, self.devices[k], handle=out[k]) for k in range(self.num_local)]

    def update_triangles(self, scenes, p0, p1, p2, root=0):
        _check(load_library().rt_comm_update_triangles(self.h, self._arr(scenes), _ptr(p0), _ptr(p1), _ptr(p2), root, None), "rt_comm_update_triangles")

    def build_lbvh(self, scenes, bits=30, flags=BOUNDS_COMPUTED):
        ms = C.c_float()
        _check(load_library().rt_comm_build_lbvh(self.h, self._arr(scenes), bits, flags, C.byref(ms)), "rt_comm_build_lbvh")
        return ms.value

    def gather_framebuffer(self, images, w, h, tile_w, tile_h, root=0):
        """images: one device image (torch tensor / pointer) per local GPU"""
        _check(load_library().rt_framebuffer_gather(self.h, self._arr(images), w, h, tile_w, tile_h, root, None), "rt_framebuffer_gather")

    def render(self, scenes, w, h, aa, images, flags=SPP_REF, tile_w=0, tile_h=0, root=0, host_out=None):
        cnt = Counters()
        _check(load_library().rt_render_multi(self.h, self._arr(scenes), w, h, aa, flags, tile_w, tile_h, root, self._arr(images), _ptr(host_out),
                                              C.byref(cnt)), "rt_render_multi")
        return cnt


def tile_shape(ranks, width):
    tw, th = C.c_int(), C.c_int()
    load_library().rt_tile_shape(ranks, width if width else 0, C.byref(tw), C.byref(th))
    return tw.value, th.value
