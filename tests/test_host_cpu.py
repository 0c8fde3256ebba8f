"""Host side of the drop-in surface, no GPU needed: the C-ABI library loads and exports every symbol
include/rt_api.h declares, the scene parser and the host precompute agree with the oracle bit for bit, the PNG
writer round-trips, compute entry points fail loudly without a B200, and the multi-GPU image partition +
framebuffer gather works across two gloo ranks."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import feature_scenes as FS
import oracle_lib as O
from cuda_ray_tracer_b200 import api, parallel, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
have_scenes = os.path.exists(os.path.join(O.SCENE_DIR, "tri.txt"))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "rt_api.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    L = api.load_library()
    for n in sorted(names):
        assert hasattr(L, n), "libb200rt.so does not export " + n
    assert set(api.EXPORTS) <= names
    assert b"sm_100a" in L.rt_version()


def test_abi_struct_sizes_match_header():
    # the ctypes mirrors must have the C layout: compile a probe against the header
    src = '#include "rt_api.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu\\n", sizeof(rt_scene_desc), sizeof(rt_tile_set), sizeof(rt_counters));}'
    exe = "/tmp/rt_abi_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src.encode(), check=True)
    sizes = [int(x) for x in subprocess.run([exe], capture_output=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(api.SceneDesc), C.sizeof(api.TileSet), C.sizeof(api.Counters)]


@pytest.mark.skipif(not have_scenes, reason="scene fixtures not present")
@pytest.mark.parametrize("name", ["tri", "redchair", "tenthousand", "spiral"])
def test_parser_matches_oracle(name):
    _check_parser_against_oracle(os.path.join(O.SCENE_DIR, name + ".txt"))


@pytest.mark.parametrize("kind", FS.KINDS)
def test_parser_matches_oracle_on_feature_scenes(kind, tmp_path):
    """bulb / fisheye / panorama / dof / gi / expose / up / per-primitive shininess, transparency, ior, roughness lines
    (tests/feature_scenes.py; the oracle's parse of the same files is pinned to parse.cpp in test_oracle_cpu.py)"""
    _check_parser_against_oracle(FS.write(kind, str(tmp_path / (kind + ".txt"))))


def _check_parser_against_oracle(path):
    hs = api.parse_scene_file(path); d = hs.desc
    o = O.OracleScene(path); c = o.config(); a = o.arrays(); n = o.counts()
    assert (d.width, d.height, d.bounces, d.aa, d.gi, d.fisheye, d.panorama) == (c["width"], c["height"], c["bounces"], c["aa"], c["gi"], int(c["fisheye"]), int(c["panorama"]))
    assert hs.filename == c["filename"]
    for k in ("color", "forward", "right", "up", "eye", "target_up", "trans", "shine"):
        assert np.array_equal(np.array(getattr(d, k)[:], np.float32).view(np.uint32), np.asarray(c[k], np.float32).view(np.uint32)), k
    for k in ("dof_focus", "dof_lens", "expose", "ior", "rough"):
        assert np.float32(getattr(d, k)) == np.float32(c[k]) or (np.isinf(getattr(d, k)) and np.isinf(c[k]))
    assert (d.num_spheres, d.num_triangles, d.num_prims, d.num_suns, d.num_bulbs, d.num_planes) == (
        n["spheres"], n["triangles"], n["prims"], n["suns"], n["bulbs"], n["planes"])
    eq = lambda x, y: np.array_equal(x.view(np.uint32), y.view(np.uint32))
    assert eq(hs.array("sphere_c", (n["spheres"], 3)), a["sphere_c"]) and eq(hs.array("sphere_r", (n["spheres"],)), a["sphere_r"])
    assert eq(hs.array("sphere_mat", (n["spheres"], 11)), a["sphere_mat"]) and eq(hs.array("tri_mat", (n["triangles"], 11)), a["tri_mat"])
    for k, f in enumerate(("tri_p0", "tri_p1", "tri_p2", "tri_nor", "tri_e1", "tri_e2")):
        assert eq(hs.array(f, (n["triangles"], 3)), a["tri"][k]), f
    assert np.array_equal(hs.array("refs", (n["prims"], 2), np.int32), a["refs"])
    assert eq(hs.array("suns", (n["suns"], 6)), a["suns"]) and eq(hs.array("planes", (n["planes"], 21)), a["planes"])
    mn, mx = hs.bounds(); omn, omx = o.bounds(O.BOUNDS_COMPUTED)
    assert eq(mn, omn) and eq(mx, omx)
    # the descriptor carries the reference's never-assigned bounds
    assert all(np.isinf(v) and v > 0 for v in d.scene_min[:]) and all(np.isinf(v) and v < 0 for v in d.scene_max[:])


def test_parser_errors(tmp_path):
    with pytest.raises(api.RtError, match="Error opening file"):
        api.parse_scene_file(str(tmp_path / "nope.txt"))
    for bad in ("sphere 1 2 3\n", "png 10 10\n", "frobnicate 1\n", "tri 1 2 9\n", "eye 1 2 3 4\n", "aa x\n"):
        p = tmp_path / "bad.txt"; p.write_text("png 4 4 out.png\nxyz 0 0 0\n" + bad)
        with pytest.raises(api.RtError, match="One of the lines are not valid"):
            api.parse_scene_file(str(p))
    p = tmp_path / "ok.txt"; p.write_text("png 4 4 out.png\r\n\n  \t\nshininess 0.5\ntransparency 0.1 0.2 0.3\nsphere 0 0 -1 0.5\n")
    hs = api.parse_scene_file(str(p))
    m = hs.array("sphere_mat", (1, 11))
    assert np.allclose(m[0, 3:9], [0.5, 0.5, 0.5, 0.1, 0.2, 0.3])


def test_synthetic_scene_text_roundtrip(tmp_path):
    cfg, a = synth.soup(300, seed=3, width=64, height=32, aa=4)
    a = synth.finish_triangles(a)
    p = str(tmp_path / "soup.txt")
    synth.write_scene_text(p, cfg, a, "soup.png")
    hs = api.parse_scene_file(p)
    for k, f in enumerate(("tri_p0", "tri_p1", "tri_p2", "tri_nor", "tri_e1", "tri_e2")):
        assert np.array_equal(hs.array(f, (300, 3)).view(np.uint32), a["tri"][k].view(np.uint32)), f
    assert np.array_equal(hs.array("tri_mat", (300, 11)).view(np.uint32), a["tri_mat"].view(np.uint32))
    o = O.OracleScene(p)
    assert np.array_equal(o.arrays()["tri"][4].view(np.uint32), a["tri"][4].view(np.uint32))


def test_triangle_setup_matches_oracle():
    rng = np.random.RandomState(0)
    p0, p1, p2 = [rng.uniform(-3, 3, (5000, 3)).astype(np.float32) for _ in range(3)]
    p2[:50] = p1[:50]                      # degenerate
    p2[50:100] = p0[50:100] + (p1[50:100] - p0[50:100]) * np.float32(2)  # collinear
    mine = api.triangle_setup(p0, p1, p2); ref = O.triangle_setup(p0, p1, p2)
    for x, y in zip(mine, ref):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


def test_xorwow_init_matches_oracle():
    for seed, sub in [(1234, 0), (1234, 1), (1234, 255), (1234, 256), (1234, 123457), (1234 + 99, 31), (2**33 + 7, 5)]:
        assert np.array_equal(api.xorwow_init(seed, sub), O.xorwow_init(seed, sub))


def test_png_roundtrip(tmp_path):
    rng = np.random.RandomState(1)
    for shape in ((1, 1), (7, 13), (100, 100), (33, 257)):
        img = rng.randint(0, 256, shape + (4,)).astype(np.uint8)
        p = str(tmp_path / "x.png")
        api.png_write(p, img)
        assert open(p, "rb").read(8) == b"\x89PNG\r\n\x1a\n"
        assert np.array_equal(api.png_read(p), img)
    try:
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(p).convert("RGBA")), img)  # an independent decoder agrees
    except ImportError:
        pass


def test_compute_fails_loudly_without_a_b200():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg, a = synth.soup(10, seed=1, width=8, height=8, aa=0)
    a = synth.finish_triangles(a)
    desc, keep = api.make_desc(cfg, a)
    with pytest.raises(api.RtError, match="rt_scene_create failed"):
        api.Scene(desc, 0)


def test_missing_library_is_an_error(monkeypatch):
    monkeypatch.setattr(api, "_lib", None)
    monkeypatch.setattr(api, "LIB_PATH", "/nonexistent/libb200rt.so")
    with pytest.raises(api.RtError, match="no CPU fallback"):
        api.load_library()


def test_tile_owner_partition():
    for (w, h, world) in ((100, 70, 2), (3840, 2160, 8), (33, 9, 4), (5, 5, 3)):
        own = parallel.tile_owner(w, h, world)
        assert own.shape == (h, w) and own.min() == 0 and own.max() == min(world, ((w + 31) // 32) * ((h + 7) // 8)) - 1
        counts = np.bincount(own.ravel(), minlength=world)
        if w * h > 100000:
            assert counts.max() / counts.min() < 1.02  # interleaved tiles balance the ranks


def test_tile_shape_avoids_column_ownership():
    import math
    for w in (3840, 1920, 800, 100, 4096):
        for g in (2, 4, 8):
            tw, th = parallel.tile_shape(g, w)
            assert th == 8 and 32 <= tw <= 1280
            own = parallel.tile_owner(w, 64, g, tw, th)
            # no rank owns whole columns of tiles: every tile-wide band of the image sees more than one rank
            if (w + tw - 1) // tw >= g:
                assert len(np.unique(own[:, : min(w, tw)])) > 1 or math.gcd((w + tw - 1) // tw, g) > 1
    assert parallel.tile_shape(8, 3840) == (256, 8) and parallel.tile_shape(1, 3840) == (32, 8)
    counts = np.bincount(parallel.tile_owner(3840, 2160, 8, 256, 8).ravel(), minlength=8)
    assert counts.max() / counts.min() < 1.02


_WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from cuda_ray_tracer_b200 import parallel
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
w, h = 100, 70
full = torch.from_numpy(np.random.RandomState(5).randint(0, 256, (h, w, 4)).astype(np.uint8))
scene = torch.zeros(1000) if rank else torch.arange(1000, dtype=torch.float32)
parallel.broadcast_scene([scene], 0)                       # scene replication
assert torch.equal(scene, torch.arange(1000, dtype=torch.float32))
own = torch.from_numpy(parallel.tile_owner(w, h, world))
ts = parallel.tile_set(world, rank); assert (ts.num_ranks, ts.rank) == (world, rank)
img = torch.where((own == rank)[..., None], full, torch.zeros_like(full))   # this rank's tiles, zero elsewhere
parallel.gather_framebuffer(img, 0)                        # framebuffer gather
if rank == 0:
    assert torch.equal(img, full), "gathered framebuffer differs"
# sharded upload: every rank contributes 1/world of a replicated host array, all-gather rebuilds it everywhere
host = torch.arange(1001 * 3, dtype=torch.float32).reshape(1001, 3)
rows = (1001 + world - 1) // world
fullbuf = torch.zeros((rows * world, 3)); part = torch.zeros((rows, 3))
parallel.upload_sharded(host, fullbuf, part, rank, world)
assert torch.equal(fullbuf[:1001], host), "sharded upload differs"
dist.barrier(); dist.destroy_process_group()
sys.stdout.write("ok%d\n" % rank); sys.stdout.flush()
'''


def test_two_rank_partition_and_gather_gloo(tmp_path):
    script = tmp_path / "worker.py"; script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29571", str(script), ROOT], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ok0" in r.stdout and "ok1" in r.stdout


def _run_cli(args, cwd):
    import subprocess
    exe = os.path.join(os.path.dirname(api.__file__), "raytracer")
    assert os.path.exists(exe), "cuda_ray_tracer_b200/raytracer is not built"
    return subprocess.run([exe] + args, capture_output=True, text=True, timeout=60, cwd=str(cwd))


def test_cli_error_paths(tmp_path):
    """parse.cpp:22-25 / 218-221: a missing file and an invalid line end the program with exit code 1 and the reference's text."""
    r = _run_cli([str(tmp_path / "nope.txt")], tmp_path)
    assert r.returncode == 1 and "Error opening file" in r.stdout
    bad = tmp_path / "bad.txt"
    bad.write_text("png 8 8 x.png\nsphere 0 0\n")
    r = _run_cli([str(bad)], tmp_path)
    assert r.returncode == 1 and "One of the lines are not valid" in r.stdout
    r = _run_cli([], tmp_path)
    assert r.returncode == 1 and "usage" in r.stderr


# ---- chunk-parallel parser --------------------------------------------------------------
HOST_ARRAYS = [("sphere_c", "num_spheres", 3), ("sphere_r", "num_spheres", 1), ("sphere_mat", "num_spheres", 11), ("tri_p0", "num_triangles", 3),
               ("tri_p1", "num_triangles", 3), ("tri_p2", "num_triangles", 3), ("tri_nor", "num_triangles", 3), ("tri_e1", "num_triangles", 3),
               ("tri_e2", "num_triangles", 3), ("tri_mat", "num_triangles", 11), ("suns", "num_suns", 6), ("bulbs", "num_bulbs", 6),
               ("planes", "num_planes", 21)]


def _parse_with(path, threads, chunk_bytes, monkeypatch):
    monkeypatch.setenv("RT_PARSE_THREADS", str(threads))
    monkeypatch.setenv("RT_PARSE_CHUNK_BYTES", str(chunk_bytes))
    return api.parse_scene_file(path)


def _same_host_scene(a, b):
    da, db = a.desc, b.desc
    assert bytes(da)[:C.sizeof(api.SceneDesc)] != b"" and a.filename == b.filename
    for f in ("width", "height", "bounces", "aa", "gi", "fisheye", "panorama", "num_spheres", "num_triangles", "num_prims", "num_suns",
              "num_bulbs", "num_planes"):
        assert getattr(da, f) == getattr(db, f), f
    for f in ("color", "shine", "trans", "forward", "right", "up", "eye", "target_up"):
        assert np.array_equal(np.array(getattr(da, f)[:], np.float32).view(np.uint32), np.array(getattr(db, f)[:], np.float32).view(np.uint32)), f
    for f in ("dof_focus", "dof_lens", "expose", "ior", "rough"):
        assert np.array_equal(np.float32(getattr(da, f)).view(np.uint32), np.float32(getattr(db, f)).view(np.uint32)), f
    for name, cnt, w in HOST_ARRAYS:
        n = getattr(da, cnt)
        assert np.array_equal(a.array(name, (n, w)).view(np.uint32), b.array(name, (n, w)).view(np.uint32)), name
    assert np.array_equal(a.array("refs", (da.num_prims, 2), np.int32), b.array("refs", (da.num_prims, 2), np.int32))
    for x, y in zip(a.bounds(), b.bounds()):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


def _stateful_scene_text(n, seed):
    """Every kind of line, material changes between primitives, absolute and relative vertex indices that reach far
    back (across chunk borders), camera lines in the middle of the file."""
    rng = np.random.RandomState(seed)
    out = ["png 64 48 stateful.png", "eye 0 0 3", "bounces 3"]
    nv = 0
    for i in range(n):
        r = rng.randint(0, 100)
        if r < 45 or nv < 3:
            out.append("xyz %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 3).astype(np.float32))); nv += 1
        elif r < 70:
            if rng.randint(2):
                idx = rng.randint(1, nv + 1, 3)          # absolute, 1-based, anywhere in the file so far
            else:
                idx = -rng.randint(1, min(nv, 400) + 1, 3)  # relative to the current end of the list
            out.append("tri %d %d %d" % tuple(idx))
        elif r < 78:
            out.append("sphere %.9g %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 4).astype(np.float32)))
        elif r < 84:
            out.append("color %.9g %.9g %.9g" % tuple(rng.uniform(0, 1, 3).astype(np.float32)))
        elif r < 87:
            out.append("shininess %.9g" % rng.uniform(0, 1) if rng.randint(2) else "shininess %.9g %.9g %.9g" % tuple(rng.uniform(0, 1, 3)))
        elif r < 90:
            out.append("transparency %.9g" % rng.uniform(0, 1) if rng.randint(2) else "transparency %.9g %.9g %.9g" % tuple(rng.uniform(0, 1, 3)))
        elif r < 92:
            out.append("ior %.9g" % rng.uniform(1, 2))
        elif r < 94:
            out.append("roughness %.9g" % rng.uniform(0, 0.3))
        elif r < 95:
            out.append("sun %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 3)))
        elif r < 96:
            out.append("bulb %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 3)))
        elif r < 97:
            out.append("plane %.9g %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 4)))
        elif r < 98:
            out.append(["forward %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 3)), "up %.9g %.9g %.9g" % tuple(rng.uniform(-1, 1, 3)),
                        "expose %.9g" % rng.uniform(1, 3), "aa %d" % rng.randint(1, 40), "dof 2.5 0.01", "gi 2"][rng.randint(6)])
        else:
            out.append(["", "   \t ", " "][rng.randint(3)])
    return "\n".join(out) + "\n"


@pytest.mark.parametrize("seed", [1, 2])
def test_parser_chunked_equals_serial(seed, tmp_path, monkeypatch):
    p = tmp_path / "stateful.txt"
    p.write_text(_stateful_scene_text(6000, seed))
    serial = _parse_with(str(p), 1, 1 << 40, monkeypatch)
    assert serial.desc.num_triangles > 500 and serial.desc.num_spheres > 100 and serial.desc.num_planes > 10
    for threads, chunk in ((4, 4096), (3, 700), (8, 64)):
        _same_host_scene(serial, _parse_with(str(p), threads, chunk, monkeypatch))
    # and against the reference's own parseInput (oracle/_ref/libref_host.so = src/parse.cpp compiled for the host)
    import ref_lib
    if ref_lib.available():
        r = ref_lib.RefScene(str(p)); ra = r.arrays(); d = serial.desc
        assert (d.num_spheres, d.num_triangles, d.num_prims, d.num_suns, d.num_bulbs, d.num_planes) == tuple(
            r.n[k] for k in ("spheres", "triangles", "prims", "suns", "bulbs", "planes"))
        eq = lambda x, y: np.array_equal(np.ascontiguousarray(x).view(np.uint32), np.ascontiguousarray(y).view(np.uint32))
        for name, cnt, w in HOST_ARRAYS:
            if name.startswith("tri_") and name != "tri_mat":
                continue
            assert eq(serial.array(name, (getattr(d, cnt), w)).reshape(ra[name].shape), ra[name]), name
        for k, f in enumerate(("tri_p0", "tri_p1", "tri_p2", "tri_nor", "tri_e1", "tri_e2")):
            assert eq(serial.array(f, (d.num_triangles, 3)), ra["tri"][k]), f
        assert np.array_equal(serial.array("refs", (d.num_prims, 2), np.int32), ra["refs"])


@pytest.mark.skipif(not have_scenes, reason="scene fixtures not present")
@pytest.mark.parametrize("name", ["redchair", "tenthousand", "spiral"])
def test_parser_chunked_equals_serial_on_reference_scenes(name, tmp_path, monkeypatch):
    path = os.path.join(O.SCENE_DIR, name + ".txt")
    _same_host_scene(_parse_with(path, 1, 1 << 40, monkeypatch), _parse_with(path, 4, 2048, monkeypatch))


def test_parser_chunked_reports_the_first_bad_line(tmp_path, monkeypatch):
    lines = _stateful_scene_text(3000, 7).split("\n")[:-1]
    n = len(lines)
    cases = {"unknown keyword": (n // 3, "frobnicate 1 2"), "bad float in a primitive": (n // 2, "sphere 0 0 zero 1"),
             "bad vertex": (n // 5, "xyz 1 2 three"), "index out of range": (n // 4, "tri 1 2 999999"), "bad camera line": (n // 6, "aa many"),
             "bad material": (n // 7, "color 1 x 1")}
    for what, (at, text) in cases.items():
        bad = list(lines); bad[at] = text
        bad[at + 40] = "another bad line"  # a later error must not win
        p = tmp_path / "bad.txt"; p.write_text("\n".join(bad) + "\n")
        for threads, chunk in ((1, 1 << 40), (4, 1024), (8, 64)):
            with pytest.raises(api.RtError, match=r"One of the lines are not valid. \(.*:%d\)" % (at + 1)):
                _parse_with(str(p), threads, chunk, monkeypatch)


def test_parser_number_syntax_follows_stof(tmp_path):
    """std::stof / std::stoi (parse.cpp:47-215) take strtof/strtol syntax: leading '+', hex floats, inf/nan, trailing junk after a valid prefix."""
    p = tmp_path / "n.txt"
    p.write_text("png +8 0x8 n.png\neye +1.5 0x1p-1 -.5e1\nexpose 2.5abc\nior inf\nroughness 1e-50\nsphere 1. .5 -0 1e0\n")
    hs = api.parse_scene_file(str(p)); d = hs.desc
    assert (d.width, d.height) == (8, 0)  # stoi("0x8") stops after the 0, base 10
    assert tuple(d.eye[:]) == (1.5, 0.5, -5.0) and d.expose == 2.5 and np.isinf(d.ior) and d.rough == 0.0
    assert np.array_equal(hs.array("sphere_c", (1, 3)), np.array([[1.0, 0.5, -0.0]], np.float32)) and np.signbit(hs.array("sphere_c", (1, 3))[0, 2])
    for bad in ("eye +-1 0 0\n", "eye . 0 0\n", "eye e5 0 0\n", "aa +\n"):
        p.write_text(bad)
        with pytest.raises(api.RtError):
            api.parse_scene_file(str(p))


# ---- row-block-parallel PNG encoder ------------------------------------------------------
def _decode_png_independently(path):
    """Minimal PNG reader on Python's zlib (checks every chunk CRC and, through zlib.decompress, the Adler-32 of the
    stitched stream); only filter types 0/1, which is what the writer emits."""
    import struct
    import zlib
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos = 8; idat = b""; ihdr = None; kinds = []
    while pos < len(b):
        (n,) = struct.unpack(">I", b[pos:pos + 4]); typ = b[pos + 4:pos + 8]; data = b[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + data) == crc, typ
        kinds.append(typ)
        if typ == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", data)
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    assert kinds[0] == b"IHDR" and kinds[-1] == b"IEND" and pos == len(b)
    w, h, depth, ctype, comp, filt, inter = ihdr
    assert (depth, ctype, comp, filt, inter) == (8, 6, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * 4 + 1)
    assert set(np.unique(raw[:, 0])) <= {0, 1}
    px = raw[:, 1:].reshape(h, w, 4)
    out = np.cumsum(px.astype(np.uint32), axis=1).astype(np.uint8)  # Sub filter: running sum over the pixels of a row, mod 256
    return np.where((raw[:, 0] == 1)[:, None, None], out, px)


@pytest.mark.parametrize("shape,threads", [((1, 1), 4), ((7, 3), 8), ((300, 257), 1), ((1100, 640), 3), ((2160, 1024), 8)])
def test_png_parallel_encoder_is_a_valid_png(shape, threads, tmp_path, monkeypatch):
    h, w = shape
    rng = np.random.RandomState(h * 131 + w)
    img = rng.randint(0, 256, (h, w, 4)).astype(np.uint8)
    img[h // 3:, : w // 2] = (10, 200, 30, 255)  # flat areas and noise
    monkeypatch.setenv("RT_PNG_THREADS", str(threads))
    p = str(tmp_path / "p.png")
    api.png_write(p, img)
    assert np.array_equal(_decode_png_independently(p), img)
    assert np.array_equal(api.png_read(p), img)
    monkeypatch.setenv("RT_PNG_THREADS", "1")
    q = str(tmp_path / "q.png")
    api.png_write(q, img)
    assert np.array_equal(_decode_png_independently(q), img)
