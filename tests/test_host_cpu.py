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
    path = os.path.join(O.SCENE_DIR, name + ".txt")
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
            assert tw * th == 256
            own = parallel.tile_owner(w, 64, g, tw, th)
            # no rank owns whole columns of tiles: every 256-pixel-wide band of the image sees more than one rank
            if (w + tw - 1) // tw >= g:
                assert len(np.unique(own[:, : min(w, tw)])) > 1 or math.gcd((w + tw - 1) // tw, g) > 1
    assert parallel.tile_shape(8, 3840) == (256, 1) and parallel.tile_shape(1, 3840) == (32, 8)
    counts = np.bincount(parallel.tile_owner(3840, 2160, 8, 256, 1).ravel(), minlength=8)
    assert counts.max() / counts.min() < 1.01


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
