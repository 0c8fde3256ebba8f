"""Multi-GPU entry points of the C ABI (csrc/comm.cu): rt_comm_init, rt_scene_broadcast, rt_comm_update_triangles,
rt_comm_build_lbvh, rt_render_multi / rt_framebuffer_gather and `raytracer --gpus N`.

The image must not depend on the number of GPUs (RNG is keyed by the global pixel index, tiles are disjoint): the
N-GPU bytes are compared with the 1-GPU bytes.  The one-rank communicator runs on any box; everything with N >= 2
needs that many GPUs (`gpurun --gpus 2`) and is skipped otherwise.  No GPU at all: the calls must fail loudly."""
import os
import subprocess

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from cuda_ray_tracer_b200 import api, synth


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def test_comm_fails_loudly_without_enough_gpus():
    L = api.load_library()
    h = api.C.c_void_p()
    have = _ngpus()
    rc = L.rt_comm_init(have + 1, None, api.C.byref(h))
    assert rc != 0 and b"GPUs asked for" in L.rt_last_error()
    assert L.rt_comm_init(0, None, api.C.byref(h)) != 0


def test_tile_shape_c_matches_the_rule():
    # 8-row tiles; tiles per row must share as few factors with the rank count as possible, then least padding, then smallest
    assert api.tile_shape(8, 3840) == (256, 8)      # 15 tiles per row
    assert api.tile_shape(8, 1920) == (128, 8)      # 15 tiles per row
    assert api.tile_shape(1, 3840) == (32, 8)
    widths = (32, 64, 96, 128, 160, 256, 384, 512, 640, 768, 1280)
    for ranks in (2, 3, 4, 8):
        for width in (640, 1920, 3840, 1000, 100):
            tw, th = api.tile_shape(ranks, width)
            assert th == 8 and tw in widths
            g = np.gcd((width + tw - 1) // tw, ranks)
            assert g == min(np.gcd((width + a - 1) // a, ranks) for a in widths)


def _scene_desc(n=20000, w=160, h=96, aa=4):
    cfg, a = synth.soup(n, seed=5, width=w, height=h, aa=aa)
    a = synth.finish_triangles(a)
    a["bulbs"] = np.array([[0.5, 0.8, 2.5, 3, 3, 3]], np.float32)
    desc, keep = api.make_desc(cfg, a)
    return cfg, a, desc, keep


@pytest.mark.gpu
@pytest.mark.parametrize("ndev", [1, 2, 4])
def test_render_multi_equals_single_gpu(ndev):
    """rt_comm_init + rt_scene_broadcast + rt_comm_build_lbvh + rt_render_multi over ndev GPUs of one process: image bytes
    and ray count equal rt_render on one GPU; then a frame with new vertices through rt_comm_update_triangles."""
    import torch
    if _ngpus() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    cfg, a, desc, keep = _scene_desc()
    w, h = 160, 96
    one = api.Scene(desc, 0)
    one.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    comm = api.Comm(ndev)
    assert comm.size == ndev and comm.num_local == ndev and comm.ranks == list(range(ndev))
    scenes = comm.broadcast_scene(desc, 0)
    assert all(s.n == one.n for s in scenes)
    comm.build_lbvh(scenes, 30, api.BOUNDS_COMPUTED)
    e0 = one.export_reference_layout(30)
    for s in scenes:  # every GPU built the same tree from the broadcast arrays
        e = s.export_reference_layout(30)
        assert np.array_equal(e["codes"], e0["codes"]) and np.array_equal(e["refs"], e0["refs"]) and np.array_equal(e["bbox"].view(np.uint32), e0["bbox"].view(np.uint32))
    imgs = [torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda:%d" % d) for d in comm.devices]
    host = np.zeros((h, w, 4), np.uint8)
    for aa, fl in ((0, 0), (4, api.SPP_TRUE), (32, 0)):
        ref = one.render(w, h, aa, fl)
        cnt = comm.render(scenes, w, h, aa, imgs, flags=fl, host_out=host)
        assert np.array_equal(host, ref["rgba"]), "aa=%d: %d bytes differ" % (aa, (host != ref["rgba"]).sum())
        assert np.array_equal(imgs[0].cpu().numpy(), ref["rgba"])
        assert cnt.rays == ref["counters"].rays and cnt.samples == ref["counters"].samples
    # explicit tile shapes, including one whose tiles do not divide the image
    for tw, th in ((32, 8), (64, 4), (48, 5)):
        cnt = comm.render(scenes, w, h, 4, imgs, flags=api.SPP_TRUE, tile_w=tw, tile_h=th, host_out=host)
        assert np.array_equal(host, one.render(w, h, 4, api.SPP_TRUE)["rgba"])
    # a deforming frame: vertices from the root's host arrays to every GPU
    p0, p1, p2 = (np.ascontiguousarray(x + np.float32(0.01), np.float32) for x in a["tri"][:3])
    comm.update_triangles(scenes, p0, p1, p2, 0)
    comm.build_lbvh(scenes, 30, api.BOUNDS_COMPUTED)
    one.update_triangles(p0, p1, p2); one.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    ref = one.render(w, h, 4, api.SPP_TRUE)
    comm.render(scenes, w, h, 4, imgs, flags=api.SPP_TRUE, host_out=host)
    assert np.array_equal(host, ref["rgba"])
    for s in scenes:
        s.close()
    comm.close(); one.close()


@pytest.mark.gpu
@pytest.mark.parametrize("gpus", [2, 4])
def test_cli_multi_gpu_png_equals_single_gpu_png(gpus, tmp_path):
    """`raytracer scene.txt --gpus N` writes the same PNG, byte for byte, as the single-GPU run (main.cu:25-94 behaviour)"""
    if _ngpus() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    cfg, a = synth.soup(5000, seed=11, width=200, height=120, aa=4)
    path = str(tmp_path / "scene.txt")
    synth.write_scene_text(path, cfg, synth.finish_triangles(a), name="out.png")
    exe = os.path.join(os.path.dirname(api.__file__), "raytracer")
    d1 = tmp_path / "one"; dn = tmp_path / "many"; d1.mkdir(); dn.mkdir()
    r1 = subprocess.run([exe, path], capture_output=True, text=True, timeout=600, cwd=str(d1))
    rn = subprocess.run([exe, path, "--gpus", str(gpus), "--counters"], capture_output=True, text=True, timeout=600, cwd=str(dn))
    assert r1.returncode == 0, r1.stderr
    assert rn.returncode == 0, rn.stderr + rn.stdout
    for line in ("Initialize raw config time:", "Malloc and transfer to device time:", "LBVH Build time (N=5000):", "Render time:", "Transfer to host time:", "cudaFree time:"):
        assert line in rn.stdout and line in r1.stdout, (line, rn.stdout)
    assert open(str(d1 / "out.png"), "rb").read() == open(str(dn / "out.png"), "rb").read()
