"""GPU parity of the LBVH build (K1..K6) against the CPU oracle: Morton codes, sorted permutation,
topology in the reference's node numbering, and every node's AABB — all bit-exact."""
import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from cuda_ray_tracer_b200 import api, synth

pytestmark = pytest.mark.gpu


def check_build(o, mode, bits, flags_extra=0):
    b = o.build(mode, bits)
    sc = H.gpu_scene_from_oracle(o, mode)
    flags = (api.BOUNDS_COMPUTED if mode == O.BOUNDS_COMPUTED else api.BOUNDS_FROM_DESC) | flags_extra
    ms = sc.build_lbvh_karas(bits, flags)
    e = sc.export_reference_layout(bits)
    n = b["n"]
    assert np.array_equal(e["codes"], b["codes"]), "sorted Morton codes differ"
    assert np.array_equal(e["perm"], b["perm"]), "sorted permutation differs"
    assert np.array_equal(e["refs"], b["refs"]), "sorted PrimitiveReference list differs"
    assert np.array_equal(e["left"][: n - 1], b["left"]) and np.array_equal(e["right"][: n - 1], b["right"]), "topology differs"
    assert np.array_equal(e["nleaf"], np.r_[np.zeros(n - 1, np.uint16), np.ones(n, np.uint16)])
    assert np.array_equal(e["prim"][n - 1:], np.arange(n, dtype=np.uint32))
    assert np.array_equal(e["bbox"].view(np.uint32), b["bbox"].view(np.uint32)), "node AABBs differ"
    sc.close()
    return ms


@pytest.mark.parametrize("name", H.SCENES)
@pytest.mark.parametrize("mode", [O.BOUNDS_SHIPPED, O.BOUNDS_COMPUTED])
@pytest.mark.parametrize("topdown", [0, api.BUILD_TOPDOWN, api.BUILD_LOCAL_CLIMB, api.BUILD_GLOBAL_CLIMB])
def test_build_scene_files(name, mode, topdown):
    p = H.scene_path(name)
    if p is None:
        pytest.skip("scene fixtures not present (oracle/_ref/scenes)")
    check_build(O.OracleScene(p), mode, 30, topdown)


def test_build_computed_bounds_with_host_bounds_in_desc():
    # RT_BOUNDS_FROM_DESC with the real box in the descriptor == RT_BOUNDS_COMPUTED
    p = H.scene_path("redchair")
    if p is None:
        pytest.skip("scene fixtures not present")
    o = O.OracleScene(p); b = o.build(O.BOUNDS_COMPUTED, 30)
    sc = H.gpu_scene_from_oracle(o, O.BOUNDS_COMPUTED)
    sc.build_lbvh_karas(30, api.BOUNDS_FROM_DESC)
    e = sc.export_reference_layout(30)
    assert np.array_equal(e["codes"], b["codes"]) and np.array_equal(e["perm"], b["perm"])


@pytest.mark.parametrize("n", [1, 2, 3, 7, 255, 256, 257, 511, 512, 513, 1024, 1025, 4095, 4096, 4097, 10000, 100003])
@pytest.mark.parametrize("bits", [30, 63])
def test_build_soup_sizes(n, bits):
    cfg, a = synth.soup(n, seed=n, width=32, height=32, aa=0)
    a = synth.finish_triangles(a)
    o = H.oracle_from_arrays(cfg, a)
    if n == 1:  # the reference crashes for N=1; here the root is the single leaf
        sc = H.gpu_scene_from_oracle(o, O.BOUNDS_COMPUTED)
        sc.build_lbvh_karas(bits, api.BOUNDS_COMPUTED)
        e = sc.export_reference_layout(bits)
        assert e["nleaf"][0] == 1 and e["perm"][0] == 0
        return
    for td in (0, api.BUILD_TOPDOWN, api.BUILD_LOCAL_CLIMB, api.BUILD_GLOBAL_CLIMB):
        check_build(o, O.BOUNDS_COMPUTED, bits, td)
    check_build(o, O.BOUNDS_SHIPPED, bits)


def test_build_mixed_duplicates():
    cfg, a = H.mixed_scene(5000)
    o = H.oracle_from_arrays(cfg, a)
    for bits in (30, 63):
        for td in (0, api.BUILD_TOPDOWN, api.BUILD_LOCAL_CLIMB, api.BUILD_GLOBAL_CLIMB):
            check_build(o, O.BOUNDS_COMPUTED, bits, td)


def test_build_1m_soup_properties_and_parity():
    n = 1_000_000
    cfg, a = synth.soup(n, seed=1, width=64, height=64, aa=0)
    a = synth.finish_triangles(a)
    o = H.oracle_from_arrays(cfg, a)
    ms = check_build(o, O.BOUNDS_COMPUTED, 30)
    ms_l = check_build(o, O.BOUNDS_COMPUTED, 30, api.BUILD_LOCAL_CLIMB)
    ms_g = check_build(o, O.BOUNDS_COMPUTED, 30, api.BUILD_GLOBAL_CLIMB)
    check_build(o, O.BOUNDS_COMPUTED, 63); check_build(o, O.BOUNDS_SHIPPED, 30)  # all codes equal: the deepest chains
    print("1M build ms: warp-first climb %.3f, global climb %.3f, block-local climb %.3f" % (ms, ms_g, ms_l))


def test_sort_is_stable_permutation_vs_torch():
    import torch
    n = 300_000
    cfg, a = synth.soup(n, seed=9, width=32, height=32, aa=0)
    a = synth.finish_triangles(a)
    o = H.oracle_from_arrays(cfg, a)
    sc = H.gpu_scene_from_oracle(o, O.BOUNDS_COMPUTED)
    sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    e = sc.export_reference_layout(30)
    codes_unsorted = torch.from_numpy(o.morton_unsorted(O.BOUNDS_COMPUTED, 30).astype(np.int64)).cuda()
    sk, sp = torch.sort(codes_unsorted, stable=True)
    assert np.array_equal(sp.cpu().numpy().astype(np.uint32), e["perm"])
    assert np.array_equal(sk.cpu().numpy().astype(np.uint32), e["codes"])


def test_update_triangles_matches_host_setup():
    n = 20000
    cfg, a = synth.soup(n, seed=5, width=32, height=32, aa=0)
    a = synth.finish_triangles(a)
    o = H.oracle_from_arrays(cfg, a)
    sc = H.gpu_scene_from_oracle(o, O.BOUNDS_COMPUTED)
    # move the vertices, rebuild on both sides
    rng = np.random.RandomState(0)
    d = rng.uniform(-0.01, 0.01, (n, 3)).astype(np.float32)
    p0, p1, p2 = a["tri"][0] + d, a["tri"][1] + d, a["tri"][2] - d
    a2 = dict(a); a2["tri"] = [p0, p1, p2, None, None, None]; a2 = synth.finish_triangles(a2)
    o2 = H.oracle_from_arrays(cfg, a2); b2 = o2.build(O.BOUNDS_COMPUTED, 30)
    sc.update_triangles(p0, p1, p2)
    sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    e = sc.export_reference_layout(30)
    assert np.array_equal(e["perm"], b2["perm"]) and np.array_equal(e["bbox"].view(np.uint32), b2["bbox"].view(np.uint32))


@pytest.mark.parametrize("bits", [30, 63])
def test_rebuild_sequence_equals_fresh_builds(bits):
    """A frame loop on ONE scene (update_triangles -> rt_build_lbvh, six frames of different vertices, no synchronisation in
    between): every frame's tree equals the tree a fresh scene builds from the same vertices.  The build's kernels are
    launched as a programmatic-dependent chain (csrc/build.cuh): blocks of the next kernel are resident while the previous
    one finishes, so anything read before the dependency wait, or served from a stale cache line, would show here."""
    n = 200_003
    cfg, a = synth.soup(n, seed=21, width=32, height=32, aa=0)
    a = synth.finish_triangles(a)
    desc, keep = api.make_desc(cfg, a)
    loop = api.Scene(desc, 0)
    rng = np.random.RandomState(4)
    for frame in range(6):
        shift = rng.uniform(-0.3, 0.3, (n, 3)).astype(np.float32) * np.float32(frame > 0)
        p0, p1, p2 = (np.ascontiguousarray(x + shift, np.float32) for x in a["tri"][:3])
        loop.update_triangles(p0, p1, p2)
        loop.build_lbvh_karas(bits, api.BOUNDS_COMPUTED, timed=False)
        loop.update_triangles(p0, p1, p2)                      # and once more right behind it: the frame loop's back-to-back case
        loop.build_lbvh_karas(bits, api.BOUNDS_COMPUTED, timed=False)
        e = loop.export_reference_layout(bits)
        cfg2, b = synth.soup(n, seed=21, width=32, height=32, aa=0)
        b["tri"] = [p0, p1, p2] + list(api.triangle_setup(p0, p1, p2))
        d2, k2 = api.make_desc(cfg2, b)
        fresh = api.Scene(d2, 0)
        fresh.build_lbvh_karas(bits, api.BOUNDS_COMPUTED | api.BUILD_TOPDOWN)
        f = fresh.export_reference_layout(bits)
        for key in ("codes", "perm", "left", "right"):
            assert np.array_equal(e[key], f[key]), "frame %d: %s differs from a fresh build" % (frame, key)
        assert np.array_equal(e["bbox"].view(np.uint32), f["bbox"].view(np.uint32)), "frame %d: boxes differ" % frame
        fresh.close()
    loop.close()
