"""rt_trace_static — rt_trace for callers that cannot allocate (no counterpart in the reference).
Its scratch is one fixed file-scope device array instead of buffers cudaMalloc'ed on demand.

Collected last on purpose (a failure here must not hide other tests).  The output must equal
the existing entry point's, rt_trace, bit for bit (ids and distances: the tolerance every trace test here uses is zero), on
two calls in a row with different rays (the array keeps the last call's contents), alone in a fresh process or after every
other test."""
import numpy as np
import pytest

import helpers as H
from cuda_ray_tracer_b200 import api

pytestmark = pytest.mark.gpu


def _rays(n, seed):
    rng = np.random.RandomState(seed)
    org = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    d = rng.uniform(-1, 1, (n, 3)).astype(np.float32) - org
    return org, np.ascontiguousarray(d / np.linalg.norm(d, axis=1, keepdims=True), np.float32)


def _trace(sc, org, d, static, flags=0):
    import torch
    to, td = torch.from_numpy(org).cuda(), torch.from_numpy(d).cuda()
    ids = torch.full((len(org),), -7, dtype=torch.int32, device="cuda"); ts = torch.full((len(org),), -7.0, dtype=torch.float32, device="cuda")
    if static:
        sc.trace_static(len(org), to, td, ids, ts, flags=flags)
    else:
        sc.trace_device(len(org), to, td, ids, ts, flags=flags)
    torch.cuda.synchronize()
    return ids.cpu().numpy(), ts.cpu().numpy()


def test_trace_static_equals_rt_trace_on_two_calls_in_a_row():
    import torch
    L = api.load_library()
    nmax = L.rt_trace_static_max_rays()
    assert nmax == 131072
    cfg, a = H.mixed_scene(4000)   # triangles + spheres, duplicate Morton codes
    a["planes"] = np.array([[0, 1, 0, 1.2] + [0, 1, 0] + [0, -1.2, 0] + [0.5, 0.5, 0.5, 0, 0, 0, 0, 0, 0, 1.458, 0]], np.float32)
    desc, keep = api.make_desc(cfg, a)
    sc = api.Scene(desc, 0)
    sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    # two calls in a row (a full-size batch, then a smaller one with other rays: stale contents of the shared array would show),
    # then the other lanes
    for call, (n, seed, flags) in enumerate(((nmax, 1, 0), (50_001, 2, 0), (nmax, 3, api.TRAVERSE_REF_ORDER), (30_000, 4, api.TRAVERSE_PACKET))):
        org, d = _rays(n, seed)
        si, st = _trace(sc, org, d, static=True, flags=flags)
        ri, rt_ = _trace(sc, org, d, static=False, flags=flags)
        assert (ri >= 0).mean() > 0.1 and (ri == -1).any()
        assert np.array_equal(si, ri), "call %d: %d hit ids differ from rt_trace" % (call, (si != ri).sum())
        assert np.array_equal(st.view(np.uint32), rt_.view(np.uint32)), "call %d: hit distances differ from rt_trace" % call
    # the call itself takes no device memory (inputs and outputs exist, the kernels are loaded by now)
    org, d = _rays(nmax, 6)
    to, td = torch.from_numpy(org).cuda(), torch.from_numpy(d).cuda()
    ids = torch.empty(nmax, dtype=torch.int32, device="cuda"); ts = torch.empty(nmax, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    sc.trace_static(nmax, to, td, ids, ts)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "rt_trace_static changed the free device memory"
    # beyond the static scratch: a clear error, nothing launched
    org, d = _rays(nmax + 1, 5)
    with pytest.raises(api.RtError, match="exceed the static scratch"):
        _trace(sc, org, d, static=True)
    # n == 0 and an unbuilt scene behave like rt_trace
    sc.trace_static(0, torch.zeros(3, device="cuda"), torch.zeros(3, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda"))
    sc2 = api.Scene(desc, 0)
    with pytest.raises(api.RtError, match="rt_build_lbvh has not run"):
        _trace(sc2, org[:10], d[:10], static=True)
    sc.close(); sc2.close()
