#!/usr/bin/env python
"""bench.py — headline benchmark: Mrays/s on a 1M-triangle soup rendered at 4K, LBVH rebuilt every frame.

One "step" = one frame of the hot path: new vertex positions -> rt_scene_update_triangles (Triangle
constructor on the device) -> rt_build_lbvh (K1..K6) -> rt_render (generate / traverse / shade / resolve).
`value` times it with the frame's vertices already resident in HBM and the image left on the device;
`e2e` times the same frame through the host-buffer entry points (pinned host vertices in, image out).
With N > 1 the scene is replicated, the image is split into interleaved 32x8 tiles (tile t -> rank t % N)
and the framebuffer is summed onto rank 0 over NCCL inside the timed region: total work is fixed, so
scaling is "strong".

`--impl reference` times the reference's own CUDA program (oracle/_ref/ref_gpu_fast: unmodified kernels compiled
for sm_100, the baseline BASELINE.json names) on the same GPU and scene, and carries the reference's traversal + shading
sources compiled for the host (oracle/_ref/libref_host.so, all host threads, a bounded sample of rows) as its
cpu_baseline leg; the fork has no CPU renderer of its own."""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "Mrays/s @4K, 1M tris + LBVH build ms"
# the reference keeps the vertex index of a `tri` line in a float (src/parse.cpp:175-180): past 2^24 vertices it reads wrong or
# out-of-range vertices, so its program cannot load a soup of more than 5 592 405 triangles
REF_MAX_VERTICES = 1 << 24
REF_VERTEX_LIMIT_NOTE = ("the reference cannot load this scene: its float-typed vertex index (src/parse.cpp:175-180) is exact only below 2^24 vertices "
                         "(5 592 405 soup triangles); the reference ratio for a large scene is taken at 5.5M triangles, profiles/vs_gpu_reference_5p5m_r02.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--near-first", action="store_true", help="RT_TRAVERSE_NEAR_FIRST lane (not bit-exact in rare cases)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--overlap-build", action="store_true", help="device leg: RT_RENDER_OVERLAP_BUILD (the frame's ray generation runs beside update+build), as the e2e leg always does")
    ap.add_argument("--wide4", action="store_true", help="EXPERIMENTAL: RT_BUILD_WIDE4, the near-first lanes step through a 4-wide tree of grandchildren")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip timing the reference's own CUDA build (oracle/_ref/ref_gpu_fast) on this GPU")
    return ap.parse_args()


def workload_name(a):
    return "soup(N=%d,seed=1) %dx%d %dspp (true-spp), 1 sun, LBVH rebuilt every frame (vertices jittered per frame)" % (
        a.tris, a.width, a.height, a.spp)


def frame_vertices(arrays, frame):
    """vertex positions of frame `frame`: the soup displaced by a deterministic per-frame offset field"""
    p0, p1, p2 = arrays["tri"][:3]
    ph = np.float32(0.37 * (frame + 1))
    d = (np.float32(0.002) * np.sin(p0 * np.float32(7.0) + ph)).astype(np.float32)
    return (p0 + d).astype(np.float32), (p1 + d).astype(np.float32), (p2 + d).astype(np.float32)


class ClockSampler:
    """SM clock and throttle reasons of the timed region.  NVML is queried in this process BETWEEN the timed steps:
    right after a step's closing synchronize (the GPU has just been under load for the whole step, the clock domain has
    not moved yet) and before the L2 flush, i.e. inside the timed loop but outside every CUDA-event interval.
    A concurrent sampler (a thread polling NVML every 20 ms, or a helper process every 25 ms) was measured to
    slow a 7.5 ms frame to 8.3 - 13.8 ms: NVML calls contend with the render loop's CUDA calls for driver locks, and
    the loop waits on the host several times per pass.  Fallback without pynvml: nvidia-smi (the recipe's clocks
    line) at the same points, first and last step only (it needs ~100 ms per query)."""
    REASONS = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]

    def __init__(self, index):
        self.index = index; self.sm = []; self.mx = []; self.mask = 0; self.how = "nvml, after every timed step"; self.h = None; self.calls = 0
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)))
        except Exception:
            self.h = None; self.how = "nvidia-smi, after the first and the last timed step"

    def sample(self, last=False):
        self.calls += 1
        if self.h is not None:
            try:
                self.sm.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                try:
                    self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            return
        if self.calls > 1 and not last:
            return
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=5).stdout.strip()
            r = [x.strip() for x in out.split(",")]
            self.sm.append(float(r[0])); self.mx.append(float(r[1]))
            for k, (_, bit) in enumerate(self.REASONS):
                if r[2 + k].lower().startswith("active"):
                    self.mask |= bit
        except Exception:
            pass

    def summary(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": [n for n, bit in self.REASONS if self.mask & bit], "samples": len(self.sm), "how": self.how}


# ------------------------------------------------------------------------------------------------
def run_reference(a):
    """Reference arm.  BASELINE.json names the baseline: "the reference's own CUDA build compiled unmodified for sm_100
    in the same run" — the reference is a CUDA program with no CPU renderer.  So `value` is oracle/_ref/ref_gpu_fast
    (the reference's unmodified kernels, nvcc -O3 defaults, sm_100; oracle/ref_gpu_driver.cu in place of main.cu) on
    this box's GPU, one step = its build_lbvh_karas + render() of the same soup at the same image size; none of this
    library's kernels run inside the timed steps.  The `cpu_baseline` leg is the reference's traversal + shading
    sources compiled for the HOST (oracle/_ref/libref_host.so) on every host core, on a bounded sample of rows.
    Without a GPU or without the prebuilt binary the CPU leg is the line's value."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib as O
    from cuda_ray_tracer_b200 import synth
    cfg, arrays = synth.soup(a.tris, 1, a.width, a.height, a.spp)
    p0, p1, p2 = frame_vertices(arrays, 0)
    host_arrays = {"tri_mat": arrays["tri_mat"], "refs": arrays["refs"], "suns": arrays["suns"]}
    cpu = cpu_baseline(a, cfg, dict(host_arrays, tri=arrays["tri"][:3]))  # displaces the vertices to frame 0 itself
    W, Hh = a.width, a.height
    aa = 32 if a.spp >= 2 else a.spp
    total = a.steps + a.warmup
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_fast")
    gpu = None
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    if have_gpu and os.path.exists(exe) and 3 * a.tris >= REF_MAX_VERTICES:
        gpu = {"unavailable": REF_VERTEX_LIMIT_NOTE}
    elif have_gpu and os.path.exists(exe):
        from cuda_ray_tracer_b200 import api
        nor, e1, e2 = O.triangle_setup(p0, p1, p2)
        arr = dict(host_arrays); arr["tri"] = [p0, p1, p2, nor, e1, e2]
        path = "/tmp/bench_soup_%d.txt" % a.tris
        synth.write_scene_text(path, cfg, arr, "bench_soup.png")
        p = subprocess.run([exe, path, "--bounds", "computed", "--w", str(W), "--h", str(Hh), "--aa", str(aa), "--repeat", str(total)],
                           capture_output=True, text=True, timeout=3000, cwd="/tmp")
        if p.returncode == 0:
            info = json.loads(p.stdout.strip().splitlines()[-1])
            # the reference carries no ray counter; the ray count is a property of (scene, seed, sample semantics): counted once,
            # outside the timed steps, by this library rendering the same file with the reference's sample semantics
            hs = api.parse_scene_file(path)
            mine = api.Scene(hs.desc, 0)
            mine.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
            rays = mine.render(W, Hh, aa, 0)["counters"].rays
            mine.close()
            step_s = [r + b * 1e-3 for r, b in zip(info["render_s"], info["build_ms"])][a.warmup:]
            gpu = {"value": rays / float(np.mean(step_s)) / 1e6, "ms_per_step": float(np.mean(step_s)) * 1e3, "rays_per_step": rays,
                   "render_s": info["render_s"][a.warmup:], "build_ms": info["build_ms"][a.warmup:], "parse_s": info["parse_s"]}
        else:
            gpu = {"error": (p.stdout[-300:] + p.stderr[-300:])}
    if gpu is not None and "value" in gpu:
        v = gpu["value"]; ms = gpu["ms_per_step"]; kind = "reference-cuda"
        sample = ("the reference's CUDA program on this GPU: build_lbvh_karas + render() of soup frame 0 at %dx%d, aa=%d (it renders 32 samples per pixel "
                  "whenever aa >= 2, SURVEY F3; Mrays/s = hitNearest calls of that frame / step time), scene bounds stored (as shipped every Morton code is 0)" % (W, Hh, aa))
    else:
        v = cpu["value"]; ms = None; kind = cpu["kind"]; sample = cpu["sample"]
    line = {"metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference", "reference_kind": kind,
            "config": {"workload": workload_name(a), "reference_arm": sample},
            "cpu_baseline": cpu,
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if gpu is not None:
        line["reference_cuda"] = gpu
        if "build_ms" in gpu:
            line["lbvh_build_ms"] = float(np.mean(gpu["build_ms"]))
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
def cpu_baseline(a, cfg, arrays):
    import oracle_lib as O
    import ref_lib as R
    import helpers as H
    cores = os.cpu_count() or 1
    p0, p1, p2 = frame_vertices(arrays, 0)
    nor, e1, e2 = O.triangle_setup(p0, p1, p2)
    arr = dict(arrays); arr["tri"] = [p0, p1, p2, nor, e1, e2]
    o = H.oracle_from_arrays(cfg, arr)
    b = o.build(O.BOUNDS_COMPUTED, 30)
    kind = "port"; ref = None
    if R.available():
        kind = "reference"; ref = R.scene_from_arrays(cfg, arr); ref.set_bvh(b)
    W, Hh = a.width, a.height
    aa = a.spp if a.spp < 2 else 32
    mid = Hh // 2

    def go(y0, y1):
        t = time.time()
        if ref is not None:
            ref.render(W, Hh, aa, y0, y1, cores)
        else:
            o.render(W, Hh, aa, O.SPP_REF, y0, y1, cores)
        return time.time() - t

    t_probe = go(mid, mid + 1)
    rows = max(1, min(Hh // 8, int(a.cpu_seconds / max(t_probe, 1e-3))))
    dt = go(mid, mid + rows)
    if dt < 0.6 * a.cpu_seconds and rows < Hh // 8:  # the 1-row probe overestimates (thread start-up): rescale once
        rows = max(rows, min(Hh // 8, int(rows * a.cpu_seconds / max(dt, 1e-3))))
        dt = go(mid, mid + rows)
    rays = o.render(W, Hh, aa, O.SPP_REF, mid, mid + rows, cores)["counters"].rays
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
            "sample": "%d image rows x %d px x %d spp (reference render() semantics) = %d rays in %.1f s" % (rows, W, aa, rays, dt)}


def gpu_baseline(a, sc, cfg, arrays, api, synth):
    """BASELINE.json's reported baseline: the reference's own CUDA program, compiled unmodified for sm_100
    (oracle/_ref/ref_gpu_fast: nvcc defaults, FMA contraction on), on this GPU, same scene.  The reference renders 32
    samples per pixel whenever aa >= 2 and has no ray counter, so both sides render the frame-0
    scene at aa = 32 in reference sample semantics (identical images, tests/test_vs_reference_gpu.py) and the
    ray count of this library's counters is divided by each side's time.  Rank 0, N = 1 only."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_fast")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_gpu_fast was not built (needs /root/reference at build time)"}
    if 3 * a.tris >= REF_MAX_VERTICES:
        return {"unavailable": REF_VERTEX_LIMIT_NOTE}
    import oracle_lib as O
    p0, p1, p2 = frame_vertices(arrays, 0)
    nor, e1, e2 = O.triangle_setup(p0, p1, p2)
    arr = dict(arrays); arr["tri"] = [p0, p1, p2, nor, e1, e2]
    path = "/tmp/bench_soup_%d.txt" % a.tris
    synth.write_scene_text(path, cfg, arr, "bench_soup.png")
    W, Hh = a.width, a.height
    aa = 32 if a.spp >= 2 else a.spp
    p = subprocess.run([exe, path, "--bounds", "computed", "--w", str(W), "--h", str(Hh), "--aa", str(aa), "--repeat", "3"],
                       capture_output=True, text=True, timeout=1200, cwd="/tmp")
    if p.returncode != 0:
        return {"error": (p.stdout[-300:] + p.stderr[-300:])}
    info = json.loads(p.stdout.strip().splitlines()[-1])
    hs = api.parse_scene_file(path)          # the same text file through this library's parser
    mine = api.Scene(hs.desc, sc.device)
    mine.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    best = None
    for _ in range(3):
        r = mine.render(W, Hh, aa, 0)
        c = r["counters"]
        best = c.render_ms if best is None else min(best, c.render_ms)
    rays = c.rays
    return {"kind": "reference CUDA build (unmodified kernels, sm_100, nvcc -O3 defaults), same B200, same scene file",
            "workload": "soup(N=%d) frame 0, %dx%d, aa=%d in reference sample semantics (32 spp), %d rays" % (a.tris, W, Hh, aa, rays),
            "reference_mrays_s": rays / min(info["render_s"]) / 1e6, "reference_render_s": min(info["render_s"]),
            "reference_build_ms": min(info["build_ms"]), "reference_parse_s": info["parse_s"],
            "b200_mrays_s": rays / best / 1e3, "b200_render_s": best * 1e-3, "host_cores": os.cpu_count()}


def measure_l2_peaks():
    """L2 ceilings of this GPU for the traversal kernel's access shapes, measured in this run by scripts/microbench/l2bw
    (built by __graft_entry__.build(); 64 MB L2-resident array, all SMs): a coalesced stream of 32-byte loads, 64-byte
    records at random indices with one record per lane, and the same with lane pairs sharing a record's line."""
    exe = os.path.join(ROOT, "scripts", "microbench", "l2bw")
    if not os.path.exists(exe):
        return {"unavailable": "scripts/microbench/l2bw is not built"}
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout
        import re
        g = lambda pat: float(re.search(pat + r"[^:]*:\s*([0-9.]+) GB/s", out).group(1))
        return {"l2_stream_gbs": g("L2 stream"), "l2_gather64_gbs": g("L2 gather, 64-byte records at random"),
                "l2_gather64_pair_shfl_gbs": g(r"lane pairs \+ 8 shuffles"), "hbm_gather64_gbs": g("HBM gather"), "how": "scripts/microbench/l2bw, this run"}
    except Exception as e:
        return {"unavailable": repr(e)}


def run_b200(a):
    import torch
    import torch.distributed as dist
    from cuda_ray_tracer_b200 import api, synth, parallel
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    api.load_library()  # raises when libb200rt.so is missing: there is no fallback
    sampler = ClockSampler(local) if (rank == 0 and not os.environ.get("BENCH_NO_SAMPLER")) else None

    cfg, arrays = synth.soup(a.tris, 1, a.width, a.height, a.spp)
    W, Hh, spp, n = a.width, a.height, a.spp, a.tris
    # scene arrays: rank 0 generates, NCCL broadcasts (scene replication)
    p0, p1, p2 = arrays["tri"][:3]
    tp = [torch.from_numpy(x).to(dev) for x in (p0, p1, p2)]
    tm = torch.from_numpy(arrays["tri_mat"]).to(dev)
    parallel.broadcast_scene(tp + [tm], 0)
    nor, e1, e2 = api.triangle_setup(p0, p1, p2)
    arrays["tri"] = [tp[0], tp[1], tp[2]] + [torch.from_numpy(x).to(dev) for x in (nor, e1, e2)]
    arrays["tri_mat"] = tm
    arrays["refs"] = torch.from_numpy(arrays["refs"]).to(dev)
    desc, keep = api.make_desc(cfg, arrays)
    sc = api.Scene(desc, local)

    nframes = 4
    frames_h = [frame_vertices({"tri": [p0, p1, p2]}, f) for f in range(nframes)]
    frames_pinned = [[torch.from_numpy(x).pin_memory() for x in fr] for fr in frames_h]
    frames_d = [[torch.from_numpy(x).to(dev) for x in fr] for fr in frames_h]
    img_d = torch.zeros((Hh, W, 4), dtype=torch.uint8, device=dev)
    img_h = torch.zeros((Hh, W, 4), dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tiles = parallel.tile_set(world, rank, width=W)
    # N > 1: a rank renders 1/N of the image but updates and builds the whole scene, so the frame's ray generation runs beside
    # update + build (RT_RENDER_OVERLAP_BUILD) on the device leg too; at N = 1 generation fills the GPU anyway and the plain
    # stream order keeps lbvh_build_ms the time of the build running alone
    overlap_dev = a.overlap_build or world > 1
    rflags = api.SPP_TRUE | (api.TRAVERSE_NEAR_FIRST if a.near_first else 0) | (api.RENDER_OVERLAP_BUILD if overlap_dev else 0)
    # launch durations of the dominant kernel (roofline) are taken with one pass at a time: with two passes in flight
    # (the default, and what `value` and `e2e` run) a launch's event interval includes waiting for the other pass's CTAs
    rflags_serial = rflags | api.TIME_TRAVERSE | api.RENDER_SERIAL_PASSES
    # e2e leg: the frame's ray generation starts beside the vertex upload (copy engine) and the build.  Measured on the
    # device leg it gains nothing (generation already fills the GPU, the build just stretches), so that leg keeps the
    # plain stream order and its lbvh_build_ms is the build running alone.
    rflags_e2e = rflags | api.RENDER_OVERLAP_BUILD
    BUILD_LAUNCHES = 1 + 8 + (1 if a.wide4 else 0)  # update_tris + init, bounds, morton, 4 onesweep passes, bottom-up (+ collapse)
    bflags = api.BOUNDS_COMPUTED | (api.BUILD_WIDE4 if a.wide4 else 0)
    node_bytes = 128.0 if a.wide4 else 64.0  # bytes of one node step

    def step_device(f, cnt, flags=None):
        fr = frames_d[f % nframes]
        sc.update_triangles(fr[0], fr[1], fr[2])
        sc.build_lbvh_karas(30, bflags, timed=False)  # asynchronous: its device time is read after the frame
        sc.render_device(W, Hh, spp, img_d, flags=rflags if flags is None else flags, tiles=tiles, counters=cnt)
        parallel.gather_framebuffer(img_d, 0)  # tiles are disjoint and the rest is zero: sum == gather
        return sc.build_ms()

    def step_device_serial(f, cnt):
        return step_device(f, cnt, rflags_serial)

    rows_per_rank = (n + world - 1) // world
    if world > 1:
        up_full = [torch.empty((rows_per_rank * world, 3), dtype=torch.float32, device=dev) for _ in range(3)]
        up_part = [torch.empty((rows_per_rank, 3), dtype=torch.float32, device=dev) for _ in range(3)]

    def step_host(f, cnt):
        fr = frames_pinned[f % nframes]
        if world > 1:  # every rank uploads 1/N of the vertices over its own PCIe link, NCCL all-gather over NVLink
            sc.begin_frame()
            full = [parallel.upload_sharded(fr[k], up_full[k], up_part[k], rank, world) for k in range(3)]
            sc.update_triangles(full[0], full[1], full[2])
        else:
            sc.update_triangles(fr[0], fr[1], fr[2])   # host pointers: H2D inside
        sc.build_lbvh_karas(30, bflags, timed=False)
        if world > 1:
            sc.render_device(W, Hh, spp, img_d, flags=rflags_e2e, tiles=tiles, counters=cnt)
            parallel.gather_framebuffer(img_d, 0)
            if rank == 0:
                img_h.copy_(img_d)
        else:
            sc.render_host_into(img_h, W, Hh, spp, flags=rflags_e2e, counters=cnt)
        return sc.build_ms()

    def timed(fn, steps, warmup, smp=None):
        cnt = api.Counters()
        if world > 1:
            dist.barrier()  # the first barrier of a leg has one-time costs (observed: a 9.9 ms first timed step at 8 GPUs): pay them before the warm-up
        for f in range(warmup):
            fn(f, cnt); flush.fill_(f & 255)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

        ms = []; rays = 0; bms = []; trav_ms = 0.0; trav_launches = 0; launches = 0
        for f in range(steps):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            b = fn(warmup + f, cnt)
            e1.record(); torch.cuda.synchronize()
            if smp is not None:
                smp.sample(last=(f == steps - 1))  # between the timed intervals: see ClockSampler
            ms.append(e0.elapsed_time(e1)); bms.append(b); rays += cnt.rays
            trav_ms += cnt.traverse_ms; trav_launches += cnt.traverse_launches; launches += cnt.kernel_launches + BUILD_LAUNCHES
            flush.fill_(f & 255)  # L2 flush between timed iterations (256 MiB > 126 MB L2), outside the timed events
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        total_ms = sum(ms)
        t = torch.tensor([total_ms, float(rays), trav_ms, float(trav_launches), float(launches), sum(bms)], dtype=torch.float64, device=dev)
        if world > 1:
            mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            return dict(ms=mx[0].item(), rays=sm[1].item(), trav_ms=mx[2].item(), trav_launches=t[3].item(), launches=sm[4].item(), build_ms=mx[5].item() / steps, steps_ms=ms)
        return dict(ms=total_ms, rays=float(rays), trav_ms=trav_ms, trav_launches=float(trav_launches), launches=float(launches), build_ms=sum(bms) / steps, steps_ms=ms)

    dev_res = timed(step_device, a.steps, a.warmup, sampler)
    e2e_res = timed(step_host, a.steps, max(1, a.warmup - 1))
    ser_res = timed(step_device_serial, a.steps, 1)  # same frames, one pass at a time, traverse launches timed

    # algorithmic bytes of the dominant kernel (traverse_kernel): one instrumented frame, same frame data
    cnt = api.Counters()
    fr = frames_d[a.warmup % nframes]
    sc.update_triangles(fr[0], fr[1], fr[2]); torch.cuda.synchronize()
    build_alone = min(sc.build_lbvh_karas(30, bflags) for _ in range(3))  # nothing else on the GPU
    sc.render_device(W, Hh, spp, img_d, flags=(rflags & ~api.TIME_TRAVERSE) | api.COUNTERS, tiles=tiles, counters=cnt)
    torch.cuda.synchronize()
    alg = torch.tensor([node_bytes * cnt.nodes_visited + 48.0 * cnt.leaf_tests, float(cnt.nodes_visited), float(cnt.leaf_tests), float(cnt.rays),
                        float(cnt.packet_steps)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(alg, op=dist.ReduceOp.SUM)
    # diagnostics for the scaling figure: per-rank render time of one frame (load balance) and the cost of the gather alone
    diag = {}
    if world > 1:
        c2 = api.Counters()
        sc.render_device(W, Hh, spp, img_d, flags=rflags, tiles=tiles, counters=c2)
        t = torch.zeros(world, dtype=torch.float64, device=dev); t[rank] = c2.render_ms
        dist.all_reduce(t)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        parallel.gather_framebuffer(img_d, 0); torch.cuda.synchronize(); dist.barrier()
        e0.record()
        for _ in range(5):
            parallel.gather_framebuffer(img_d, 0)
        e1.record(); torch.cuda.synchronize()
        diag = {"rank_render_ms": [round(x, 3) for x in t.tolist()], "gather_ms": e0.elapsed_time(e1) / 5}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None; traffic_src = None  # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this command
    try:
        prof = "ncu_traverse_r02.json" if a.tris == 1_000_000 else "ncu_traverse_%dm_r02.json" % (a.tris // 1_000_000)
        pj = json.load(open(os.path.join(ROOT, "profiles", prof)))
        traffic = pj["dram_bytes_per_launch_mean"]; traffic_src = "profiles/%s (captured at commit %s)" % (prof, pj.get("commit", "?"))
    except Exception:
        pass
    # per-launch figures: algorithmic bytes of one frame / traverse launches of one frame; duration = live CUDA-event average
    launches_per_frame = ser_res["trav_launches"] / a.steps
    bytes_per_launch = alg[0].item() / world / max(launches_per_frame, 1)   # per rank, like the per-rank event time
    ms_per_launch = ser_res["trav_ms"] / max(ser_res["trav_launches"], 1)
    achieved = bytes_per_launch / (ms_per_launch * 1e-3) / 1e9 if ms_per_launch > 0 else 0.0
    l2 = measure_l2_peaks()  # after the timed loops: same GPU, same run
    value = dev_res["rays"] / (dev_res["ms"] * 1e-3) / 1e6
    e2e = e2e_res["rays"] / (e2e_res["ms"] * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": dev_res["ms"] / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(a), "lane": ("near-first (not bit-exact)" if a.near_first else "exact near-first (bit-exact vs reference order)") + (", 4-wide tree (experimental)" if a.wide4 else ""),
                   "l2": "256 MiB flush write between timed steps; per-frame working set (state pool + scene) exceeds the 126 MB L2",
                   "parallelism": "scene replicated, image in interleaved %dx%d tiles over %d GPU(s), NCCL broadcast of the scene + gather of each rank's own tiles (rt_framebuffer_gather); two passes in flight per GPU%s" % (parallel.tile_shape(world, W) + (world, "; ray generation overlaps update + build" if overlap_dev else ""))},
        "ms_steps_rank0": {"device": [round(x, 3) for x in dev_res["steps_ms"]], "e2e": [round(x, 3) for x in e2e_res["steps_ms"]],
                           "serial_passes": [round(x, 3) for x in ser_res["steps_ms"]]},  # this rank's own event time of every timed step
        "ms_per_step_serial_passes": ser_res["ms"] / a.steps,
        "lbvh_build_ms": dev_res["build_ms"],  # inside the timed frames (asynchronous, read back after each frame)
        "lbvh_build_ms_alone": build_alone,   # best of 3 with an idle GPU and a synchronising call
        "rays_per_step": dev_res["rays"] / a.steps,
        "nodes_per_ray": alg[1].item() / max(alg[3].item(), 1), "leaf_tests_per_ray": alg[2].item() / max(alg[3].item(), 1),
        "packet_lanes_per_node_step": (alg[1].item() / alg[4].item()) if alg[4].item() > 0 else None,  # of 32: lanes that wanted the node their warp fetched
        "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": 3 * 12 * n, "d2h_bytes_per_step": W * Hh * 4,
                "ms_per_step": e2e_res["ms"] / a.steps},
        "gpu_launches": int(dev_res["launches"]),
        "roofline": {"bound": "l2", "achieved": achieved, "peak": l2.get("l2_stream_gbs"), "unit": "GB/s",
                     "frac": (achieved / l2["l2_stream_gbs"]) if l2.get("l2_stream_gbs") else None,
                     "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "traverse_near_kernel<EXACT=%d>" % (0 if a.near_first else 1),
                     "peak_source": "L2 stream rate of a 64 MB resident array, measured in this run (scripts/microbench/l2bw); MEASURED_PEAKS.json has no L2 figure",
                     "bytes_per_launch": bytes_per_launch, "ms_per_launch": ms_per_launch,
                     "share_of_step": ser_res["trav_ms"] / ser_res["ms"],
                     "timed_how": "CUDA events around every traverse launch of %d frames rendered one pass at a time (RT_RENDER_SERIAL_PASSES), %.2f ms per frame; the headline frames keep two passes in flight, where a launch interval would include waiting for the other pass" % (a.steps, ser_res["ms"] / a.steps),
                     "hbm_peak": peak, "frac_hbm": achieved / peak, "hbm_peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                     "l2_gather_peak": l2.get("l2_gather64_gbs"), "frac_l2_gather": (achieved / l2["l2_gather64_gbs"]) if l2.get("l2_gather64_gbs") else None,
                     "l2_gather_pair_peak": l2.get("l2_gather64_pair_shfl_gbs"), "l2_measured": l2,
                     "note": "algorithmic bytes = 64 B per internal node a ray visits + 48 B per triangle test (SURVEY 8d), counted by an instrumented frame. The 64 MB of nodes and 96 MB of leaf records are served by L1 and L2 (DRAM traffic = the ray records in and the hits out, `traffic`), so the bound is the L2 and `frac` is against its measured stream rate; frac_hbm is kept for the record and exceeds 1 because HBM is not what feeds the kernel. l2_gather_peak is what 32 lanes reading 32 unrelated 64-byte records reach from L2 (one L1TEX wavefront per lane per 32-byte load): the pixel-major queue order lets the kernel EXCEED it, because lanes of a warp that trace samples of the same pixel read the same lines (one wavefront for several lanes, L1 hits)"},
        "clocks": sampler.summary() if sampler is not None else None,
    }
    if diag:
        line["scaling_diag"] = diag
    if not a.no_cpu_baseline and world == 1:
        try:
            line["cpu_baseline"] = cpu_baseline(a, cfg, {"tri": [p0, p1, p2], "tri_mat": arrays["tri_mat"].cpu().numpy(),
                                                       "refs": arrays["refs"].cpu().numpy(), "suns": arrays["suns"]})
        except Exception as e:  # the baseline is a reported figure, never the product path
            line["cpu_baseline"] = {"error": repr(e)}
    if not a.no_gpu_baseline and world == 1:
        try:
            line["gpu_baseline"] = gpu_baseline(a, sc, cfg, {"tri": [p0, p1, p2], "tri_mat": arrays["tri_mat"].cpu().numpy(),
                                                            "refs": arrays["refs"].cpu().numpy(), "suns": arrays["suns"]}, api, synth)
        except Exception as e:
            line["gpu_baseline"] = {"error": repr(e)}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
