"""A small pass over every kernel family of the library through the C ABI: every build variant and key width at sizes around
the warp / block / tile borders, the frame-loop update, the renderer through every traversal lane, rt_trace / rt_trace_static,
and the generated feature scenes (bulbs, mirrors, glass, rough, GI, DOF, fisheye, panorama).  Prints one line per step with a
running CRC of the outputs.

Used by tests/test_determinism_gpu.py: two fresh processes must print the same lines.  The reference's refit has an unfenced
race (src/lbvh_builder.cu:329-397); a race in the builder's climb, the sort's look-back or the ray queues would show
here as a checksum that changes from run to run.  (It was written to run under compute-sanitizer; that tool is closed on this
GPU pool, so run-to-run identity plus the oracle comparisons of tests/ are the evidence there is.)

Small on purpose: soups of up to 20 000 primitives, images of 96 x 64; about 5 s including the imports.  torch is used only
for the device buffers of rt_trace (--no-torch skips that part)."""
import os, sys, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cuda_ray_tracer_b200 import api, synth
import feature_scenes as F

no_torch = "--no-torch" in sys.argv
crc = 0


def note(what, *arrays):
    global crc
    for a in arrays:
        crc = zlib.crc32(np.ascontiguousarray(a).tobytes(), crc)
    print("%-64s crc %08x" % (what, crc), flush=True)


# ---- build: every variant, both key widths, sizes around the tile / block / warp borders ----
for n in (1, 2, 33, 257, 3000, 20000):
    cfg, a = synth.soup(n, seed=n, width=96, height=64, aa=0)
    a = synth.finish_triangles(a)
    desc, keep = api.make_desc(cfg, a)
    sc = api.Scene(desc, 0)
    for bits in (30, 63):
        for fl in (0, api.BUILD_GLOBAL_CLIMB, api.BUILD_LOCAL_CLIMB, api.BUILD_TOPDOWN, api.BUILD_WIDE4):
            sc.build_lbvh_karas(bits, api.BOUNDS_COMPUTED | fl)
            e = sc.export_reference_layout(bits)
            note("build n=%d bits=%d flags=%d" % (n, bits, fl), e["codes"], e["perm"], e["left"], e["right"], e["bbox"])
    sc.build_lbvh_karas(30, api.BOUNDS_FROM_DESC)  # the shipped reference: every code equal
    note("build n=%d shipped bounds" % n, sc.export_reference_layout(30)["left"])
    if n >= 3000:  # a frame loop step: new vertices, rebuild
        p0, p1, p2 = (np.ascontiguousarray(x + np.float32(0.01), np.float32) for x in a["tri"][:3])
        sc.update_triangles(p0, p1, p2); sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
        note("update_triangles + build n=%d" % n, sc.export_reference_layout(30)["perm"])
    sc.close()

# ---- render: the soup through every lane; feature scenes (bulbs, mirrors, glass, rough, GI, DOF, fisheye, panorama) ----
cfg, a = synth.soup(20000, seed=5, width=96, height=64, aa=4)
a = synth.finish_triangles(a)
a["bulbs"] = np.array([[0.5, 0.8, 2.5, 3, 3, 3]], np.float32)
desc, keep = api.make_desc(cfg, a)
sc = api.Scene(desc, 0)
for bfl in (0, api.BUILD_WIDE4):
    sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED | bfl)
    for aa, fl in ((0, 0), (4, api.SPP_TRUE), (4, api.SPP_TRUE | api.TRAVERSE_REF_ORDER), (4, api.SPP_TRUE | api.TRAVERSE_PACKET),
                   (4, api.SPP_TRUE | api.TRAVERSE_NEAR_FIRST), (4, api.SPP_TRUE | api.RENDER_SERIAL_PASSES), (32, api.COUNTERS)):
        out = sc.render(96, 64, aa, fl, want_ids=True)
        note("render soup build_flags=%d aa=%d flags=%d" % (bfl, aa, fl), out["rgba"], out["ids"])
if not no_torch:
    import torch
    rng = np.random.RandomState(1)
    n = 30000
    org = rng.uniform(-3, 3, (n, 3)).astype(np.float32); d = rng.uniform(-1, 1, (n, 3)).astype(np.float32) - org
    d = np.ascontiguousarray(d / np.linalg.norm(d, axis=1, keepdims=True), np.float32)
    to, td = torch.from_numpy(org).cuda(), torch.from_numpy(d).cuda()
    for fl in (0, api.TRAVERSE_REF_ORDER, api.TRAVERSE_PACKET, api.TRAVERSE_NEAR_FIRST):
        for static in (False, True):
            ids = torch.full((n,), -7, dtype=torch.int32, device="cuda"); ts = torch.zeros(n, dtype=torch.float32, device="cuda")
            (sc.trace_static if static else sc.trace_device)(n, to, td, ids, ts, flags=fl)
            torch.cuda.synchronize()
            note("trace flags=%d static=%d" % (fl, static), ids.cpu().numpy(), ts.cpu().numpy())
sc.close()
tmp = os.environ.get("TMPDIR", "/tmp")
for kind in F.KINDS:
    path = F.write(kind, os.path.join(tmp, "sanitize_%s.txt" % kind), w=64, h=48)
    hs = api.parse_scene_file(path)
    sc = api.Scene(hs.desc, 0)
    sc.build_lbvh_karas(30, api.BOUNDS_COMPUTED)
    for fl in (0, api.TRAVERSE_REF_ORDER):
        note("render feature scene %s flags=%d" % (kind, fl), sc.render(flags=fl)["rgba"])
    sc.close()
print("workload done crc %08x" % crc)
