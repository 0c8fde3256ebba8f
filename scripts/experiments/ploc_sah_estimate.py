import sys, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, helpers as H, oracle_lib as O
from cuda_ray_tracer_b200 import synth
n = int(sys.argv[1]); R = int(sys.argv[2]) if len(sys.argv) > 2 else 8
cfg, a = synth.soup(n, seed=1, width=64, height=64, aa=0); a = synth.finish_triangles(a)
o = H.oracle_from_arrays(cfg, a); b = o.build(O.BOUNDS_COMPUTED, 30)
bb = b["bbox"].reshape(-1, 6)  # xmin,xmax,ymin,ymax,zmin,zmax ; internal [0,n-1), leaves [n-1,2n-1)
def area(lo, hi):
    d = hi - lo; return 2 * (d[:, 0] * d[:, 1] + d[:, 1] * d[:, 2] + d[:, 0] * d[:, 2])
lo = bb[:, 0::2]; hi = bb[:, 1::2]
A = area(lo, hi)
root = A[0]
print("LBVH  SAH (sum internal area / root area): %.2f   leaves: %.2f" % (A[: n - 1].sum() / root, A[n - 1:].sum() / root))
# PLOC over the leaves in Morton order
clo = lo[n - 1:].copy(); chi = hi[n - 1:].copy()
tot = 0.0; it = 0; t0 = time.time()
while len(clo) > 1:
    m = len(clo); it += 1
    best = np.full(m, np.inf); nn = np.full(m, -1)
    for off in list(range(-R, 0)) + list(range(1, R + 1)):
        j = np.arange(m) + off
        ok = (j >= 0) & (j < m)
        jj = np.clip(j, 0, m - 1)
        ar = area(np.minimum(clo, clo[jj]), np.maximum(chi, chi[jj]))
        ar[~ok] = np.inf
        better = ar < best  # ties: smaller j wins because offsets ascend
        best[better] = ar[better]; nn[better] = jj[better]
    i = np.arange(m)
    mutual = nn[nn] == i
    lead = mutual & (i < nn)
    drop = mutual & (i > nn)
    tot += best[lead].sum()
    nlo = clo.copy(); nhi = chi.copy()
    nlo[lead] = np.minimum(clo[lead], clo[nn[lead]]); nhi[lead] = np.maximum(chi[lead], chi[nn[lead]])
    clo = nlo[~drop]; chi = nhi[~drop]
    if it <= 40 or len(clo) < 10: print("  iter %d: %d -> %d" % (it, m, len(clo)))
print("PLOC r=%d SAH: %.2f  (%d iterations, %.1f s)" % (R, tot / root, it, time.time() - t0))
