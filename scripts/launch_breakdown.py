"""Per-kernel totals of one timed frame from an ncu launch list (--metrics gpu__time_duration.sum --csv).
usage: python scripts/launch_breakdown.py gpurun_out/launches.csv [frame_index]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ik = hdr.index('Kernel Name'); iv = hdr.index('Metric Value'); ig = hdr.index('Grid Size')
seq = [(r[ik].split('(')[0].replace('<unnamed>::', '').replace('void ', ''), float(r[iv]) / 1e3, r[ig]) for r in rows[1:]]
idx = [i for i, s in enumerate(seq) if s[0].startswith('update_tris')]
f = int(sys.argv[2]) if len(sys.argv) > 2 else 2
frame = seq[idx[f]:idx[f + 1]]
agg = collections.OrderedDict()
for s in frame:
    a = agg.setdefault(s[0], [0, 0.0]); a[0] += 1; a[1] += s[1]
tot = sum(a[1] for a in agg.values())
for k, a in agg.items():
    print("%-60s n=%3d %9.1f us %5.1f%%" % (k[:60], a[0], a[1], 100 * a[1] / tot))
print("frame total %.1f us (serialised, cold-cache ncu times)" % tot)
if len(sys.argv) > 3:
    for s in frame: print("%-60s %9.1f us %s" % s)
