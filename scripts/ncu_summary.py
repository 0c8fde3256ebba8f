"""Text + JSON summary of an ncu --set full report of the traversal kernels, for profiles/.
usage: python scripts/ncu_summary.py gpurun_out/prof_traverse_r01.ncu-rep profiles/ncu_traverse_r01"""
import collections
import csv
import io
import json
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]


def ncu(*args):
    return subprocess.run(["ncu", "-i", rep] + list(args), capture_output=True, text=True).stdout


rows = list(csv.reader(io.StringIO(ncu("--page", "raw", "--csv"))))
hdr, units, data = rows[0], rows[1], rows[2:]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"]
ik = hdr.index("Kernel Name")
txt = ["ncu --set full --clock-control none, report %s" % rep, ""]
js = {"report": rep, "launches": []}
for li, r in enumerate(data):
    name = r[ik].split("(")[0]
    txt.append("launch %d: %s" % (li, name))
    d = {"kernel": name}
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            txt.append("  %-86s %14s %s" % (k, r[i], units[i]))
            try:
                d[k] = float(r[i].replace(",", ""))
                d[k + ".unit"] = units[i]
            except ValueError:
                d[k] = r[i]
    js["launches"].append(d)
    txt.append("")


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


tr = [to_bytes(d["dram__bytes_read.sum"], d["dram__bytes_read.sum.unit"]) + to_bytes(d["dram__bytes_write.sum"], d["dram__bytes_write.sum.unit"])
      for d in js["launches"]]
js["dram_bytes_per_launch_mean"] = sum(tr) / len(tr)
txt.append("DRAM bytes per launch (read + write), mean over the captured launches: %.1f MB" % (js["dram_bytes_per_launch_mean"] / 1e6))
txt.append("")

for li in range(min(2, len(data))):
    src = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(li), "--launch-count", "1"))))
    h = [i for i, r in enumerate(src) if r and r[0] == "Address"][0]
    sh = src[h]; isrc = sh.index("Source"); ie = sh.index("Instructions Executed"); it = sh.index("Thread Instructions Executed")
    ops = collections.Counter(); opt = collections.Counter(); tot = 0; tt = 0
    for r in src[h + 1:]:
        try:
            e = int(r[ie]); t = int(r[it])
        except (ValueError, IndexError):
            continue
        op = r[isrc].strip().split()
        if op and op[0].startswith("@"):
            op = op[1:]
        o = op[0].split(".")[0] if op else "?"
        ops[o] += e; opt[o] += t; tot += e; tt += t
    txt.append("launch %d opcode mix (%d static SASS instructions, %.3g warp instructions, %.1f active lanes on average):" % (li, len(src) - h - 1, tot, tt / max(tot, 1)))
    for o, c in ops.most_common(16):
        txt.append("  %-10s %5.1f%% of warp instructions, %.1f lanes" % (o, 100.0 * c / tot, opt[o] / max(c, 1)))
    ctl = sum(ops[o] for o in ("BRA", "BSSY", "BSYNC", "BREAK", "ISETP", "PLOP3", "WARPSYNC", "VOTE"))
    txt.append("  control flow / predicate logic (BRA BSSY BSYNC BREAK ISETP PLOP3 VOTE): %.1f%%" % (100.0 * ctl / tot))
    txt.append("")
    srcl = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", str(li), "--launch-count", "1"))))
    cur = None; lines = []
    for r in srcl:
        if r and r[0] == "File Path":
            cur = r[1].split("/")[-1]; continue
        if len(r) > 9 and r[0].isdigit():
            try:
                lines.append((int(r[6]), int(r[7]), int(r[8]), cur, r[0], r[1].strip()[:100]))
            except ValueError:
                pass
    ts = sum(x[0] for x in lines); ti = sum(x[1] for x in lines)
    txt.append("launch %d hottest source lines (stall samples, share of warp instructions, active lanes):" % li)
    for s_, i_, t_, f_, ln, code in sorted(lines, key=lambda x: -x[0])[:14]:
        txt.append("  %5.1f%% smp %5.1f%% inst %4.1f lanes | %s:%s %s" % (100.0 * s_ / max(ts, 1), 100.0 * i_ / max(ti, 1), t_ / max(i_, 1), f_, ln, code))
    txt.append("")
open(out + ".txt", "w").write("\n".join(txt))
json.dump(js, open(out + ".json", "w"), indent=1)
print("\n".join(txt[:40]))
