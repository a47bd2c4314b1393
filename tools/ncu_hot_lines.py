"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump per CUDA source line.
usage: ncu -i rep --page source --print-source cuda,sass --csv --kernel-name regex:K | python tools/ncu_hot_lines.py [N]"""
import csv
import sys

top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
sort_col = {"samples": 1, "instr": 2, "smem": 8}[sys.argv[2]] if len(sys.argv) > 2 else 1
rows = list(csv.reader(sys.stdin))
hdr = None
cur_file = ""
lines = {}
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        cur_file = r[1].split("/")[-1]
        continue
    if len(r) > 10 and r[0] == "Line No":
        hdr = r
        ix = {n: hdr.index(n) for n in ("# Samples", "Instructions Executed", "stall_long_sb", "stall_short_sb", "stall_wait",
                                        "stall_barrier", "stall_branch_resolving", "L1 Wavefronts Shared", "L1 Wavefronts Shared Excessive")}
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0]:  # a CUDA line summary row
        try:
            lines[(cur_file, int(r[0]))] = [r[1].strip()] + [int(r[ix[k]] or 0) for k in ix]
        except ValueError:
            pass
tot_s = sum(v[1] for v in lines.values()) or 1
tot_i = sum(v[2] for v in lines.values()) or 1
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
print("smem wavefronts %d (excessive %d)" % (sum(v[8] for v in lines.values()), sum(v[9] for v in lines.values())))
print("%7s %6s %10s %6s %7s %7s %6s %6s %9s %9s  source" % ("samples", "%", "instr", "%", "long_sb", "shortsb", "wait", "bar",
                                                           "smem_wf", "smem_exc"))
for (f, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][sort_col])[:top]:
    print("%7d %5.1f%% %10d %5.1f%% %7d %7d %6d %6d %9d %9d  %s:%d: %s" % (v[1], 100.0 * v[1] / tot_s, v[2], 100.0 * v[2] / tot_i,
                                                                         v[3], v[4], v[5], v[6], v[8], v[9], f, ln, v[0][:80]))
