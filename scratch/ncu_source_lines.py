"""Aggregate an ncu `--page source --csv --print-source sass,cuda` export by CUDA source line:
share of executed warp instructions, share of stall samples, active lanes per instruction, long-scoreboard share.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > src.csv; python scratch/ncu_source_lines.py src.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
cur = None
agg = collections.defaultdict(lambda: [0, 0, 0, 0, ""])
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        ii, ti, si, li = (r.index(k) for k in ("Instructions Executed", "Thread Instructions Executed", "# Samples", "stall_long_sb"))
        continue
    try:
        ln = int(r[0])
        a = agg[(cur, ln)]
        a[0] += int(r[ii] or 0); a[1] += int(r[ti] or 0); a[2] += int(r[si] or 0); a[3] += int(r[li] or 0); a[4] = r[1]
    except (ValueError, IndexError):
        pass
tot_i = sum(a[0] for a in agg.values())
tot_s = sum(a[2] for a in agg.values())
print(f"warp instructions {tot_i}, samples {tot_s}")
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, _), a in agg.items():
    b = byfile[f]; b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
for f, b in sorted(byfile.items(), key=lambda x: -x[1][2]):
    print(f"{f:30s} inst {100 * b[0] / tot_i:5.1f}%  samples {100 * b[2] / tot_s:5.1f}%  lanes/inst {b[1] / max(b[0], 1):5.1f}")
print()
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f:22s}:{l:4d} inst {100 * a[0] / tot_i:5.2f}% smp {100 * a[2] / tot_s:5.2f}% lanes {a[1] / max(a[0], 1):5.1f} "
          f"long_sb {100 * a[3] / max(a[2], 1):3.0f}% | {a[4].strip()[:86]}")
