"""Per-source-line totals from `ncu -i rep --page source --csv --print-source cuda,sass`:
executed warp instructions and stall samples by (file, line).  Usage: ncu_line_summary.py file.csv [top_n]"""
import csv
import collections
import os
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
H = None
acc = collections.OrderedDict()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = os.path.basename(r[1])
    elif r[0] == "Line No":
        H = {k: i for i, k in enumerate(r)}
    elif H and r[0].isdigit():
        try:
            ins = float(r[H["Instructions Executed"]])
            smp = float(r[H["# Samples"]])
        except ValueError:
            continue
        key = (cur_file, int(r[0]))
        a = acc.setdefault(key, [0.0, 0.0, r[1].strip()])
        a[0] += ins
        a[1] += smp
tot_i = sum(a[0] for a in acc.values())
tot_s = sum(a[1] for a in acc.values())
print(f"total executed warp-instr {tot_i:.0f}, samples {tot_s:.0f}")
byfile = collections.defaultdict(lambda: [0.0, 0.0])
for (f, l), a in acc.items():
    byfile[f][0] += a[0]
    byfile[f][1] += a[1]
for f, (i, s) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:28s} {100 * i / tot_i:5.1f}% instr  {100 * s / max(tot_s, 1):5.1f}% samples")
print(f"\ntop {top_n} lines by executed instructions:")
for (f, l), a in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top_n]:
    print(f"{f:22s}:{l:4d} {100 * a[0] / tot_i:5.1f}% instr {100 * a[1] / max(tot_s, 1):5.1f}% smp  {a[2][:90]}")
