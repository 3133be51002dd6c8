"""Sums gpu__time_duration per kernel name over the second half of an ncu launch-list CSV (the first half is warm-up).
usage: python tools/launch_summary.py launches.csv"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5 and r[-1].replace(".", "").isdigit()]
half = rows[len(rows) // 2:]
acc = collections.OrderedDict()
for r in half:
    a = acc.setdefault(r[4][:90], [0, 0.0])
    a[0] += 1
    a[1] += float(r[-1]) / 1e3
for k, a in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print(f"{a[1]:9.1f} us {a[0]:4d}x {a[1] / a[0]:8.1f} us each  {k}")
print(f"total {sum(a[1] for a in acc.values()):.1f} us")
