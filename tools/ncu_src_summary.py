"""Summarise `ncu --page source --csv` output: instruction and stall-sample totals per SASS opcode,
and the hottest instructions.  Usage: python tools/ncu_src_summary.py file.csv [top_n]"""
import csv
import collections
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
ix = {k: i for i, k in enumerate(H)}
body = [r for r in rows[hdr + 1:] if len(r) >= len(H) - 2]


def num(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0


def opcode(s):
    t = s.split()
    if not t:
        return "?"
    if t[0].startswith("@"):
        t = t[1:]
    return t[0].split(".")[0] if t else "?"


tot_i = sum(num(r, "Instructions Executed") for r in body)
tot_s = sum(num(r, "# Samples") for r in body)
print(f"SASS instructions: {len(body)}  executed warp-instr: {tot_i:.0f}  samples: {tot_s:.0f}")
by = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in body:
    o = opcode(r[ix["Source"]])
    by[o][0] += 1
    by[o][1] += num(r, "Instructions Executed")
    by[o][2] += num(r, "# Samples")
print(f"{'opcode':12s} {'static':>6s} {'executed':>12s} {'%exec':>6s} {'samples':>8s} {'%smp':>6s}")
for o, (n, e, s) in sorted(by.items(), key=lambda kv: -kv[1][1])[:top_n]:
    print(f"{o:12s} {n:6d} {e:12.0f} {100 * e / tot_i:6.1f} {s:8.0f} {100 * s / max(tot_s, 1):6.1f}")
stalls = [k for k in H if k.startswith("stall_") and "Not Issued" not in k]
print("\nstall totals:", ", ".join(f"{k[6:]}={sum(num(r, k) for r in body):.0f}" for k in stalls))
print("\nhottest instructions by samples:")
order = sorted(range(len(body)), key=lambda i: -num(body[i], "# Samples"))[:top_n]
for i in sorted(order):
    r = body[i]
    st = sorted(((num(r, k), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{i:5d} {num(r, '# Samples'):6.0f} {num(r, 'Instructions Executed'):9.0f}  {r[ix['Source']].strip()[:70]:70s} "
          + " ".join(f"{k}={v:.0f}" for v, k in st if v))
