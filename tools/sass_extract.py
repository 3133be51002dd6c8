"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / bulk-TMA / cp.async / packed-fp32 code in the
built library.  usage: python tools/sass_extract.py [lib.so] > profiles/rN_sass_extract.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "rq_vae_recommender_b200/librqengine.so"
MN = ["UTCHMMA", "LDTM", "UTCBAR", "UBLKCP", "LDGSTS", "SYNCS", "FFMA2", "FADD2", "FMNMX3", "REDG", "ATOMS", "STG.E.128",
      "LDG.E.128"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
counts, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        for k in MN:
            if re.search(r"\b" + re.escape(k), line):
                counts[cur][k] += 1
print("# cuobjdump -sass %s (release build, sm_100a): per-kernel counts of the mnemonics that prove tcgen05 / TMEM /" % lib)
print("# bulk-TMA / cp.async / packed fp32 code (tools/sass_extract.py).")
print("kernel," + ",".join(MN))
for fn, c in sorted(counts.items(), key=lambda kv: demangle(kv[0])):
    name = re.sub(r"\(.*", "", re.sub(r"\((int|bool)\)", "", demangle(fn)))
    print(name + "," + ",".join(str(c[k]) for k in MN))
