"""`ncu -i rep --page raw --csv` -> metric,unit,value lines for the metrics the profiles/ summaries keep.
Usage: python tools/ncu_full_extract.py report.ncu-rep > profiles/xxx.csv"""
import csv
import subprocess
import sys

KEEP = ("gpu__time_duration", "dram__bytes", "dram__throughput", "sm__throughput", "sm__inst_executed_pipe_tensor",
        "sm__pipe_tensor", "smsp__issue_active", "smsp__inst_executed.sum", "sm__warps_active", "launch__",
        "l1tex__data_bank_conflicts", "lts__t_bytes", "lts__throughput", "smsp__cycles_active.avg",
        "sm__cycles_elapsed.max", "smsp__warp_issue_stalled", "l1tex__throughput", "sm__inst_executed.sum",
        "smsp__pcsamp", "launch__registers", "tcgen05", "sm__mem_tensor")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units, vals = rows[0], rows[1], rows[2]
print("metric,unit,value")
for k, u, v in zip(h, units, vals):
    if k == "Kernel Name":
        print(f"kernel,,{v}")
    if any(k.startswith(p) for p in KEEP):
        print(f"{k},{u},{v}")
