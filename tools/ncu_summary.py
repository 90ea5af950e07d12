"""Summarise an `ncu --page source --csv` dump: opcode histogram (warp-level instructions executed),
stall-sample totals, SIMT efficiency.  Usage: ncu_summary.py file.src.csv"""
import csv
import re
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
ops, ops_thr, samples = Counter(), Counter(), Counter()
stalls = Counter()
tot_i = tot_t = 0
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    sass = r[ix["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", sass)
    op = m.group(2).split(".")[0] if m else "?"
    n = int(r[ix["Instructions Executed"]] or 0)
    t = int(r[ix["Thread Instructions Executed"]] or 0)
    ops[op] += n
    ops_thr[op] += t
    samples[op] += int(r[ix["# Samples"]] or 0)
    tot_i += n
    tot_t += t
    for c in stall_cols:
        stalls[c] += int(r[ix[c]] or 0)
print(f"warp instructions {tot_i:,}  thread instructions {tot_t:,}  avg active threads {tot_t / max(tot_i, 1):.2f}")
print("opcode            warp-inst      share   avg-thr  samples")
for op, n in ops.most_common(28):
    print(f"{op:14s} {n:14,} {100 * n / tot_i:8.2f}% {ops_thr[op] / max(n, 1):8.2f} {samples[op]:8d}")
ts = sum(stalls.values())
print("stall samples:", ", ".join(f"{k[6:]} {100 * v / ts:.1f}%" for k, v in stalls.most_common(10)))
