"""Turn one `ncu --set full` capture of render_kernel (exported as raw / sass / cuda,sass CSV pages) into the
committed summary under profiles/ and the per-workload figures bench.py reads from profiles/traffic.json.

  python tools/ncu_digest.py <prefix> <workload> <out_summary.txt> "<title line>"
  e.g. python tools/ncu_digest.py gpurun_out/r05_c4 c4 profiles/r01_v5_c4_ncu_summary.txt "kernel v5 ..."
reads <prefix>.raw.csv, <prefix>.sass.csv, <prefix>.cuda.csv
"""
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
prefix, workload, out_path, title = sys.argv[1:5]

WANT = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__warps_eligible.avg.per_cycle_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]

rows = list(csv.reader(open(prefix + ".raw.csv")))
names, units, vals = rows[0], rows[1], rows[2]
raw = {n: (units[i], vals[i]) for i, n in enumerate(names)}


def to_bytes(unit, v):
    return float(v) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


traffic = to_bytes(*raw["dram__bytes_read.sum"]) + to_bytes(*raw["dram__bytes_write.sum"])

# FP64-pipe lane instructions from the SASS page: sum over D* opcodes of thread instructions executed
srows = list(csv.reader(open(prefix + ".sass.csv")))
hi = next(i for i, r in enumerate(srows) if r and r[0] == "Address")
ix = {h: i for i, h in enumerate(srows[hi])}
fp64_lane = warp_inst = 0
for r in srows[hi + 1:]:
    if len(r) < len(srows[hi]):
        continue
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ix["Source"]].strip())
    op = m.group(2).split(".")[0] if m else "?"
    n, t = int(r[ix["Instructions Executed"]] or 0), int(r[ix["Thread Instructions Executed"]] or 0)
    warp_inst += n
    if op in ("DADD", "DMUL", "DFMA", "DSETP"):
        fp64_lane += t


def tool(name, *args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "tools", name), *args], capture_output=True, text=True).stdout


with open(out_path, "w") as f:
    f.write("# " + title + "\n")
    for n in sorted(WANT):
        if n in raw:
            f.write(f"{n} {raw[n][0]} {raw[n][1]}\n")
    f.write("\n" + tool("ncu_summary.py", prefix + ".sass.csv"))
    f.write("\n" + tool("ncu_sections.py", prefix + ".cuda.csv"))
    f.write("\n" + "\n".join(l[:170] for l in tool("ncu_lines.py", prefix + ".cuda.csv", "40").split("\n")))

tp = os.path.join(ROOT, "profiles", "traffic.json")
prof = json.load(open(tp)) if os.path.exists(tp) else {}
prof[workload] = traffic
prof[workload + "_fp64_lane_inst"] = float(fp64_lane)
prof[workload + "_warp_inst"] = float(warp_inst)
prof[workload + "_issue_active_pct"] = float(raw["smsp__issue_active.avg.pct_of_peak_sustained_active"][1])
prof[workload + "_fp64_pipe_active_pct"] = float(raw["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"][1])
prof[workload + "_avg_active_lanes"] = float(raw["smsp__thread_inst_executed_per_inst_executed.ratio"][1])
prof[workload + "_kernel_ms_under_ncu"] = float(raw["gpu__time_duration.sum"][1]) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(raw["gpu__time_duration.sum"][0], 1.0)
try:   # the capture is assumed to be of the library as built now (tools/r02_*.sh profile right after building)
    prof[workload + "_kernel_id"] = open(os.path.join(ROOT, "rust-ray-tracer_b200", "csrc", "librrt_b200.kernel_id")).read().strip()
except OSError:
    prof[workload + "_kernel_id"] = None
prof[workload + "_source"] = os.path.relpath(out_path, ROOT)
prof["source"] = ("per workload: <w> = dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of render_kernel (ncu --set full); "
                  "<w>_fp64_lane_inst = sum over DADD/DMUL/DFMA/DSETP of thread instructions executed; see <w>_source")
json.dump(prof, open(tp, "w"), indent=1)
print(f"{workload}: traffic {traffic / 1e9:.3f} GB/launch, FP64 lane inst {fp64_lane / 1e9:.1f} G, warp inst {warp_inst / 1e9:.2f} G -> {out_path}")
