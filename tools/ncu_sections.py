"""Group an `ncu --page source --csv --print-source cuda,sass` dump by code section of the traversal
(line ranges read from landmarks in the current rrt_trace.cuh).  Usage: ncu_sections.py file.cuda.csv"""
import collections
import csv
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(os.path.join(ROOT, "rust-ray-tracer_b200", "csrc", "rrt_trace.cuh")).read().split("\n")


def line_of(marker):
    return next(i + 1 for i, l in enumerate(src) if marker in l)


marks = [(1, "trace: min/max helpers"), (line_of("struct AxisRay"), "trace: axis setup / markstein / slab_axis"),
         (line_of("conservative culling in f32"), "trace: cull ray + cull_box (f32, all hierarchies)"),
         (line_of("slab_hit_bit(double"), "trace: child hit predicate"),
         (line_of("bool test_record"), "trace: test_record"), (line_of("void walk("), "walk: set-up"),
         (line_of("---- own triangles"), "walk: node header + list hierarchy + leaf records"),
         (line_of("---- children"), "walk: children planes + slabs"),
         (line_of("double nxy"), "walk: children tmin/tmax/key/hit + subtree boxes"),
         (line_of("as repeated selection of the smallest"), "walk: child sort + push"),
         (line_of("---- next node"), "walk: pop next node"),
         (line_of("void child_interval("), "search: way up (child_interval)"),
         (line_of("SearchResult scene_search("), "search: set-up + hierarchy steps + leaf bookkeeping"),
         (line_of("(d): up from"), "search: way up (child_interval)"),
         (line_of("void traverse("), "trace: traverse dispatch")]


def section(f, l):
    if f != "rrt_trace.cuh":
        return f
    name = marks[0][1]
    for start, n in marks:
        if l >= start:
            name = n
    return name


rows = list(csv.reader(open(sys.argv[1])))
fp = None
inst, thr, smp = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fp = r[1].split("/")[-1]
    elif r[0].isdigit() and len(r) > 8:
        try:
            n, t, s = int(r[7]), int(r[8]), int(r[6])
        except ValueError:
            continue
        k = section(fp, int(r[0]))
        inst[k] += n
        thr[k] += t
        smp[k] += s
T, S = sum(inst.values()), sum(smp.values())
print(f"total warp instructions {T:,}, stall samples {S:,}")
for k, n in inst.most_common():
    print(f"{k:44s} {100 * n / T:6.2f}% inst  {100 * smp[k] / S:6.2f}% smp  avg active threads {thr[k] / max(n, 1):5.1f}")
