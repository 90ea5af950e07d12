"""Development aid (CPU only): static SASS instructions per SOURCE LINE of one kernel of a built library
(nvdisasm --print-line-info on the embedded cubin; the library is built with -lineinfo).  For straight-line code
that runs once per octree-node visit (the child-slab block of traverse()) the static count is the dynamic count
per visit, which is what an instruction-issue-bound kernel pays — so two compile-time variants can be compared
without a GPU.
  python tools/sass_lines.py [lib.so] [--kernel render_kernelILb0ELb1] [--file rrt_trace.cuh] [--from L --to L]"""
import argparse
import collections
import os
import re
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(lib, kernel):
    """-> list of (file, line, opcode) for every instruction of the first kernel whose name contains `kernel`
    (device functions it calls are separate entries of the same section: they are included, tagged by their own lines)."""
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=td, check=True, capture_output=True)
        cubin = next(os.path.join(td, f) for f in sorted(os.listdir(td)) if f.startswith("rrt_render."))
        txt = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True, check=True).stdout
    rows, on, cur = [], False, ("?", 0)
    for ln in txt.splitlines():
        if ln.startswith("\t.section\t.text."):
            on = kernel in ln
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m:
            rows.append((cur[0], cur[1], m.group(1)))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("lib", nargs="?", default=os.path.join(ROOT, "rust-ray-tracer_b200", "csrc", "librrt_b200.so"))
    ap.add_argument("--kernel", default="render_kernelILb0ELb1")
    ap.add_argument("--file", default="rrt_trace.cuh")
    ap.add_argument("--from", dest="lo", type=int, default=0)
    ap.add_argument("--to", dest="hi", type=int, default=10 ** 9)
    ap.add_argument("--ops", action="store_true")
    a = ap.parse_args()
    rows = line_table(a.lib, a.kernel)
    per = collections.Counter()
    ops = collections.defaultdict(collections.Counter)
    for f, l, op in rows:
        per[(f, l)] += 1
        ops[(f, l)][op] += 1
    src = {}
    path = os.path.join(ROOT, "rust-ray-tracer_b200", "csrc", a.file)
    if os.path.exists(path):
        src = dict(enumerate(open(path).read().splitlines(), 1))
    tot = 0
    for (f, l), n in sorted(per.items()):
        if f != a.file or not (a.lo <= l <= a.hi):
            continue
        tot += n
        extra = ("   " + " ".join(f"{k}:{v}" for k, v in ops[(f, l)].most_common(6))) if a.ops else ""
        print(f"{n:5d}  {f}:{l:<4d} {src.get(l, '').strip()[:100]}{extra}")
    print(f"{tot:5d}  total in {a.file}:{a.lo}-{a.hi}   (kernel total {len(rows)})")


if __name__ == "__main__":
    main()
