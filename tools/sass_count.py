"""Development aid (CPU only): static SASS figures of a built librrt_b200.so — instruction count and opcode histogram
per kernel, and whether two builds contain the same instruction stream for a kernel (register names and addresses
included) — to check on a box without a GPU that a refactoring left the default kernel untouched, and what a
compile-time variant does to the code size of the (instruction-cache sensitive) render kernel.
  python tools/sass_count.py [lib.so] [--kernel render_kernelILb0ELb1] [--diff other.so] [--ops]"""
import argparse
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INST = re.compile(r"^\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*")


def kernels(lib):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    res, name = collections.OrderedDict(), None
    for line in out.splitlines():
        if "Function :" in line:
            name = line.split("Function :")[1].strip()
            res[name] = []
            continue
        m = INST.match(line)
        if m and name:
            res[name].append(m.group(2).strip())
    return res


def opcode(inst):
    t = inst.split()
    if t[0].startswith("@"):
        t = t[1:]
    return t[0].split(".")[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("lib", nargs="?", default=os.path.join(ROOT, "rust-ray-tracer_b200", "csrc", "librrt_b200.so"))
    ap.add_argument("--kernel", default="render_kernel")
    ap.add_argument("--diff")
    ap.add_argument("--ops", action="store_true")
    a = ap.parse_args()
    ks = kernels(a.lib)
    other = kernels(a.diff) if a.diff else None
    for name, insts in ks.items():
        if a.kernel not in name:
            continue
        short = re.sub(r"_GLOBAL__N__[0-9a-f]+_\d+_", "", name)[:90]
        line = f"{short}: {len(insts)} instructions"
        if other is not None:
            o = next((v for k, v in other.items() if re.sub(r"_GLOBAL__N__[0-9a-f_a-z]+?cu_[0-9a-f]+", "", k) ==
                      re.sub(r"_GLOBAL__N__[0-9a-f_a-z]+?cu_[0-9a-f]+", "", name)), None)
            line += "  | other: " + ("absent" if o is None else f"{len(o)} instructions, " + ("IDENTICAL" if o == insts else "different"))
        print(line)
        if a.ops:
            h = collections.Counter(opcode(i) for i in insts)
            print("   " + "  ".join(f"{k}:{v}" for k, v in h.most_common(24)))


if __name__ == "__main__":
    main()
