"""In-tree build of csrc/librrt_b200.so for sm_100a (called by __graft_entry__.build()).

nvcc cross-compiles without a GPU.  Flags that matter for parity with the reference's Rust f64
arithmetic: -fmad=false (no FMA contraction), -prec-div=true, -prec-sqrt=true, never --use_fast_math.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("RRT_OUT") or os.path.join(CSRC, "librrt_b200.so")   # RRT_OUT: variant builds (tools/)
SOURCES = ["rrt_api.cu", "rrt_render.cu", "rrt_microbench.cu", "rrt_hostframe.cu", "rrt_octree.cpp", "rrt_loader.cpp"]
HEADERS = ["rrt_device.cuh", "rrt_trace.cuh", "rrt_shade.cuh", "rrt_raygen.cuh", "rrt_render.h", "rrt_octree.h",
           os.path.join("..", "..", "include", "rrt.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-fno-fast-math,-Wall,-fopenmp", "-lgomp",
    "-Xptxas", "-v",
]


def nvcc_path():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= t for d in deps)


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    cmd = [nvcc_path(), "-shared"] + NVCC_FLAGS + os.environ.get("RRT_EXTRA_FLAGS", "").split()
    if os.path.exists("/usr/bin/g++"):
        cmd += ["-ccbin", "/usr/bin/g++"]
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", OUT]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = r.stdout + r.stderr
    with open(OUT[:-3] + ".log" if os.environ.get("RRT_OUT") else os.path.join(CSRC, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if r.returncode != 0:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed")
    # what device code this library holds: sha1 over the kernel sources and the flags they were compiled with
    # (bench.py reports it next to the kernel id of the committed ncu capture it takes its roofline counts from)
    import hashlib
    h = hashlib.sha1(" ".join(NVCC_FLAGS + os.environ.get("RRT_EXTRA_FLAGS", "").split()).encode())
    for f in ["rrt_render.cu", "rrt_device.cuh", "rrt_trace.cuh", "rrt_shade.cuh", "rrt_raygen.cuh", "rrt_octree.h"]:
        h.update(open(os.path.join(CSRC, f), "rb").read())
    with open(OUT[:-3] + ".kernel_id", "w") as f:
        f.write(h.hexdigest()[:16] + "\n")
    if verbose:
        print(log)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
