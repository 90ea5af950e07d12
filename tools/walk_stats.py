"""Development aid (CPU only): histograms of the product's octree walk on whole emulated frames (primary, shadow and
mirror rays in their real proportions) — the traversal headers compiled for the host with -DRRT_WALK_STATS
(tests/host_emulation + tools/walk_stats_hook.cpp).  Answers "what does a visited node look like" without a GPU.
  python tools/walk_stats.py c4 [c2 ...] [--width 480]"""
import argparse
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import bench  # noqa: E402
import rrt_b200  # noqa: E402
import test_host_emulation as T  # noqa: E402

LIB = os.path.join(ROOT, "tools", "libwalk_stats.so")
KINDS = ["children kept (child_mask)", "children the exact slabs hit", "children left after their subtree boxes",
         "inner culling steps per own-list walk", "leaf (exact M-T) steps per own-list walk",
         "inner steps before the first accepted record (+1; 0 = none accepted)",
         "node kind: 1 own list | 2 children | 4 own hit"]


def build():
    flags = ["-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared", "-w", "-DRRT_WALK_STATS"] + T.EMU_FLAGS
    if " fma " in open("/proc/cpuinfo").read():
        flags.append("-mfma")
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    subprocess.run(["g++", *flags, "-I" + cuda_inc, "-o", LIB, T.EMU_SRC, os.path.join(T.CSRC, "rrt_octree.cpp"),
                    os.path.join(ROOT, "tools", "walk_stats_hook.cpp")], check=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workloads", nargs="*", default=["c4"])
    ap.add_argument("--width", type=int, default=480)
    a = ap.parse_args()
    build()
    L = C.CDLL(LIB)
    vp = C.c_void_p
    L.emu_scene_create.argtypes = [vp, C.c_uint64, vp, vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(vp)]
    L.emu_scene_destroy.argtypes = [vp]
    L.emu_scene_destroy.restype = None
    L.emu_scene_set_shading.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_uint32, vp, C.c_uint32, vp]
    L.emu_render_shard.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    L.walk_stats_get.argtypes = [vp]
    for name in a.workloads:
        desc, W, H = bench.WORKLOADS[name]
        w, h = a.width, max(2, int(round(a.width * H / W)) & ~1)
        sd = bench.load_workload(name)
        es = T.EmuScene(L, sd, parallel=True)
        p = rrt_b200.scene.make_params(w, h) if hasattr(rrt_b200.scene, "make_params") else None
        if p is None:
            raise SystemExit("make_params not found")
        L.walk_stats_reset()
        es.render_shard(p, want_hit_ids=False)
        hist = np.zeros((8, 64), dtype=np.uint64)
        L.walk_stats_get(hist.ctypes.data)
        print(f"== {name}: {desc}, emulated at {w}x{h} ({4 * w * h} primary rays), flags {' '.join(T.EMU_FLAGS) or '-'}")
        ar = np.arange(64, dtype=float)
        nray = 4.0 * w * h
        print(f"  totals per primary ray (with its shadow / mirror rays): node visits {hist[6].sum() / nray:.2f}, with children "
              f"{hist[0].sum() / nray:.2f}, subtree-box tests {(hist[1] * ar).sum() / nray:.2f}, own-list walks {hist[3].sum() / nray:.2f}, "
              f"inner culling steps {(hist[3] * ar).sum() / nray:.2f}, exact-test steps {(hist[4] * ar).sum() / nray:.2f}")
        if hist[7].sum():
            f = hist[7].astype(float) / nray
            print(f"  shadow filter per primary ray: node visits {f[0]:.2f}, list boxes {f[1]:.2f}, inner culling steps {f[5]:.2f}, "
                  f"subtree-box tests {f[6]:.2f}, exact tests {f[2]:.2f}; queries answered (clear) {f[3]:.3f}, handed to the exact walk {f[4]:.3f}")
        if hist[7][8:15].sum():
            f = hist[7].astype(float) / nray
            print(f"  sparse closest-hit per primary ray: inner steps {f[9]:.2f}, exact tests {f[8]:.2f}, ancestor boxes {f[12]:.2f}; "
                  f"answered: none {f[10]:.3f} + hit {f[14]:.3f}; handed over: second record within the margin {f[11]:.4f}, "
                  f"path box (miss / origin inside / point contact) {f[13]:.4f}")
        for k, label in enumerate(KINDS):
            row = hist[k].astype(float)
            tot = row.sum()
            if tot == 0:
                continue
            nz = np.nonzero(row)[0]
            mean = (row * np.arange(64)).sum() / tot
            cells = "  ".join(f"{b}:{100 * row[b] / tot:.1f}%" for b in nz if row[b] / tot >= 0.002)
            print(f"  [{k}] {label}: n={int(tot)} ({tot / (4 * w * h):.2f}/primary ray) mean {mean:.2f}\n        {cells}")
        es.close()


if __name__ == "__main__":
    main()
