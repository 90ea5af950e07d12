"""Experiment driver (needs a library built with -DRRT_PROBE=1): what does each traversal flavour cost as a kernel of
its own?   RRT_LIB_PATH=variants/probe_x.so [RRT_GLOBAL_CULL=1] python tools/probe.py c4 [c2 ...]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200 import abi  # noqa: E402

L = abi.lib()
L.rrt_debug_probe_scene.argtypes = [C.c_void_p, C.POINTER(abi.RenderParams), C.c_int, C.c_int, C.c_int,
                                    C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
for name in sys.argv[1:] or ["c4"]:
    desc, W, H = bench.WORKLOADS[name]
    rt = rrt_b200.RayTracer(bench.load_workload(name))
    p = rt.params(W, H)
    for kind, kname in ((0, "camera rays"), (1, "2nd ray (shadow/mirror)")):
        for mode, mname in ((0, "octree walk"), (1, "search only")):
            ms = (C.c_float * 2)()
            out = (C.c_uint64 * 3)()
            rc = L.rrt_debug_probe_scene(rt.handle, C.byref(p), kind, mode, 4, ms, out)
            print(f"  {name} {kname:24s} {mname:12s} rc {rc}  fill {ms[0]:8.3f} ms  trace {ms[1]:8.3f} ms  rays {out[0]:>10d} "
                  f"answered {out[1]:>10d} ({100.0 * out[1] / max(1, out[0]):5.1f} %)  sum {out[2]}", flush=True)
    rt.close()
