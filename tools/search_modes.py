"""Experiment: every combination of the two searches round the walk, and the adaptive switch, per workload.
   python tools/search_modes.py c2 c4 ..."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200 import abi  # noqa: E402

L = abi.lib()
for name in sys.argv[1:] or ["c4"]:
    desc, W, H = bench.WORKLOADS[name]
    rt = rrt_b200.RayTracer(bench.load_workload(name))
    sc = rrt_b200.Scene(W, H)
    kw = {}
    if name == "c5":
        o, r = rrt_b200.orbit_camera(5)
        kw = dict(origin=o, camera_rot=r)
    for label, mode, f, s in [x for x in (("walk only", 0, 0, 0), ("filter", 3, 1, 0), ("sparse", 3, 0, 1), ("both", 3, 1, 1), ("adaptive", 1, 0, 0)) if not os.environ.get("RRT_MODES") or x[0] in os.environ["RRT_MODES"].split(",")]:
        abi.check(L.rrt_scene_search_mode(rt.handle, mode, f, s))
        ms = []
        for _ in range(6):
            sc.draw_scene(rt, **kw)
            ms.append(sc.device_ms)
        st = (C.c_uint32 * 12)()
        abi.check(L.rrt_scene_search_state(rt.handle, st))
        print(f"  {name} {label:10s} {min(ms):8.3f} ms   (frames {[round(m, 2) for m in ms]})  state {list(st)}", flush=True)
    rt.close()
