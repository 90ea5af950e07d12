"""Development aid (CPU only): work counters of the product's traversal, compiled for the host
(tests/host_emulation), on camera rays of a workload — to compare culling-hierarchy builders without a GPU.
  python tools/cull_quality.py c4 [c2 ...]"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import bench  # noqa: E402
import test_host_emulation as T  # noqa: E402

T._build_emu()
L = C.CDLL(T.EMU_LIB)
vp = C.c_void_p
L.emu_scene_create.argtypes = [vp, C.c_uint64, vp, vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(vp)]
L.emu_scene_destroy.argtypes = [vp]
L.emu_scene_destroy.restype = None
L.emu_scene_set_shading.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_uint32, vp, C.c_uint32, vp]
L.emu_trace_counts.argtypes = [vp, C.c_uint64, vp, vp, C.c_int, vp]
for name in sys.argv[1:] or ["c4"]:
    desc, W, H = bench.WORKLOADS[name]
    sd = bench.load_workload(name)
    t0 = time.time()
    es = T.EmuScene(L, sd, parallel=True)
    build = time.time() - t0
    rng = np.random.default_rng(0)
    n = 400_000
    x = rng.uniform(-W / 2, W / 2, n) / W
    y = rng.uniform(-H / 2, H / 2, n) / H
    o = np.tile(np.array([0.0, 2.0, -10.0]), (n, 1))
    d = np.stack([x, y, np.ones(n)], 1)
    out = (C.c_uint64 * 4)()
    t0 = time.time()
    L.emu_trace_counts(es.h, n, o.ctypes.data, d.ctypes.data, 0, out)
    print(f"{name}: scene build {build:.2f} s; per primary ray (closest hit only): M-T {out[0] / n:.2f}  child slabs {out[1] / n:.1f}  "
          f"culling boxes {out[2] / n:.1f}  nodes {out[3] / n:.2f}   ({time.time() - t0:.2f} s)")
    es.close()
