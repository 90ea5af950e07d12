"""Development aid: drain profile of render_kernel built with -DRRT_TIMELINE=1 (RRT_LIB_PATH=that library).
For the whole frame and for each shard of a G-way split: when the sample queue ran dry and when the last warp left."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200 import abi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
desc, W, H = bench.WORKLOADS[name]
rt = rrt_b200.RayTracer(bench.load_workload(name))
p = rt.params(W, H)
L = abi.lib()
tl = (C.c_ulonglong * 260)()
fb = torch.zeros((H, W), dtype=torch.int32, device="cuda:0")
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda:0")
for hint, g, r in [(h, g, r) for h in (0, 1) for (g, r) in [(1, 0)] + [(G, r) for r in range(G)]]:
    sh = abi.Shard(8, g, r)
    abi.check(L.rrt_scene_order_hint(rt.handle, hint))
    for it in range(4):
        flush.zero_()
        torch.cuda.synchronize()
        L.rrt_debug_timeline(None, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        abi.check(L.rrt_render_device_into_frame(rt.handle, C.byref(p), C.byref(sh), fb.data_ptr(), None))
        e1.record()
        torch.cuda.synchronize()
        L.rrt_debug_timeline(tl, 0)
    t0, tq, te, nw = tl[0], tl[1], tl[2], tl[3]
    hist = list(tl[4:])
    cum, marks = 0, {}
    for i, h in enumerate(hist):
        cum += h
        for f in (0.5, 0.9, 0.99, 0.999):
            if f not in marks and cum >= f * nw:
                marks[f] = (i + 1) * 5
    print(f"{name} hint {hint} shard {r}/{g}: event {e0.elapsed_time(e1):.3f} ms  kernel {1e-6 * (te - t0):.3f} ms  queue dry at {1e-6 * (tq - t0):.3f}  "
          f"tail {1e-3 * (te - tq):.0f} us  warps {nw}  exits after dry: 50% {marks.get(0.5)} us 90% {marks.get(0.9)} 99% {marks.get(0.99)} 99.9% {marks.get(0.999)}", flush=True)
