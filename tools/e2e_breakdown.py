"""Development aid: where the end-to-end time of one rrt_render call goes (c4, one GPU): kernel time (CUDA events inside
the call) and wall time per call, for a page-locked canvas (direct host stores) and a pageable one (render + copy)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
import rrt_b200  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
desc, W, H = bench.WORKLOADS[name]
rt = rrt_b200.RayTracer(bench.load_workload(name))
for kind in ("pinned", "pageable", "pinned", "pageable"):
    sc = rrt_b200.Scene(W, H)
    if kind == "pinned":
        host = torch.empty((H, W), dtype=torch.int32, pin_memory=True)
        sc.canvas.buffer = host.numpy().view(np.uint32).reshape(-1)
    for _ in range(4):
        sc.draw_scene(rt)
    torch.cuda.synchronize()
    ks, t0 = [], time.perf_counter()
    for _ in range(20):
        sc.draw_scene(rt)
        ks.append(sc.device_ms)
    wall = (time.perf_counter() - t0) / 20 * 1e3
    print(f"{name} {kind:9s}: kernel {sum(ks) / len(ks):.3f} ms (min {min(ks):.3f})  wall per call {wall:.3f} ms  -> outside the kernel {wall - sum(ks) / len(ks):.3f} ms", flush=True)
rt.close()
