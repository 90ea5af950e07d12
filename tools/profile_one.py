"""Renders a few frames of one workload through the C ABI (what ncu wraps)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import rrt_b200  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 4
desc, W, H = bench.WORKLOADS[name]
rt = rrt_b200.RayTracer(bench.load_workload(name))
sc = rrt_b200.Scene(W, H)
for _ in range(frames):
    sc.draw_scene(rt)
print(name, W, H, "last frame device ms", sc.device_ms)
